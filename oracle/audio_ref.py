"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's mel <-> audio path.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package; the product path
(neural_sound_generation_b200/audio.py + csrc/audio.cu) never does.

What is restated (numpy, float64 unless said otherwise), each function citing /root/reference/src/audio_tacotron.py:

* ``preemphasis`` / ``inv_preemphasis``      :22-30   scipy.signal.lfilter([1, -k], [1]) / ([1], [1, -k])
* ``melspectrogram``                         :70-78   pre-emphasis -> STFT -> |.| -> mel basis -> dB -> normalise
* ``inv_mel_spectrogram``                    :99-116  denormalise -> amplitude -> pinv(mel basis), floor 1e-10 -> ** power
                                                      -> Griffin-Lim -> inverse pre-emphasis
* ``griffin_lim``                            :142-152 60 iterations of  angles = exp(1j * angle(stft(y)));  y = istft(S * angles)
* ``stft`` / ``istft``                       :154-161 ``librosa.stft(y, n_fft, hop_length)`` / ``librosa.istft(D, hop_length)``
* ``mel_basis``                              :205-216 ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)``
* ``amp_to_db`` ``db_to_amp`` ``normalize`` ``denormalize``   :218-256 (hparams: allow_clipping_in_normalization = True,
                                                      symmetric_mels = False, max_abs_value = 1, min_level_db = -100,
                                                      ref_level_db = 20, power = 1.5; hparams_tacotron.py:95-117)

**Parity status: unpinned against librosa.**  The arithmetic of this path lives in librosa (un-vendored,
``requirements.txt`` lists it WITHOUT a version) and scipy; librosa is not installed in this image and cannot be
installed (no network), and the reference ships no audio fixtures.  The functions below restate librosa's published
algorithms as they were when the reference was written (0.6.x): ``stft`` = centred frames (``np.pad(mode='reflect')`` by
n_fft // 2), periodic Hann window (``scipy.signal.get_window('hann', n_fft, fftbins=True)``), ``1 + len(y) // hop``
frames, rfft; ``istft`` = irfft, synthesis window, overlap-add, division by the window sum-of-squares envelope where it
exceeds ``tiny(float32)``, trimming n_fft // 2 at both ends; ``filters.mel`` = Slaney mel scale (linear below 1 kHz, log
above), triangular filters, Slaney area normalisation.  Two independent anchors are kept in tests/test_oracle_audio.py:
``scipy.signal.stft`` / ``istft`` (boundary='even' is librosa's reflect padding; scaling undone) must agree with
``stft`` / ``istft`` here to 1e-12, and scipy.signal.lfilter itself is what ``preemphasis`` calls.  The one place where
the reference is random -- the initial phases ``np.random.rand(*S.shape)`` (:146) -- is an explicit argument
(``angles0``), so that the product and the oracle start from the same phases.
"""
import numpy as np
from scipy import signal

# hparams_tacotron.py:65-117 (the values main.py's audio branch runs with; main.py:176-179 passes 22050/1024/256/80)
HP = dict(sample_rate=22050, fft_size=1024, hop_size=256, num_mels=80, fmin=125, fmax=7600, min_level_db=-100.0,
          ref_level_db=20.0, max_abs_value=1.0, power=1.5, griffin_lim_iters=60, preemphasis=0.97)


def preemphasis(wav, k=HP['preemphasis']):
    return signal.lfilter([1, -k], [1], wav)                                  # :22-25


def inv_preemphasis(wav, k=HP['preemphasis']):
    return signal.lfilter([1], [1, -k], wav)                                  # :27-30


def hann_periodic(n):
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)                 # get_window('hann', n, fftbins=True)


def stft(y, n_fft=1024, hop=256):
    """librosa.stft(y, n_fft, hop_length) -> (1 + n_fft // 2, 1 + len(y) // hop) complex."""
    y = np.asarray(y, dtype=np.float64)
    yp = np.pad(y, n_fft // 2, mode='reflect')
    n_frames = 1 + (len(yp) - n_fft) // hop
    idx = np.arange(n_fft)[:, None] + hop * np.arange(n_frames)[None, :]
    frames = yp[idx] * hann_periodic(n_fft)[:, None]
    return np.fft.rfft(frames, axis=0)


def istft(D, hop=256):
    """librosa.istft(D, hop_length) -> hop * (frames - 1) samples."""
    n_fft = 2 * (D.shape[0] - 1)
    n_frames = D.shape[1]
    win = hann_periodic(n_fft)
    total = n_fft + hop * (n_frames - 1)
    y = np.zeros(total)
    wss = np.zeros(total)
    ytmp = np.fft.irfft(D, n=n_fft, axis=0) * win[:, None]
    for i in range(n_frames):
        y[i * hop:i * hop + n_fft] += ytmp[:, i]
        wss[i * hop:i * hop + n_fft] += win ** 2
    nz = wss > np.finfo(np.float32).tiny
    y[nz] /= wss[nz]
    return y[n_fft // 2:total - n_fft // 2]


def griffin_lim(S, angles0, n_iter=HP['griffin_lim_iters'], n_fft=1024, hop=256, trace=None):
    """:142-152.  S (1 + n_fft/2, T) magnitudes, angles0 = np.random.rand(*S.shape) of :146."""
    Sc = np.abs(S).astype(np.complex128)
    y = istft(Sc * np.exp(2j * np.pi * angles0), hop)
    for _ in range(n_iter):
        if trace is not None:
            trace.append(y.copy())
        ang = np.exp(1j * np.angle(stft(y, n_fft, hop)))
        y = istft(Sc * ang, hop)
    return y


def _hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    mels = f / (200.0 / 3)
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / (200.0 / 3), np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f = m * (200.0 / 3)
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / (200.0 / 3), np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f)


def mel_basis(sr=22050, n_fft=1024, n_mels=80, fmin=HP['fmin'], fmax=HP['fmax']):
    """librosa.filters.mel (Slaney scale and normalisation) -> (n_mels, 1 + n_fft // 2) float32, as librosa returns it."""
    fftfreqs = np.linspace(0, sr / 2.0, 1 + n_fft // 2)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, 1 + n_fft // 2))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0, np.minimum(lower, upper))
    w *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]
    return w.astype(np.float32)


def amp_to_db(x):
    min_level = np.exp(HP['min_level_db'] / 20 * np.log(10))                  # :218-220
    return 20 * np.log10(np.maximum(min_level, x))


def db_to_amp(x):
    return np.power(10.0, x * 0.05)                                           # :222-223


def normalize(S):
    return np.clip(HP['max_abs_value'] * ((S - HP['min_level_db']) / (-HP['min_level_db'])), 0, HP['max_abs_value'])   # :231


def denormalize(D):
    return (np.clip(D, 0, HP['max_abs_value']) * -HP['min_level_db'] / HP['max_abs_value']) + HP['min_level_db']       # :249


def mel_to_linear(mel, sr=22050, n_fft=1024, n_mels=80):
    inv = np.linalg.pinv(mel_basis(sr, n_fft, n_mels))                        # :199-203
    return np.maximum(1e-10, np.dot(inv, mel))


def melspectrogram(wav, sr=22050, n_fft=1024, hop=256, n_mels=80):
    D = stft(preemphasis(wav), n_fft, hop)                                    # :70-78
    S = amp_to_db(np.dot(mel_basis(sr, n_fft, n_mels), np.abs(D))) - HP['ref_level_db']
    return normalize(S)


def inv_mel_magnitudes(mel, sr=22050, n_fft=1024, n_mels=80):
    """:101-107 and the ``** hparams.power`` of :115: what Griffin-Lim is handed."""
    return mel_to_linear(db_to_amp(denormalize(mel) + HP['ref_level_db']), sr, n_fft, n_mels) ** HP['power']


def inv_mel_spectrogram(mel, angles0, sr=22050, n_fft=1024, hop=256, n_mels=80, n_iter=HP['griffin_lim_iters']):
    return inv_preemphasis(griffin_lim(inv_mel_magnitudes(mel, sr, n_fft, n_mels), angles0, n_iter, n_fft, hop))   # :115-116


def save_wav_scale(wav):
    """:16-19: int16 samples save_wav writes."""
    w = wav * (32767 / max(0.01, np.max(np.abs(wav))))
    return w.astype(np.int16)
