"""CPU tests of the audio oracle (oracle/audio_ref.py).  librosa -- where the reference's arithmetic
for this path lives -- is not installed here, so the restatement is anchored on independent implementations instead:
scipy.signal.stft / istft (boundary='even' = librosa's reflect padding), numpy's FFT, and algebraic properties."""
import numpy as np
import scipy.signal as ss

from oracle import audio_ref as A


def test_stft_istft_against_scipy():
    rng = np.random.RandomState(0)
    for n_fft, hop, n in ((1024, 256, 256 * 40), (512, 128, 128 * 37 + 5), (256, 64, 3000)):
        y = rng.randn(n)
        D = A.stft(y, n_fft, hop)
        assert D.shape == (n_fft // 2 + 1, 1 + n // hop)
        _, _, Z = ss.stft(y, window='hann', nperseg=n_fft, noverlap=n_fft - hop, nfft=n_fft, boundary='even', padded=False)
        Z = Z * A.hann_periodic(n_fft).sum()                       # scipy scales by 1 / sum(window)
        m = min(D.shape[1], Z.shape[1])
        assert np.abs(D[:, :m] - Z[:, :m]).max() < 1e-10
        yi = A.istft(D, hop)
        assert len(yi) == hop * (D.shape[1] - 1)
        assert np.abs(yi - y[:len(yi)]).max() < 1e-10             # perfect reconstruction (Hann, 75 % overlap)


def test_mel_basis_properties():
    mb = A.mel_basis()
    assert mb.shape == (80, 513) and mb.dtype == np.float32
    assert (mb >= 0).all() and (mb.sum(1) > 0).all()
    freqs = np.linspace(0, 11025, 513)
    peak = freqs[mb.argmax(1)]
    assert (np.diff(peak) > 0).all() and peak[0] >= 125 and peak[-1] <= 7600
    # Slaney area normalisation: every filter integrates to ~1 over frequency (in units of the bin spacing)
    area = mb.sum(1) * (freqs[1] - freqs[0])
    assert np.abs(area - 1).max() < 0.2
    # the scale itself: linear below 1 kHz, 1 kHz <-> mel 15, 6.4 kHz <-> mel 42
    assert abs(float(A._hz_to_mel(1000.0)) - 15.0) < 1e-12 and abs(float(A._hz_to_mel(6400.0)) - 42.0) < 1e-9
    assert np.allclose(A._mel_to_hz(A._hz_to_mel(np.array([60.0, 999.0, 1000.0, 4000.0]))), [60, 999, 1000, 4000])


def test_level_conversions_and_filters():
    x = np.array([1e-9, 1e-5, 1e-3, 0.5, 1.0, 3.0])
    db = A.amp_to_db(x)
    assert np.allclose(db[0], -100.0) and np.allclose(A.db_to_amp(db[2:]), x[2:])
    n = A.normalize(db - 20.0)
    assert n.min() >= 0 and n.max() <= 1
    assert np.allclose(A.denormalize(n[2:5]), (db - 20.0)[2:5])
    w = np.random.RandomState(1).randn(5000)
    assert np.allclose(A.inv_preemphasis(A.preemphasis(w)), w, atol=1e-9)
    assert np.array_equal(A.save_wav_scale(np.array([0.5, -1.0, 0.25])), np.array([16383, -32767, 8191], dtype=np.int16))


def test_griffin_lim_reduces_spectral_error_and_round_trip():
    rng = np.random.RandomState(3)
    t = np.arange(256 * 30) / 22050.0
    wav = 0.3 * np.sin(2 * np.pi * 440 * t) + 0.2 * np.sin(2 * np.pi * 1330 * t + 1.0) + 0.01 * rng.randn(len(t))
    S = np.abs(A.stft(wav))
    ang = rng.rand(*S.shape)
    trace = []
    y = A.griffin_lim(S, ang, n_iter=20, trace=trace)
    assert y.shape == (256 * 30,)
    err = [np.linalg.norm(np.abs(A.stft(v)) - S) / np.linalg.norm(S) for v in (trace[0], trace[5], y)]
    assert err[2] < err[1] < err[0] and err[2] < 0.35
    mel = A.melspectrogram(wav)
    assert mel.shape == (80, 31) and 0 <= mel.min() and mel.max() <= 1
    out = A.inv_mel_spectrogram(mel, ang, n_iter=5)
    assert out.shape == (256 * 30,) and np.isfinite(out).all()
