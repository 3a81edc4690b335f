"""GPU tests of the mel <-> audio kernels (csrc/audio.cu through neural_sound_generation_b200/audio.py)
against the oracle restatement of the reference's librosa-based path (oracle/audio_ref.py; float64).

Bars (fp32 kernels against a float64 oracle): transforms / filters 2e-5 relative in the 2-norm; ONE Griffin-Lim
iteration from the same phases 1e-4; the full 60-iteration run is an iterated, phase-renormalising map (errors are
re-scaled by 1 / |X| at weak bins every iteration), so it is held to the oracle's own spectral convergence instead of
sample-exactness, with the sample error reported."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from neural_sound_generation_b200 import audio as G
    from oracle import audio_ref as R

TOL = 2e-5


def rel(a, ref):
    a, ref = np.asarray(a, dtype=np.complex128), np.asarray(ref, dtype=np.complex128)
    return float(np.linalg.norm(a - ref) / max(np.linalg.norm(ref), 1e-30))


def _speechlike(n, seed=0):
    rng = np.random.RandomState(seed)
    t = np.arange(n) / 22050.0
    env = 0.5 + 0.5 * np.sin(2 * np.pi * 3.0 * t)
    return (env * (0.3 * np.sin(2 * np.pi * 220 * t) + 0.2 * np.sin(2 * np.pi * 1330 * t + 1.0)
                   + 0.1 * np.sin(2 * np.pi * 3100 * t)) + 0.02 * rng.randn(n)).astype(np.float32)


@pytest.mark.parametrize('n_fft,hop,n', [(1024, 256, 256 * 41), (1024, 256, 256 * 40 + 77), (512, 128, 5000),
                                         (2048, 512, 512 * 9), (256, 64, 777), (64, 16, 200), (1024, 300, 7000)])
def test_stft_istft_vs_oracle(n_fft, hop, n):
    y = np.random.RandomState(n).randn(n).astype(np.float32)
    D = G._stft(y, n_fft, hop)
    want = R.stft(y, n_fft, hop)
    assert D.shape == want.shape and D.dtype == np.complex64
    assert rel(D, want) < TOL
    if n_fft % hop == 0:                                 # (librosa's istft is only an inverse for COLA hops)
        yi = G._istft(want.astype(np.complex64), hop)
        wi = R.istft(want, hop)
        assert yi.shape == wi.shape and rel(yi, wi) < TOL
        assert rel(yi, y[:len(yi)]) < 5e-5
    else:
        yi = G._istft(want.astype(np.complex64), hop)
        assert rel(yi, R.istft(want, hop)) < TOL
    # CUDA tensors in -> CUDA tensors out
    Dt = G._stft(torch.tensor(y, device='cuda'), n_fft, hop)
    assert Dt.is_cuda and rel(Dt.cpu().numpy(), want) < TOL


def test_stft_with_preemphasis_and_odd_frame_count():
    y = _speechlike(256 * 20)                            # 21 frames: the last CTA works on a lone frame
    D = G._stft(y, 1024, 256, _preemph=0.97)
    assert rel(D, R.stft(R.preemphasis(y.astype(np.float64)), 1024, 256)) < TOL
    assert rel(G.preemphasis(y, 0.97), R.preemphasis(y.astype(np.float64))) < 1e-6


@pytest.mark.parametrize('n,k', [(1, 0.97), (5, 0.97), (4096, 0.97), (4097, 0.97), (256 * 860, 0.97), (30000, 0.999),
                                 (10000, -0.5)])
def test_inv_preemphasis_vs_lfilter(n, k):
    x = np.random.RandomState(n).randn(n).astype(np.float32)
    y = G.inv_preemphasis(x, k)
    want = R.inv_preemphasis(x.astype(np.float64), k)
    assert y.shape == want.shape and rel(y, want) < 2e-5
    assert rel(G.inv_preemphasis(G.preemphasis(x, k), k), x) < 2e-5


def test_mel_basis_matches_oracle_and_mel_to_linear():
    assert np.array_equal(G._build_mel_basis(22050, 1024, 80), R.mel_basis(22050, 1024, 80))
    mel = np.random.RandomState(5).rand(80, 37).astype(np.float32) * 1.2 - 0.1       # also outside [0, 1]: clipped
    want = R.inv_mel_magnitudes(mel.astype(np.float64))
    import neural_sound_generation_b200._cabi as A
    from neural_sound_generation_b200 import ops
    m = torch.tensor(mel, device='cuda')
    inv_t = G._bases(22050, 1024, 80, m.device)[1]
    mag = torch.empty(37, 513, device='cuda')
    A.check(A.load().nsg_mel_to_linear(A.ptr(m), 80, 37, A.ptr(inv_t), 513, -100.0, 20.0, 1.0, 1.5, A.ptr(mag),
                                       A.stream_ptr()), 'nsg_mel_to_linear')
    got = mag.t().cpu().numpy()
    # the pseudo-inverse has negative lobes: bins where the product cancels to ~0 are floored at 1e-10 ** 1.5 on both
    # sides; compare where the oracle's value is well above the floor, and the floor set itself
    big = want > 1e-6 * want.max()
    assert np.abs(got[big] / want[big] - 1).max() < 2e-3
    assert rel(got, want) < 1e-4
    assert ops.launch_count > 0


def test_melspectrogram_vs_oracle():
    for n in (256 * 50, 256 * 33 + 100):
        wav = _speechlike(n, seed=n)
        mel = G.melspectrogram(wav, 22050, 1024, 256, 80)
        want = R.melspectrogram(wav.astype(np.float64))
        assert mel.shape == want.shape == (80, 1 + n // 256)
        assert mel.min() >= 0 and mel.max() <= 1
        assert np.abs(mel - want).max() < 2e-4           # values in [0, 1] (1 = 100 dB): 0.02 dB


def test_griffin_lim_one_iteration_and_step0_vs_oracle():
    wav = _speechlike(256 * 24, seed=1)
    S = np.abs(R.stft(wav.astype(np.float64))) ** 1.2
    ang = np.random.RandomState(9).rand(*S.shape)
    y0 = G._griffin_lim(S.astype(np.float32), 1024, 256, angles=ang, n_iter=0)
    w0 = R.griffin_lim(S, ang, n_iter=0)
    assert y0.shape == w0.shape == (256 * 24,) and rel(y0, w0) < TOL
    y1 = G._griffin_lim(S.astype(np.float32), 1024, 256, angles=ang, n_iter=1)
    assert rel(y1, R.griffin_lim(S, ang, n_iter=1)) < 1e-4
    y3 = G._griffin_lim(S.astype(np.float32), 1024, 256, angles=ang, n_iter=3)
    assert rel(y3, R.griffin_lim(S, ang, n_iter=3)) < 1e-3
    # deterministic: the same launch twice gives the same bits
    assert np.array_equal(y3, G._griffin_lim(S.astype(np.float32), 1024, 256, angles=ang, n_iter=3))


def test_griffin_lim_full_run_converges_like_the_oracle():
    n_frames = 120
    wav = _speechlike(256 * (n_frames - 1), seed=2)
    S = np.abs(R.stft(wav.astype(np.float64)))
    ang = np.random.RandomState(4).rand(*S.shape)
    want = R.griffin_lim(S, ang, n_iter=60)
    got = G._griffin_lim(S.astype(np.float32), 1024, 256, angles=ang, n_iter=60)

    def spectral_error(y):
        return np.linalg.norm(np.abs(R.stft(np.asarray(y, dtype=np.float64))) - S) / np.linalg.norm(S)
    e_got, e_want = spectral_error(got), spectral_error(want)
    print('griffin-lim 60 iterations: spectral error ours %.4f oracle %.4f, sample error %.3e' % (e_got, e_want, rel(got, want)))
    assert np.isfinite(got).all()
    assert e_got <= 1.05 * e_want + 1e-3
    assert rel(got, want) < 0.05


def test_inv_mel_spectrogram_end_to_end():
    wav = _speechlike(256 * 63, seed=3)
    mel = R.melspectrogram(wav.astype(np.float64)).astype(np.float32)          # (80, 64), what the decoder would emit
    ang = np.random.RandomState(6).rand(513, 64)
    got = G.inv_mel_spectrogram(mel, 22050, 1024, 256, 80, angles=ang, n_iter=2)
    want = R.inv_mel_spectrogram(mel.astype(np.float64), ang, n_iter=2)
    assert got.shape == want.shape == (256 * 63,)
    assert rel(got, want) < 2e-3
    full = G.inv_mel_spectrogram(torch.tensor(mel, device='cuda'), 22050, 1024, 256, 80)    # 60 iterations, own phases
    assert full.is_cuda and full.shape == (256 * 63,) and torch.isfinite(full).all()
    with pytest.raises(Exception):
        G.inv_mel_spectrogram(mel[:40], 22050, 1024, 256, 80)
    with pytest.raises(Exception):
        G._stft(torch.zeros(4000), 1024, 256)            # CPU tensor: no CPU path


def test_save_wav(tmp_path):
    from scipy.io import wavfile
    w = _speechlike(5000)
    p = str(tmp_path / 'a' / 'x.wav')
    G.save_wav(torch.tensor(w, device='cuda'), p)
    sr, data = wavfile.read(p)
    assert sr == 22050 and data.dtype == np.int16 and np.array_equal(data, R.save_wav_scale(w))
