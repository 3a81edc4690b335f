"""mel <-> audio on the GPU: drop-in for the functions of the reference's ``src/audio_tacotron.py`` that
``main.py:164-197`` calls after decoding.

Same names, argument order and layouts as the reference: ``melspectrogram(wav, sample_rate, fft_size, hop_size, n_mels)``
-> ``(n_mels, frames)``; ``inv_mel_spectrogram(mel, sample_rate, fft_size, hop_size, n_mel)`` -> waveform;
``_stft`` / ``_istft`` / ``_griffin_lim`` / ``_linear_to_mel`` / ``_mel_to_linear`` / ``_build_mel_basis`` /
``preemphasis`` / ``inv_preemphasis`` / ``_amp_to_db`` / ``_db_to_amp`` / ``_normalize`` / ``_denormalize`` /
``save_wav``; ``hparams`` carries the values of ``hparams_tacotron.py:65-117`` that this path reads.  Inputs may be numpy
arrays (as the reference passes them) or CUDA tensors; numpy in -> numpy out, tensor in -> tensor out.

Everything per-sample / per-frame runs in ``csrc/audio.cu`` (``nsg_stft``, ``nsg_istft``, ``nsg_griffin_lim`` -- all 60
iterations in ONE persistent launch --, ``nsg_mel_to_linear``, ``nsg_melspectrogram``, ``nsg_inv_preemphasis``).  Host
side, once per (sample_rate, fft_size, n_mels): the mel filter bank (``librosa.filters.mel``'s published construction:
Slaney mel scale + area normalisation) and its pseudo-inverse (``np.linalg.pinv``, as ``audio_tacotron.py:199-203``) --
layer constants, cached and uploaded once.  There is no CPU path: without the CUDA library every call raises.
"""
import os
import types

import numpy as np
import torch

from . import _cabi as A
from . import ops

# hparams_tacotron.py:65-117 -- the values the mel <-> audio path reads
hparams = types.SimpleNamespace(
    sample_rate=22050, fft_size=1024, hop_size=256, num_mels=80, fmin=125, fmax=7600, min_level_db=-100.0,
    ref_level_db=20.0, max_abs_value=1.0, power=1.5, griffin_lim_iters=60, preemphasis=0.97, preemphasize=True,
    signal_normalization=True, allow_clipping_in_normalization=True, symmetric_mels=False, use_lws=False)

_basis_cache = {}


def _hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, 15.0 + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, f * 0.015)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    logstep = np.log(6.4) / 27.0
    return np.where(m >= 15.0, 1000.0 * np.exp(logstep * (m - 15.0)), m / 0.015)


def _build_mel_basis(sample_rate, fft_size, n_mels):
    """audio_tacotron.py:205-216 = librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax): Slaney scale (linear below 1 kHz,
    logarithmic above), triangular filters, each scaled by 2 / (its band width in Hz).  (n_mels, 1 + fft_size // 2) fp32."""
    assert hparams.fmax <= sample_rate // 2
    freqs = np.linspace(0.0, sample_rate / 2.0, 1 + fft_size // 2)
    edges = _mel_to_hz(np.linspace(_hz_to_mel(hparams.fmin), _hz_to_mel(hparams.fmax), n_mels + 2))
    width = np.diff(edges)
    dist = edges[:, None] - freqs[None, :]
    rising, falling = -dist[:-2] / width[:-1, None], dist[2:] / width[1:, None]
    tri = np.maximum(0.0, np.minimum(rising, falling))
    return (tri * (2.0 / (edges[2:] - edges[:-2]))[:, None]).astype(np.float32)


def _bases(sample_rate, fft_size, n_mels, device):
    """(basis_t [bins][n_mels], inv_basis_t [n_mels][bins]) on ``device``, built once."""
    key = (sample_rate, fft_size, n_mels, hparams.fmin, hparams.fmax, str(device))
    if key not in _basis_cache:
        basis = _build_mel_basis(sample_rate, fft_size, n_mels)
        inv = np.linalg.pinv(basis)                                            # (bins, n_mels), audio_tacotron.py:201
        _basis_cache[key] = (torch.from_numpy(np.ascontiguousarray(basis.T)).to(device),
                             torch.from_numpy(np.ascontiguousarray(inv.T.astype(np.float32))).to(device))
    return _basis_cache[key]


def _dev(x, dtype=torch.float32):
    """numpy / tensor -> contiguous CUDA tensor; second value: the caller passed numpy."""
    A.require_device()
    if isinstance(x, torch.Tensor):
        if not x.is_cuda:
            raise A.NsgError('nsg_b200.audio needs CUDA tensors or numpy arrays (got a %s tensor); there is no CPU path'
                             % x.device)
        return x.to(dtype).contiguous(), False
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x), dtype=np.float32 if dtype == torch.float32 else None)
                            ).to('cuda', dtype), True


def _out(t, as_numpy):
    return t.cpu().numpy() if as_numpy else t


def _launch(rc, what):
    A.check(rc, what)
    ops.launch_count += 1


# ---------------------------------------------------------------------------------------------- filters
def preemphasis(wav, k, preemphasize=True):
    """audio_tacotron.py:22-25: lfilter([1, -k], [1], wav) -- an FIR tap; the STFT kernels fold it into their frame
    load, this stand-alone form exists for callers of the reference's function."""
    if not preemphasize:
        return wav
    x, as_np = _dev(wav)
    y = x.clone()
    y[1:] -= k * x[:-1]
    return _out(y, as_np)


def inv_preemphasis(wav, k, inv_preemphasize=True):
    """audio_tacotron.py:27-30: lfilter([1], [1, -k], wav)."""
    if not inv_preemphasize:
        return wav
    x, as_np = _dev(wav)
    y = torch.empty_like(x)
    _launch(A.load().nsg_inv_preemphasis(A.ptr(x), x.numel(), float(k), A.ptr(y), A.stream_ptr()), 'nsg_inv_preemphasis')
    return _out(y, as_np)


# ---------------------------------------------------------------------------------------------- transforms
def _stft(y, fft_size, hop_size, _preemph=0.0):
    """audio_tacotron.py:154-158 (use_lws False): librosa.stft -> (1 + fft_size // 2, frames) complex64."""
    x, as_np = _dev(y)
    frames = 1 + x.numel() // hop_size
    spec = torch.empty(frames, fft_size // 2 + 1, 2, dtype=torch.float32, device=x.device)
    _launch(A.load().nsg_stft(A.ptr(x), x.numel(), fft_size, hop_size, float(_preemph), A.ptr(spec), A.stream_ptr()),
            'nsg_stft')
    D = torch.view_as_complex(spec).t()
    return _out(D, as_np)


def _istft(y, hop_size):
    """audio_tacotron.py:160-161: librosa.istft of a (1 + fft_size // 2, frames) spectrum."""
    as_np = not isinstance(y, torch.Tensor)
    D = torch.as_tensor(np.asarray(y)) if as_np else y
    A.require_device()
    D = D.to('cuda', torch.complex64).t().contiguous()               # frame-major
    frames, fft_size = D.shape[0], 2 * (D.shape[1] - 1)
    spec = torch.view_as_real(D).contiguous()
    ws = torch.empty(A.load().nsg_istft_workspace_bytes(frames, fft_size), dtype=torch.uint8, device=spec.device)
    out = torch.empty(hop_size * (frames - 1), dtype=torch.float32, device=spec.device)
    _launch(A.load().nsg_istft(A.ptr(spec), frames, fft_size, hop_size, A.ptr(ws), A.ptr(out), A.stream_ptr()), 'nsg_istft')
    ops.launch_count += 1
    return _out(out, as_np)


def _griffin_lim(S, fft_size, hop_size, angles=None, n_iter=None, _frame_major=False):
    """audio_tacotron.py:142-152.  S: (1 + fft_size // 2, frames) magnitudes.  ``angles`` (same shape, fractions of a
    turn) stands for the reference's ``np.random.rand(*S.shape)`` (:146); None draws them with numpy's global stream
    exactly as the reference does."""
    as_np = not isinstance(S, torch.Tensor)
    if _frame_major:
        mag = S
    else:
        mag, _ = _dev(S)
        mag = mag.abs().t().contiguous()
    frames, bins = mag.shape
    if angles is None:
        angles = np.random.rand(bins, frames)
    ang, _ = _dev(angles)
    ang = ang.t().contiguous()
    iters = hparams.griffin_lim_iters if n_iter is None else n_iter
    ws = torch.empty(A.load().nsg_griffin_lim_workspace_bytes(frames, fft_size), dtype=torch.uint8, device=mag.device)
    out = torch.empty(hop_size * (frames - 1), dtype=torch.float32, device=mag.device)
    _launch(A.load().nsg_griffin_lim(A.ptr(mag), A.ptr(ang), frames, fft_size, hop_size, iters, A.ptr(ws), A.ptr(out),
                                     A.stream_ptr()), 'nsg_griffin_lim')
    return _out(out, as_np)


# ---------------------------------------------------------------------------------------------- conversions
def _linear_to_mel(spectogram, sample_rate, fft_size, n_mels, _mel_basis=None):
    """audio_tacotron.py:193-197.  (A small dense product of a cached constant: stays a torch matmul; the fused path
    is ``melspectrogram``.)"""
    x, as_np = _dev(spectogram)
    basis = _bases(sample_rate, fft_size, n_mels, x.device)[0].t() if _mel_basis is None else _dev(_mel_basis)[0]
    return _out(basis @ x, as_np)


def _mel_to_linear(mel_spectrogram, sample_rate, fft_size, n_mels, _inv_mel_basis=None):
    """audio_tacotron.py:199-203: max(1e-10, pinv(basis) @ mel)."""
    x, as_np = _dev(mel_spectrogram)
    inv = _bases(sample_rate, fft_size, n_mels, x.device)[1].t() if _inv_mel_basis is None else _dev(_inv_mel_basis)[0]
    return _out(torch.clamp_min(inv @ x, 1e-10), as_np)


def _amp_to_db(x):
    min_level = np.exp(hparams.min_level_db / 20 * np.log(10))
    if isinstance(x, torch.Tensor):
        return 20 * torch.log10(torch.clamp_min(x, min_level))
    return 20 * np.log10(np.maximum(min_level, x))


def _db_to_amp(x):
    return torch.pow(10.0, x * 0.05) if isinstance(x, torch.Tensor) else np.power(10.0, x * 0.05)


def _normalize(S):
    clip = torch.clamp if isinstance(S, torch.Tensor) else np.clip
    return clip(hparams.max_abs_value * ((S - hparams.min_level_db) / (-hparams.min_level_db)), 0, hparams.max_abs_value)


def _denormalize(D):
    clip = torch.clamp if isinstance(D, torch.Tensor) else np.clip
    return (clip(D, 0, hparams.max_abs_value) * -hparams.min_level_db / hparams.max_abs_value) + hparams.min_level_db


# ---------------------------------------------------------------------------------------------- the two entry points
def melspectrogram(wav, sample_rate, fft_size, hop_size, n_mels):
    """audio_tacotron.py:70-78 in one launch -> (n_mels, 1 + len(wav) // hop_size), values in [0, max_abs_value]."""
    x, as_np = _dev(wav)
    basis_t = _bases(sample_rate, fft_size, n_mels, x.device)[0]
    frames = 1 + x.numel() // hop_size
    mel = torch.empty(frames, n_mels, dtype=torch.float32, device=x.device)
    k = hparams.preemphasis if hparams.preemphasize else 0.0
    _launch(A.load().nsg_melspectrogram(A.ptr(x), x.numel(), fft_size, hop_size, float(k), A.ptr(basis_t), n_mels,
                                        float(hparams.min_level_db), float(hparams.ref_level_db),
                                        float(hparams.max_abs_value), A.ptr(mel), A.stream_ptr()), 'nsg_melspectrogram')
    return _out(mel.t(), as_np)


def inv_mel_spectrogram(mel_spectrogram, sample_rate, fft_size, hop_size, n_mel, angles=None, n_iter=None):
    """audio_tacotron.py:99-116 (use_lws False): three launches -- mel -> linear magnitudes ** power, Griffin-Lim (all
    iterations), inverse pre-emphasis.  ``mel_spectrogram``: (n_mel, frames) as main.py:181 passes it."""
    mel, as_np = _dev(mel_spectrogram)
    if mel.dim() != 2 or mel.shape[0] != n_mel:
        raise A.NsgError('inv_mel_spectrogram expects a (%d, frames) mel spectrogram, got %s' % (n_mel, tuple(mel.shape)))
    frames, bins = mel.shape[1], fft_size // 2 + 1
    inv_t = _bases(sample_rate, fft_size, n_mel, mel.device)[1]
    mag = torch.empty(frames, bins, dtype=torch.float32, device=mel.device)
    _launch(A.load().nsg_mel_to_linear(A.ptr(mel), n_mel, frames, A.ptr(inv_t), bins, float(hparams.min_level_db),
                                       float(hparams.ref_level_db), float(hparams.max_abs_value), float(hparams.power),
                                       A.ptr(mag), A.stream_ptr()), 'nsg_mel_to_linear')
    y = _griffin_lim(mag, fft_size, hop_size, angles=angles, n_iter=n_iter, _frame_major=True)
    y = inv_preemphasis(y, hparams.preemphasis, hparams.preemphasize)
    return _out(y, as_np)


def save_wav(wav, path):
    """audio_tacotron.py:16-19: peak-normalise to int16 and write a RIFF/WAVE file (scipy.io.wavfile.write)."""
    from scipy.io import wavfile
    w = wav.detach().float().cpu().numpy() if isinstance(wav, torch.Tensor) else np.asarray(wav)
    w = w * (32767 / max(0.01, float(np.max(np.abs(w)))))
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    wavfile.write(path, hparams.sample_rate, w.astype(np.int16))
