"""GPU test of ``nsg_istft_static`` -- the allocation-free variant of ``nsg_istft`` whose frame scratch is a file-scope
device array inside the library (include/nsg_b200.h; no counterpart in the reference).
Called through the C-ABI and compared with the existing entry point on two calls in a row; the bar is the
audio tests' 2e-5 (tests/test_gpu_audio.py), the two are expected to agree bit for bit (same kernels, another scratch
address).  Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from neural_sound_generation_b200 import _cabi as A

TOL = 2e-5


def _rel(a, ref):
    a, ref = a.double(), ref.double()
    return float((a - ref).norm() / ref.norm().clamp_min(1e-30))


def _spectrum(frames, n_fft, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(frames, n_fft // 2 + 1, 2, generator=g).to('cuda').contiguous()


def _istft_workspace(spec, frames, n_fft, hop):
    lib = A.load()
    ws = torch.empty(lib.nsg_istft_workspace_bytes(frames, n_fft), dtype=torch.uint8, device='cuda')
    y = torch.empty(hop * (frames - 1), dtype=torch.float32, device='cuda')
    A.check(lib.nsg_istft(A.ptr(spec), frames, n_fft, hop, A.ptr(ws), A.ptr(y), A.stream_ptr()), 'nsg_istft')
    return y


def _istft_static(spec, frames, n_fft, hop):
    lib = A.load()
    y = torch.full((hop * (frames - 1),), float('nan'), dtype=torch.float32, device='cuda')
    A.check(lib.nsg_istft_static(A.ptr(spec), frames, n_fft, hop, A.ptr(y), A.stream_ptr()), 'nsg_istft_static')
    return y


def test_istft_static_matches_istft_on_two_calls_in_a_row():
    A.require_device()
    lib = A.load()
    assert lib.nsg_istft_static_max_floats() == 1 << 20
    # second call smaller than the first, with other parameters: it runs on the first call's left-over scratch
    cases = [(41, 1024, 256, 11), (23, 512, 128, 12), (1024, 1024, 256, 13)]        # the last one fills the scratch exactly
    for frames, n_fft, hop, seed in cases:
        spec = _spectrum(frames, n_fft, seed)
        want = _istft_workspace(spec, frames, n_fft, hop)
        got = _istft_static(spec, frames, n_fft, hop)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        e = _rel(got, want)
        assert e < TOL, (frames, n_fft, hop, e)
        assert torch.equal(got, want), 'same kernels on another scratch address: expected bit-identical output'
    # the same input twice in a row: identical results
    spec = _spectrum(41, 1024, 11)
    a = _istft_static(spec, 41, 1024, 256)
    b = _istft_static(spec, 41, 1024, 256)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


def test_istft_static_rejects_sizes_beyond_its_scratch():
    A.require_device()
    lib = A.load()
    frames, n_fft, hop = 1025, 1024, 256              # one frame more than 2^20 floats
    spec = _spectrum(frames, n_fft, 3)
    y = torch.zeros(hop * (frames - 1), dtype=torch.float32, device='cuda')
    rc = lib.nsg_istft_static(A.ptr(spec), frames, n_fft, hop, A.ptr(y), A.stream_ptr())
    assert rc == -1                                    # NSG_E_INVALID
    msg = lib.nsg_last_error().decode()
    assert 'exceeds' in msg and 'nsg_istft' in msg
    torch.cuda.synchronize()
    assert float(y.abs().max()) == 0.0                 # nothing was launched
    with pytest.raises(A.NsgError):
        A.check(lib.nsg_istft_static(A.ptr(spec), 1, n_fft, hop, A.ptr(y), A.stream_ptr()), 'nsg_istft_static')
