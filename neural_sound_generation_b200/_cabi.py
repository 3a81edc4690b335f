"""ctypes binding of libnsg_b200.so (C ABI declared in include/nsg_b200.h).

There is no CPU or PyTorch fallback: if the shared library is missing or the
device is not a B200 the import / first call fails loudly.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libnsg_b200.so')

F32, BF16 = 0, 1
PRO_NONE, PRO_RELU, PRO_AFFINE_RELU, PRO_AFFINE = 0, 1, 2, 3
MASK_NONE, MASK_POS, MASK_AFFINE = 0, 1, 2
STATS_NONE, STATS_MOMENTS, STATS_BNBWD = 0, 1, 2

EXPORTS = [
    'nsg_version', 'nsg_last_error', 'nsg_device_ok', 'nsg_conv_gemm', 'nsg_conv_wgrad', 'nsg_bn_finalize',
    'nsg_bn_eval_coeffs', 'nsg_bn_apply', 'nsg_bn_bwd_reduce', 'nsg_bn_bwd_apply', 'nsg_relu_bwd_add',
    'nsg_colsum', 'nsg_vq_forward', 'nsg_scatter_add', 'nsg_vq_backward_inputs', 'nsg_recon_loss',
    'nsg_adam_step', 'nsg_tanh_bwd', 'nsg_gather_rows', 'nsg_vq_workspace_bytes',
    'nsg_conv_pack_bytes', 'nsg_conv_pack_weights', 'nsg_conv_gemm_tc', 'nsg_conv_wgrad_tc',
    'nsg_conv_pack_many', 'nsg_conv_gemm_tma', 'nsg_conv_wgrad_tma', 'nsg_wgrad_unpack_many',
    'nsg_conv_wgrad_slices', 'nsg_conv_wgrad_tma_sliced', 'nsg_pdl', 'nsg_recon_loss_bf16', 'nsg_loss_finalize',
    'nsg_bn_act64', 'nsg_bn_bwd64', 'nsg_relu64', 'nsg_vq_bwd_stats64', 'nsg_debug_timeline',
    'nsg_c1_conv_tc', 'nsg_c1_convT_tc', 'nsg_c1_wgrad_tc', 'nsg_scatter_mode', 'nsg_ema_update', 'nsg_vq_resident',
    'nsg_conv_bn_tma', 'nsg_conv_bn_tma_fits', 'nsg_c1_conv_bn_tc', 'nsg_mel_collate', 'nsg_peer_allreduce_stats',
    'nsg_hang_report', 'nsg_bn_act64_sync', 'nsg_bn_bwd64_sync', 'nsg_conv_rect', 'nsg_conv_rect_wgrad', 'nsg_gated_act', 'nsg_vae_reparam_kl',
    'nsg_vae_reparam_kl_bwd', 'nsg_sqerr_sum', 'nsg_softmax_sample', 'nsg_stft', 'nsg_melspectrogram', 'nsg_istft', 'nsg_istft_workspace_bytes', 'nsg_istft_static', 'nsg_istft_static_max_floats', 'nsg_griffin_lim',
    'nsg_griffin_lim_workspace_bytes', 'nsg_mel_to_linear', 'nsg_inv_preemphasis',
]


class ConvGeom(C.Structure):
    _fields_ = [('batch', C.c_int), ('in_h', C.c_int), ('in_w', C.c_int), ('cin', C.c_int),
                ('out_h', C.c_int), ('out_w', C.c_int), ('cout', C.c_int),
                ('ksize', C.c_int), ('stride', C.c_int), ('pad', C.c_int), ('gather', C.c_int),
                ('w_cin_stride', C.c_longlong), ('w_cout_stride', C.c_longlong), ('dtype', C.c_int)]


class Prologue(C.Structure):
    _fields_ = [('mode', C.c_int), ('scale', C.c_void_p), ('shift', C.c_void_p)]


class Epilogue(C.Structure):
    _fields_ = [('bias', C.c_void_p), ('add_src', C.c_void_p), ('mask_mode', C.c_int), ('mask_src', C.c_void_p),
                ('mscale', C.c_void_p), ('mshift', C.c_void_p), ('stats_mode', C.c_int), ('stats', C.c_void_p),
                ('xhat_src', C.c_void_p), ('mean', C.c_void_p), ('invstd', C.c_void_p), ('tanh_out', C.c_int)]


PEER_MAX, PEER_SLOTS = 8, 32
PEER_AREA_BYTES = PEER_SLOTS * PEER_MAX * 128 * 8 + PEER_SLOTS * PEER_MAX * 4


class PeerSync(C.Structure):
    _fields_ = [('world', C.c_int), ('rank', C.c_int), ('slot', C.c_int), ('areas', C.c_void_p * PEER_MAX)]


class BnFuse(C.Structure):
    _fields_ = [('mode', C.c_int), ('relu_out', C.c_int), ('gamma', C.c_void_p), ('beta', C.c_void_p),
                ('eps', C.c_float), ('momentum', C.c_float), ('count', C.c_double), ('running_mean', C.c_void_p),
                ('running_var', C.c_void_p), ('num_batches_tracked', C.c_void_p), ('coef', C.c_void_p),
                ('res', C.c_void_p), ('barrier', C.c_void_p), ('dgamma', C.c_void_p), ('dbeta', C.c_void_p),
                ('out2', C.c_void_p), ('peer', C.POINTER(PeerSync))]


class RectGeom(C.Structure):
    _fields_ = [('batch', C.c_int), ('in_h', C.c_int), ('in_w', C.c_int), ('cin', C.c_int),
                ('out_h', C.c_int), ('out_w', C.c_int), ('cout', C.c_int),
                ('kh', C.c_int), ('kw', C.c_int), ('pad_h', C.c_int), ('pad_w', C.c_int), ('gather', C.c_int),
                ('w_cin_stride', C.c_longlong), ('w_cout_stride', C.c_longlong)]


class NsgError(RuntimeError):
    pass


_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            'neural_sound_generation_b200: %s not found. Build it with `python -c "import __graft_entry__ as g; '
            'g.build()"` or `make -C neural_sound_generation_b200/csrc`. There is no CPU fallback.' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i, ll, f, d = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double
    lib.nsg_version.restype = C.c_char_p
    lib.nsg_last_error.restype = C.c_char_p
    lib.nsg_device_ok.argtypes = [C.c_char_p, i]
    lib.nsg_conv_gemm.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Prologue), C.POINTER(Epilogue), vp, i, vp]
    lib.nsg_conv_wgrad.argtypes = [C.POINTER(ConvGeom), vp, C.POINTER(Prologue), vp, C.POINTER(Prologue), vp, vp,
                                   i, vp]
    lib.nsg_conv_pack_bytes.argtypes = [C.POINTER(ConvGeom)]
    lib.nsg_conv_pack_weights.argtypes = [C.POINTER(ConvGeom), vp, vp, vp]
    lib.nsg_conv_gemm_tc.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Prologue), C.POINTER(Epilogue), vp, i, vp]
    lib.nsg_conv_wgrad_tc.argtypes = [C.POINTER(ConvGeom), vp, C.POINTER(Prologue), vp, C.POINTER(Prologue), vp, vp,
                                      i, vp]
    lib.nsg_debug_timeline.argtypes = [i, vp]
    lib.nsg_c1_conv_tc.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Epilogue), vp, vp]
    lib.nsg_c1_conv_bn_tc.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Epilogue), C.POINTER(BnFuse), vp, vp]
    lib.nsg_c1_convT_tc.argtypes = [C.POINTER(ConvGeom), vp, vp, vp, i, vp, vp]
    lib.nsg_c1_wgrad_tc.argtypes = [C.POINTER(ConvGeom), vp, vp, vp, vp, vp]
    lib.nsg_conv_pack_many.argtypes = [i, vp, vp, vp, vp, vp, vp]
    lib.nsg_conv_gemm_tma.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Epilogue), vp, vp]
    lib.nsg_conv_bn_tma.argtypes = [C.POINTER(ConvGeom), vp, vp, C.POINTER(Epilogue), C.POINTER(BnFuse), vp, vp]
    lib.nsg_conv_bn_tma_fits.argtypes = [C.POINTER(ConvGeom)]
    lib.nsg_conv_wgrad_tma.argtypes = [C.POINTER(ConvGeom), vp, vp, vp, vp, vp]
    lib.nsg_wgrad_unpack_many.argtypes = [i, vp, vp, vp, vp, vp, vp, vp]
    lib.nsg_pdl.argtypes = [i]
    lib.nsg_recon_loss_bf16.argtypes = [vp, vp, i, i, i, i, vp, vp, vp, vp, vp]
    lib.nsg_loss_finalize.argtypes = [vp, C.c_double, C.c_double, vp, vp]
    lib.nsg_conv_wgrad_slices.argtypes = [C.POINTER(ConvGeom)]
    lib.nsg_conv_wgrad_tma_sliced.argtypes = [C.POINTER(ConvGeom), vp, vp, vp, vp, vp]
    lib.nsg_bn_act64.argtypes = [vp, vp, vp, vp, d, vp, vp, f, f, vp, vp, vp, vp, vp, vp, vp, i, ll, vp]
    lib.nsg_bn_bwd64.argtypes = [vp, vp, vp, vp, d, vp, vp, vp, vp, vp, vp, ll, vp]
    lib.nsg_bn_act64_sync.argtypes = [vp, vp, vp, vp, d, vp, vp, f, f, vp, vp, vp, vp, vp, vp, vp, i, ll,
                                      C.POINTER(PeerSync), vp, vp]
    lib.nsg_bn_bwd64_sync.argtypes = [vp, vp, vp, vp, d, vp, vp, vp, vp, vp, vp, ll, C.POINTER(PeerSync), vp, vp]
    lib.nsg_relu64.argtypes = [vp, vp, ll, vp]
    lib.nsg_vq_bwd_stats64.argtypes = [vp, vp, vp, f, vp, vp, vp, vp, vp, ll, vp]
    lib.nsg_bn_finalize.argtypes = [vp, i, d, vp, vp, f, f, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.nsg_bn_eval_coeffs.argtypes = [i, vp, vp, vp, vp, f, vp, vp, vp]
    lib.nsg_bn_apply.argtypes = [vp, vp, vp, vp, i, i, vp, ll, i, i, vp]
    lib.nsg_bn_bwd_reduce.argtypes = [vp, vp, vp, vp, vp, vp, i, vp, ll, i, i, vp]
    lib.nsg_bn_bwd_apply.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, vp, d, vp, vp, vp, ll, i, i, vp]
    lib.nsg_relu_bwd_add.argtypes = [vp, vp, vp, vp, ll, i, vp]
    lib.nsg_colsum.argtypes = [vp, vp, ll, i, i, vp]
    lib.nsg_vq_forward.argtypes = [vp, vp, ll, i, i, i, i, vp, vp, vp, vp, ll, vp]
    lib.nsg_vq_workspace_bytes.argtypes = [i, i, i]
    lib.nsg_scatter_add.argtypes = [vp, vp, vp, ll, i, i, f, vp, vp, i, vp]
    lib.nsg_vq_backward_inputs.argtypes = [vp, vp, vp, f, vp, ll, i, vp]
    lib.nsg_recon_loss.argtypes = [vp, vp, i, i, i, i, vp, vp, vp]
    lib.nsg_gather_rows.argtypes = [vp, vp, ll, i, i, vp, i, vp]
    lib.nsg_scatter_mode.argtypes = [i]
    lib.nsg_vq_resident.argtypes = [i]
    lib.nsg_ema_update.argtypes = [vp, vp, vp, vp, vp, i, i, f, f, vp, vp]
    lib.nsg_tanh_bwd.argtypes = [vp, vp, vp, ll, vp]
    lib.nsg_mel_collate.argtypes = [vp, vp, i, i, i, vp, vp, vp]
    lib.nsg_peer_allreduce_stats.argtypes = [vp, C.POINTER(PeerSync), vp]
    lib.nsg_hang_report.argtypes = [vp, i]
    lib.nsg_conv_rect.argtypes = [C.POINTER(RectGeom), vp, vp, vp, vp, i, vp, vp, vp]
    lib.nsg_conv_rect_wgrad.argtypes = [C.POINTER(RectGeom), vp, vp, vp, vp, vp]
    lib.nsg_gated_act.argtypes = [vp, vp, vp, ll, i, ll, vp, vp]
    lib.nsg_vae_reparam_kl.argtypes = [vp, vp, ll, i, vp, vp, vp]
    lib.nsg_vae_reparam_kl_bwd.argtypes = [vp, vp, vp, f, ll, i, vp, vp]
    lib.nsg_sqerr_sum.argtypes = [vp, vp, ll, vp, f, i, vp, vp]
    lib.nsg_softmax_sample.argtypes = [vp, ll, i, ll, ll, vp, vp, ll, vp]
    lib.nsg_stft.argtypes = [vp, ll, i, i, f, vp, vp]
    lib.nsg_melspectrogram.argtypes = [vp, ll, i, i, f, vp, i, f, f, f, vp, vp]
    lib.nsg_istft_workspace_bytes.argtypes = [i, i]
    lib.nsg_istft.argtypes = [vp, i, i, i, vp, vp, vp]
    lib.nsg_istft_static.argtypes = [vp, i, i, i, vp, vp]
    lib.nsg_istft_static_max_floats.argtypes = []
    lib.nsg_griffin_lim_workspace_bytes.argtypes = [i, i]
    lib.nsg_griffin_lim.argtypes = [vp, vp, i, i, i, i, vp, vp, vp]
    lib.nsg_mel_to_linear.argtypes = [vp, i, i, vp, i, f, f, f, f, vp, vp]
    lib.nsg_inv_preemphasis.argtypes = [vp, ll, f, vp, vp]
    lib.nsg_adam_step.argtypes = [vp, vp, vp, vp, ll, f, f, f, f, f, vp, i, vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ('nsg_version', 'nsg_last_error'):
            fn.restype = ll if name in ('nsg_vq_workspace_bytes', 'nsg_conv_pack_bytes', 'nsg_istft_workspace_bytes', 'nsg_istft_static_max_floats',
                                        'nsg_griffin_lim_workspace_bytes') else i
    if os.environ.get('NSG_SCATTER_MODE') in ('0', '1'):      # development aid, see include/nsg_b200.h
        lib.nsg_scatter_mode(int(os.environ['NSG_SCATTER_MODE']))
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        raise NsgError('%s failed (%d): %s' % (what, rc, load().nsg_last_error().decode()))


_device_checked = False


def require_device():
    """Fail loudly unless a CUDA device of compute capability 10.x is current."""
    global _device_checked
    if _device_checked:
        return
    if not torch.cuda.is_available():
        raise NsgError('neural_sound_generation_b200 needs a CUDA device (B200, sm_100a); there is no CPU path')
    name = C.create_string_buffer(128)
    if load().nsg_device_ok(name, 128) != 1:
        raise NsgError(load().nsg_last_error().decode())
    _device_checked = True


def hang_report(clear=False):
    """Where a kernel's bounded wait gave up (include/nsg_b200.h: nsg_hang_report), or None."""
    w = (C.c_uint * 8)()
    load().nsg_hang_report(w, 1 if clear else 0)
    if w[0] == 0:
        return None
    unit = {1: 'conv_tc.cu', 2: 'conv_tma.cu', 3: 'vq_tc.cu', 4: 'fused_elem.cu'}.get(w[1], '?')
    return ('kernel wait timed out at %s:%d, block (%d, %d) of %d, thread %d, parity/value %#x'
            % (unit, w[2], w[3] & 0xFFFFF, w[3] >> 20, w[6], w[4], w[5]))


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def dtype_code(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise NsgError('unsupported activation dtype %s (float32 / bfloat16 only)' % t.dtype)
