"""Tensor-level wrappers over the C ABI (one Python function per entry point).

Every function takes CUDA tensors that are already laid out as the ABI wants
(NHWC activations, fp32 parameters) and launches on the current stream.  No
function here computes anything in PyTorch.
"""
import ctypes as C

import torch

from . import _cabi as A
from ._cabi import (F32, BF16, PRO_NONE, PRO_RELU, PRO_AFFINE_RELU, PRO_AFFINE, MASK_NONE, MASK_POS,  # noqa: F401
                    MASK_AFFINE, STATS_NONE, STATS_MOMENTS, STATS_BNBWD)

# number of kernel launches issued through this module (bench.py reports it as gpu_launches)
launch_count = 0


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise A.NsgError('nsg_b200 ops need CUDA tensors (got %s); there is no CPU path' % t.device)


def _contig(t, what):
    if t is not None and not t.is_contiguous():
        raise A.NsgError('%s must be contiguous' % what)


def conv_geom(batch, in_h, in_w, cin, out_h, out_w, cout, ksize, stride, pad, gather, w_cin_stride,
              w_cout_stride, dtype):
    return A.ConvGeom(batch, in_h, in_w, cin, out_h, out_w, cout, ksize, stride, pad, gather, w_cin_stride,
                      w_cout_stride, dtype)


def prologue(mode=PRO_NONE, scale=None, shift=None):
    return A.Prologue(mode, A.ptr(scale), A.ptr(shift))


def conv_pack(geom, w):
    """Packed bf16 (hi + lo) tensor-core image of conv weight ``w`` for the GEMM orientation in ``geom``."""
    global launch_count
    A.require_device()
    _cuda(w)
    _contig(w, 'conv weight')
    nbytes = A.load().nsg_conv_pack_bytes(C.byref(geom))
    if nbytes < 0:
        raise A.NsgError('nsg_conv_pack_weights failed (-3): the tcgen05 kernels serve cin == cout == 64')
    packed = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
    rc = A.load().nsg_conv_pack_weights(C.byref(geom), A.ptr(w), A.ptr(packed), A.stream_ptr())
    A.check(rc, 'nsg_conv_pack_weights')
    launch_count += 1
    return packed


def conv_gemm(geom, x, w, out, pro=None, bias=None, add_src=None, mask_mode=MASK_NONE, mask_src=None,
              mscale=None, mshift=None, stats_mode=STATS_NONE, stats=None, xhat_src=None, mean=None, invstd=None,
              tanh_out=False, impl=0, passes=1, packed=None):
    """impl 0: CUDA-core kernel reading ``w``; impl 1: tcgen05 kernel reading ``packed`` (built from w if None)."""
    global launch_count
    A.require_device()
    _cuda(x, w, out, bias, add_src, mask_src, stats, xhat_src)
    _contig(x, 'conv input')
    _contig(out, 'conv output')
    _contig(w, 'conv weight')
    epi = A.Epilogue(A.ptr(bias), A.ptr(add_src), mask_mode, A.ptr(mask_src), A.ptr(mscale), A.ptr(mshift),
                     stats_mode, A.ptr(stats), A.ptr(xhat_src), A.ptr(mean), A.ptr(invstd), int(bool(tanh_out)))
    p = pro if pro is not None else prologue()
    if impl == 1:
        if packed is None:
            packed = conv_pack(geom, w)
        rc = A.load().nsg_conv_gemm_tc(C.byref(geom), A.ptr(x), A.ptr(packed), C.byref(p), C.byref(epi), A.ptr(out),
                                       passes, A.stream_ptr())
        A.check(rc, 'nsg_conv_gemm_tc')
    else:
        rc = A.load().nsg_conv_gemm(C.byref(geom), A.ptr(x), A.ptr(w), C.byref(p), C.byref(epi), A.ptr(out), 0,
                                    A.stream_ptr())
        A.check(rc, 'nsg_conv_gemm')
    launch_count += (geom.stride * geom.stride) if geom.gather == 1 else 1
    return out


def conv_wgrad(geom, p, q, dw, pro_p=None, pro_q=None, q_colsum=None, impl=0, passes=1):
    global launch_count
    A.require_device()
    _cuda(p, q, dw, q_colsum)
    _contig(p, 'wgrad P')
    _contig(q, 'wgrad Q')
    _contig(dw, 'wgrad dw')
    pp = pro_p if pro_p is not None else prologue()
    pq = pro_q if pro_q is not None else prologue()
    if impl == 1:
        rc = A.load().nsg_conv_wgrad_tc(C.byref(geom), A.ptr(p), C.byref(pp), A.ptr(q), C.byref(pq), A.ptr(dw),
                                        A.ptr(q_colsum), passes, A.stream_ptr())
        A.check(rc, 'nsg_conv_wgrad_tc')
    else:
        rc = A.load().nsg_conv_wgrad(C.byref(geom), A.ptr(p), C.byref(pp), A.ptr(q), C.byref(pq), A.ptr(dw),
                                     A.ptr(q_colsum), 0, A.stream_ptr())
        A.check(rc, 'nsg_conv_wgrad')
    launch_count += 1


def bn_finalize(stats, count, gamma, beta, eps, momentum, running_mean, running_var, nbt, mean, invstd, scale,
                shift):
    global launch_count
    A.require_device()
    _cuda(stats, gamma, beta, mean)
    rc = A.load().nsg_bn_finalize(A.ptr(stats), gamma.numel(), float(count), A.ptr(gamma), A.ptr(beta), eps,
                                  momentum, A.ptr(running_mean), A.ptr(running_var), A.ptr(nbt), A.ptr(mean),
                                  A.ptr(invstd), A.ptr(scale), A.ptr(shift), A.stream_ptr())
    A.check(rc, 'nsg_bn_finalize')
    launch_count += 1


def bn_eval_coeffs(gamma, beta, running_mean, running_var, eps, scale, shift):
    global launch_count
    A.require_device()
    _cuda(gamma, beta, running_mean, running_var, scale, shift)
    rc = A.load().nsg_bn_eval_coeffs(gamma.numel(), A.ptr(gamma), A.ptr(beta), A.ptr(running_mean),
                                     A.ptr(running_var), eps, A.ptr(scale), A.ptr(shift), A.stream_ptr())
    A.check(rc, 'nsg_bn_eval_coeffs')
    launch_count += 1


def bn_apply(h, scale, shift, y, res=None, res_relu=False, relu_out=False):
    global launch_count
    A.require_device()
    _cuda(h, scale, shift, y, res)
    c = h.shape[-1]
    rc = A.load().nsg_bn_apply(A.ptr(h), A.ptr(scale), A.ptr(shift), A.ptr(res), int(res_relu), int(relu_out),
                               A.ptr(y), h.numel() // c, c, A.dtype_code(h), A.stream_ptr())
    A.check(rc, 'nsg_bn_apply')
    launch_count += 1
    return y


def bn_bwd_reduce(g, h, mean, invstd, stats, scale=None, shift=None, relu_mask=False):
    global launch_count
    A.require_device()
    _cuda(g, h, mean, invstd, stats)
    c = h.shape[-1]
    rc = A.load().nsg_bn_bwd_reduce(A.ptr(g), A.ptr(h), A.ptr(mean), A.ptr(invstd), A.ptr(scale), A.ptr(shift),
                                    int(relu_mask), A.ptr(stats), h.numel() // c, c, A.dtype_code(h),
                                    A.stream_ptr())
    A.check(rc, 'nsg_bn_bwd_reduce')
    launch_count += 1


def bn_bwd_apply(g, h, mean, invstd, gamma, stats, count, dh, dgamma=None, dbeta=None, scale=None, shift=None,
                 relu_mask=False):
    global launch_count
    A.require_device()
    _cuda(g, h, mean, invstd, gamma, stats, dh)
    c = h.shape[-1]
    rc = A.load().nsg_bn_bwd_apply(A.ptr(g), A.ptr(h), A.ptr(mean), A.ptr(invstd), A.ptr(gamma), A.ptr(scale),
                                   A.ptr(shift), int(relu_mask), A.ptr(stats), float(count), A.ptr(dh),
                                   A.ptr(dgamma), A.ptr(dbeta), h.numel() // c, c, A.dtype_code(h),
                                   A.stream_ptr())
    A.check(rc, 'nsg_bn_bwd_apply')
    launch_count += 1
    return dh


def relu_bwd_add(g1, g2, x, dx):
    global launch_count
    A.require_device()
    _cuda(g1, g2, x, dx)
    rc = A.load().nsg_relu_bwd_add(A.ptr(g1), A.ptr(g2), A.ptr(x), A.ptr(dx), x.numel(), A.dtype_code(x),
                                   A.stream_ptr())
    A.check(rc, 'nsg_relu_bwd_add')
    launch_count += 1
    return dx


def colsum(g, out):
    global launch_count
    A.require_device()
    _cuda(g, out)
    c = g.shape[-1]
    rc = A.load().nsg_colsum(A.ptr(g), A.ptr(out), g.numel() // c, c, A.dtype_code(g), A.stream_ptr())
    A.check(rc, 'nsg_colsum')
    launch_count += 1


def vq_mode_serves(k, d, mode):
    return mode == 0 or A.load().nsg_vq_workspace_bytes(k, d, mode) >= 0


def vq_forward(z, codebook, indices, codes=None, sqerr=None, mode=0):
    """z (..., D) contiguous activations; codebook (K, D) fp32; indices int64 (N,)."""
    global launch_count
    A.require_device()
    _cuda(z, codebook, indices, codes, sqerr)
    _contig(z, 'vq input')
    _contig(codebook, 'codebook')
    k, d = codebook.shape
    n = z.numel() // d
    if codebook.dtype != torch.float32:
        raise A.NsgError('codebook must be float32')
    ws, ws_bytes = None, 0
    if mode != 0:
        ws_bytes = A.load().nsg_vq_workspace_bytes(k, d, mode)
        if ws_bytes < 0:
            raise A.NsgError('nsg_vq_forward failed (-3): mode %d does not serve K=%d, D=%d' % (mode, k, d))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=z.device)
    rc = A.load().nsg_vq_forward(A.ptr(z), A.ptr(codebook), n, k, d, A.dtype_code(z), mode, A.ptr(indices),
                                 A.ptr(codes), A.ptr(sqerr), A.ptr(ws), ws_bytes, A.stream_ptr())
    A.check(rc, 'nsg_vq_forward')
    launch_count += 1 if mode == 0 else 2


def scatter_add(rows, indices, table, alpha=1.0, sub=None, counts=None):
    global launch_count
    A.require_device()
    _cuda(rows, indices, table, sub, counts)
    _contig(rows, 'scatter rows')
    k, d = table.shape
    n = rows.numel() // d
    rc = A.load().nsg_scatter_add(A.ptr(rows), A.ptr(sub), A.ptr(indices), n, k, d, float(alpha), A.ptr(table),
                                  A.ptr(counts), A.dtype_code(rows), A.stream_ptr())
    A.check(rc, 'nsg_scatter_add')
    launch_count += 1


def ema_update(counts, sums, ema_count, ema_sum, codebook, decay=0.99, eps=1e-5):
    """EMA codebook update from one step's scatter_add statistics (extension; include/nsg_b200.h)."""
    global launch_count
    A.require_device()
    _cuda(counts, sums, ema_count, ema_sum, codebook)
    for t in (counts, sums, ema_count, ema_sum, codebook):
        if t.dtype != torch.float32 or not t.is_contiguous():
            raise A.NsgError('ema_update: fp32 contiguous tensors required')
    k, d = codebook.shape
    scratch = torch.empty(1, dtype=torch.float64, device=codebook.device)
    rc = A.load().nsg_ema_update(A.ptr(counts), A.ptr(sums), A.ptr(ema_count), A.ptr(ema_sum), A.ptr(codebook), k, d,
                                 float(decay), float(eps), A.ptr(scratch), A.stream_ptr())
    A.check(rc, 'nsg_ema_update')
    launch_count += 2


def vq_backward_inputs(g_st, z, zq, coef, dz):
    global launch_count
    A.require_device()
    _cuda(g_st, z, zq, dz)
    rc = A.load().nsg_vq_backward_inputs(A.ptr(g_st), A.ptr(z), A.ptr(zq), float(coef), A.ptr(dz), z.numel(),
                                         A.dtype_code(z), A.stream_ptr())
    A.check(rc, 'nsg_vq_backward_inputs')
    launch_count += 1
    return dz


def recon_loss(x_tilde, c, loss_sum, dpre=None):
    """x_tilde (B, H, Wt) fp32, c (B, H, W) fp32, W >= Wt (right zero pad, train.py:118-120)."""
    global launch_count
    A.require_device()
    _cuda(x_tilde, c, loss_sum, dpre)
    _contig(x_tilde, 'x_tilde')
    _contig(c, 'target mel')
    b, h, w = c.shape
    wt = x_tilde.shape[-1]
    rc = A.load().nsg_recon_loss(A.ptr(x_tilde), A.ptr(c), b, h, w, wt, A.ptr(loss_sum), A.ptr(dpre),
                                 A.stream_ptr())
    A.check(rc, 'nsg_recon_loss')
    launch_count += 1


def recon_loss_bf16(x_tilde, c, loss_sum, dpre=None, x_tilde_f32=None, dpre_sum=None):
    """x_tilde (B, H, Wt) bf16, c (B, H, W) fp32; dpre (bf16) and x_tilde_f32 (fp32) optional outputs (B, H, Wt);
    dpre_sum (fp32, one element) += sum of dpre (the output layer's bias gradient)."""
    global launch_count
    A.require_device()
    _cuda(x_tilde, c, loss_sum, dpre, x_tilde_f32, dpre_sum)
    if dpre_sum is not None and dpre_sum.dtype != torch.float32:
        raise TypeError('recon_loss_bf16: dpre_sum must be fp32')
    _bf16(x_tilde, dpre)
    _contig(x_tilde, 'x_tilde')
    _contig(c, 'target mel')
    if c.dtype != torch.float32 or loss_sum.dtype != torch.float64:
        raise TypeError('recon_loss_bf16: c must be fp32 and loss_sum fp64')
    if x_tilde_f32 is not None and (x_tilde_f32.dtype != torch.float32 or not x_tilde_f32.is_contiguous()):
        raise TypeError('recon_loss_bf16: x_tilde_f32 must be a contiguous fp32 tensor')
    b, h, w = c.shape
    wt = x_tilde.shape[-1]
    rc = A.load().nsg_recon_loss_bf16(A.ptr(x_tilde), A.ptr(c), b, h, w, wt, A.ptr(loss_sum), A.ptr(dpre),
                                      A.ptr(x_tilde_f32), A.ptr(dpre_sum), A.stream_ptr())
    A.check(rc, 'nsg_recon_loss_bf16')
    launch_count += 1


def loss_finalize(sums, recon_scale, vq_scale, out=None):
    """[sums[0]*recon_scale, sums[1]*vq_scale, sums[1]*vq_scale] as one fp64 tensor (train.py:129-134)."""
    global launch_count
    A.require_device()
    if out is None:
        out = torch.empty(3, dtype=torch.float64, device=sums.device)
    _cuda(sums, out)
    if sums.dtype != torch.float64 or out.dtype != torch.float64 or sums.numel() < 2 or out.numel() < 3:
        raise TypeError('loss_finalize: fp64 tensors of >= 2 and >= 3 elements')
    rc = A.load().nsg_loss_finalize(A.ptr(sums), float(recon_scale), float(vq_scale), A.ptr(out), A.stream_ptr())
    A.check(rc, 'nsg_loss_finalize')
    launch_count += 1
    return out


def adam_step(params, grads, exp_avg, exp_avg_sq, step_count, lr, beta1=0.9, beta2=0.999, eps=1e-8,
              grad_scale=1.0, zero_grads=False, ticket=None):
    """ticket: a zeroed int32 scalar owned by the optimiser -> the step counter advances inside the same launch."""
    global launch_count
    A.require_device()
    _cuda(params, grads, exp_avg, exp_avg_sq, step_count, ticket)
    rc = A.load().nsg_adam_step(A.ptr(params), A.ptr(grads), A.ptr(exp_avg), A.ptr(exp_avg_sq), params.numel(),
                                lr, beta1, beta2, eps, grad_scale, A.ptr(step_count), 1 if zero_grads else 0,
                                A.ptr(ticket), A.stream_ptr())
    A.check(rc, 'nsg_adam_step')
    launch_count += 1 if (ticket is not None and params.numel() > 0) else 2


def tanh_bwd(g, y, out):
    global launch_count
    A.require_device()
    _cuda(g, y, out)
    rc = A.load().nsg_tanh_bwd(A.ptr(g), A.ptr(y), A.ptr(out), y.numel(), A.stream_ptr())
    A.check(rc, 'nsg_tanh_bwd')
    launch_count += 1
    return out


def gather_rows(table, indices, out):
    """out (N, D) [fp32/bf16] = table (K, D) fp32 rows selected by int64 indices (N,)."""
    global launch_count
    A.require_device()
    _cuda(table, indices, out)
    _contig(table, 'gather table')
    k, d = table.shape
    rc = A.load().nsg_gather_rows(A.ptr(table), A.ptr(indices), indices.numel(), k, d, A.ptr(out),
                                  A.dtype_code(out), A.stream_ptr())
    A.check(rc, 'nsg_gather_rows')
    launch_count += 1
    return out


# ---------------------------------------------------------------------------------------------- TMA-fed bf16 path
def _bf16(*ts):
    for t in ts:
        if t is not None and t.dtype != torch.bfloat16:
            raise A.NsgError('the TMA kernels take bfloat16 tensors (got %s)' % t.dtype)


def _ptr_array(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


def conv_pack_many(items):
    """items: list of (w fp32, packed uint8 view of k*k*8192 bytes, ntaps, w_cin_stride, w_cout_stride); one launch."""
    global launch_count
    A.require_device()
    if not items:
        return
    n = len(items)
    rc = A.load().nsg_conv_pack_many(
        n, _ptr_array([it[0] for it in items]), _ptr_array([it[1] for it in items]),
        (C.c_int * n)(*[it[2] for it in items]), (C.c_longlong * n)(*[it[3] for it in items]),
        (C.c_longlong * n)(*[it[4] for it in items]), A.stream_ptr())
    A.check(rc, 'nsg_conv_pack_many')
    launch_count += 1


def conv_pack_tma(geom, w):
    """bf16 weight image (k*k*8192 bytes) of ``w`` for the GEMM orientation in ``geom``."""
    _cuda(w)
    _contig(w, 'conv weight')
    ntaps = geom.ksize * geom.ksize
    packed = torch.empty(ntaps * 8192, dtype=torch.uint8, device=w.device)
    conv_pack_many([(w, packed, ntaps, geom.w_cin_stride, geom.w_cout_stride)])
    return packed


def conv_gemm_tma(geom, x, packed, out, bias=None, add_src=None, mask_mode=MASK_NONE, mask_src=None, mscale=None,
                  mshift=None, stats_mode=STATS_NONE, stats=None, xhat_src=None, mean=None, invstd=None,
                  relu_out=False):
    global launch_count
    A.require_device()
    _cuda(x, packed, out, bias, add_src, mask_src, stats, xhat_src)
    _bf16(x, out, add_src, mask_src, xhat_src)
    _contig(x, 'conv input')
    _contig(out, 'conv output')
    epi = A.Epilogue(A.ptr(bias), A.ptr(add_src), mask_mode, A.ptr(mask_src), A.ptr(mscale), A.ptr(mshift),
                     stats_mode, A.ptr(stats), A.ptr(xhat_src), A.ptr(mean), A.ptr(invstd), 2 if relu_out else 0)
    rc = A.load().nsg_conv_gemm_tma(C.byref(geom), A.ptr(x), A.ptr(packed), C.byref(epi), A.ptr(out), A.stream_ptr())
    A.check(rc, 'nsg_conv_gemm_tma')
    launch_count += 1
    return out


BNF_TRAIN, BNF_EVAL, BNF_BWD = 1, 2, 3
_fits_cache = {}


def conv_bn_tma_fits(geom):
    """True when a BatchNorm-fused (grid-barrier) launch of this geometry fits: <= 8 tiles per CTA on this device."""
    A.require_device()
    key = (torch.cuda.current_device(),) + tuple(getattr(geom, f) for f, _ in geom._fields_)
    r = _fits_cache.get(key)
    if r is None:
        r = _fits_cache[key] = A.load().nsg_conv_bn_tma_fits(C.byref(geom)) == 1
    return r


def conv_bn_tma(geom, x, packed, out, bn_mode, gamma, beta=None, eps=1e-5, momentum=0.1, count=1.0, running_mean=None,
                running_var=None, nbt=None, coef=None, res=None, barrier=None, dgamma=None, dbeta=None, out2=None,
                bn_relu=False, bias=None, add_src=None, mask_mode=MASK_NONE, mask_src=None, stats=None,
                xhat_src=None, mean=None, invstd=None, relu_out=False, peer=None):
    """nsg_conv_bn_tma: the TMA convolution with the BatchNorm that follows it (forward) / that its output gradient
    flows into (backward) fused; see include/nsg_b200.h (nsg_bn_fuse).  ``out`` may be None for TRAIN / BWD."""
    global launch_count
    A.require_device()
    _cuda(x, packed, out, out2, bias, add_src, mask_src, stats, xhat_src, res, barrier)
    _bf16(x, out, out2, add_src, mask_src, xhat_src, res)
    _contig(x, 'conv input')
    for t, what in ((out, 'conv output'), (out2, 'second output'), (res, 'residual')):
        if t is not None:
            _contig(t, what)
    stats_mode = {BNF_TRAIN: STATS_MOMENTS, BNF_BWD: STATS_BNBWD}.get(bn_mode, STATS_NONE)
    epi = A.Epilogue(A.ptr(bias), A.ptr(add_src), mask_mode, A.ptr(mask_src), None, None, stats_mode, A.ptr(stats),
                     A.ptr(xhat_src), A.ptr(mean), A.ptr(invstd), 2 if relu_out else 0)
    bn = A.BnFuse(bn_mode, 1 if bn_relu else 0, A.ptr(gamma), A.ptr(beta), eps, momentum, float(count),
                  A.ptr(running_mean), A.ptr(running_var), A.ptr(nbt), A.ptr(coef), A.ptr(res), A.ptr(barrier),
                  A.ptr(dgamma), A.ptr(dbeta), A.ptr(out2), C.pointer(peer) if peer is not None else None)
    rc = A.load().nsg_conv_bn_tma(C.byref(geom), A.ptr(x), A.ptr(packed), C.byref(epi), C.byref(bn), A.ptr(out),
                                  A.stream_ptr())
    A.check(rc, 'nsg_conv_bn_tma')
    launch_count += 1


def conv_wgrad_slices(geom):
    """Number of partial images conv_wgrad_tma(..., sliced=True) writes for this geometry on the current device."""
    A.require_device()
    s = A.load().nsg_conv_wgrad_slices(C.byref(geom))
    if s <= 0:
        raise RuntimeError('nsg_conv_wgrad_slices: geometry not served by the TMA weight-gradient kernel')
    return s


def conv_wgrad_tma(geom, p, q, dw_packed, q_colsum=None, sliced=False):
    """sliced=False: dw_packed [k*k,64,64] += (atomics).  sliced=True: dw_packed [conv_wgrad_slices(geom),k*k,64,64]
    is overwritten with per-CTA partials (summed by wgrad_unpack_many)."""
    global launch_count
    A.require_device()
    _cuda(p, q, dw_packed, q_colsum)
    _bf16(p, q)
    _contig(p, 'wgrad P')
    _contig(q, 'wgrad Q')
    if sliced:
        need = conv_wgrad_slices(geom) * geom.ksize * geom.ksize * 4096
        if dw_packed.numel() < need or dw_packed.dtype != torch.float32:
            raise ValueError('conv_wgrad_tma: sliced image needs %d fp32 elements, got %d' % (need, dw_packed.numel()))
        fn, name = A.load().nsg_conv_wgrad_tma_sliced, 'nsg_conv_wgrad_tma_sliced'
    else:
        fn, name = A.load().nsg_conv_wgrad_tma, 'nsg_conv_wgrad_tma'
    rc = fn(C.byref(geom), A.ptr(p), A.ptr(q), A.ptr(dw_packed), A.ptr(q_colsum), A.stream_ptr())
    A.check(rc, name)
    launch_count += 1


def wgrad_unpack_many(items):
    """items: list of (dw_packed fp32 [S,k*k,64,64], dw fp32 parameter-layout gradient, ntaps, w_cin_stride,
    w_cout_stride[, S = 1])."""
    global launch_count
    A.require_device()
    if not items:
        return
    n = len(items)
    rc = A.load().nsg_wgrad_unpack_many(
        n, _ptr_array([it[0] for it in items]), _ptr_array([it[1] for it in items]),
        (C.c_int * n)(*[it[2] for it in items]), (C.c_int * n)(*[(it[5] if len(it) > 5 else 1) for it in items]),
        (C.c_longlong * n)(*[it[3] for it in items]), (C.c_longlong * n)(*[it[4] for it in items]), A.stream_ptr())
    A.check(rc, 'nsg_wgrad_unpack_many')
    launch_count += 1


def bn_act64(h, out, gamma, beta, stats=None, count=1.0, res=None, relu_out=True, eps=1e-5, momentum=0.1,
             running_mean=None, running_var=None, nbt=None, coef=None, peer=None, gate=None):
    """coef: optional (mean, invstd, scale, shift) fp32 [64] tensors to publish.  peer (ops.peer_sync) + gate (one
    zeroed int32 counter): the statistics are summed over the GPUs inside the launch (``count`` = global count)."""
    global launch_count
    A.require_device()
    _cuda(h, out, gamma, beta, stats, res, running_mean, running_var)
    _bf16(h, out, res)
    if h.shape[-1] != 64:
        raise A.NsgError('bn_act64 serves 64-channel tensors')
    cf = coef if coef is not None else (None, None, None, None)
    rc = A.load().nsg_bn_act64_sync(A.ptr(h), A.ptr(res), A.ptr(out), A.ptr(stats), float(count), A.ptr(gamma),
                                    A.ptr(beta), eps, momentum, A.ptr(running_mean), A.ptr(running_var), A.ptr(nbt),
                                    A.ptr(cf[0]), A.ptr(cf[1]), A.ptr(cf[2]), A.ptr(cf[3]), int(bool(relu_out)),
                                    h.numel() // 64, C.pointer(peer) if peer is not None else None, A.ptr(gate),
                                    A.stream_ptr())
    A.check(rc, 'nsg_bn_act64')
    launch_count += 1
    return out


def bn_bwd64(g, h, dh, stats, count, mean, invstd, gamma, dgamma=None, dbeta=None, dbias_prev=None, peer=None,
             gate=None):
    global launch_count
    A.require_device()
    _cuda(g, h, dh, stats, mean, invstd, gamma)
    _bf16(g, h, dh)
    rc = A.load().nsg_bn_bwd64_sync(A.ptr(g), A.ptr(h), A.ptr(dh), A.ptr(stats), float(count), A.ptr(mean),
                                    A.ptr(invstd), A.ptr(gamma), A.ptr(dgamma), A.ptr(dbeta), A.ptr(dbias_prev),
                                    h.numel() // 64, C.pointer(peer) if peer is not None else None, A.ptr(gate),
                                    A.stream_ptr())
    A.check(rc, 'nsg_bn_bwd64')
    launch_count += 1
    return dh


def relu64(x, out):
    global launch_count
    A.require_device()
    _cuda(x, out)
    _bf16(x, out)
    rc = A.load().nsg_relu64(A.ptr(x), A.ptr(out), x.numel(), A.stream_ptr())
    A.check(rc, 'nsg_relu64')
    launch_count += 1
    return out


def vq_bwd_stats64(g_st, z, zq, coef, dz, h=None, mean=None, invstd=None, stats=None):
    global launch_count
    A.require_device()
    _cuda(g_st, z, zq, dz, h, stats)
    _bf16(g_st, z, zq, dz, h)
    rc = A.load().nsg_vq_bwd_stats64(A.ptr(g_st), A.ptr(z), A.ptr(zq), float(coef), A.ptr(dz), A.ptr(h),
                                     A.ptr(mean), A.ptr(invstd), A.ptr(stats), z.numel() // 64, A.stream_ptr())
    A.check(rc, 'nsg_vq_bwd_stats64')
    launch_count += 1
    return dz


def c1_pack(w, kind):
    """8 KB operand image of a single-channel layer's weight: kind -1 (rows = 64 channels, K = 16 taps) for
    c1_conv_tc, kind -2 (rows = 16 taps, K = 64 channels) for c1_convT_tc.  Both weights are (64, 1, 4, 4)-shaped."""
    packed = torch.empty(8192, dtype=torch.uint8, device=w.device)
    conv_pack_many([(w, packed, kind, 1, 16) if kind == -1 else (w, packed, kind, 16, 1)])
    return packed


def c1_conv_tc(geom, x1, packed, out, bias=None, add_src=None, mask_mode=MASK_NONE, mask_src=None, mscale=None,
               mshift=None, stats_mode=STATS_NONE, stats=None, xhat_src=None, mean=None, invstd=None):
    global launch_count
    A.require_device()
    _cuda(x1, packed, out, bias, add_src, mask_src, stats, xhat_src)
    _bf16(x1, out, add_src, mask_src, xhat_src)
    epi = A.Epilogue(A.ptr(bias), A.ptr(add_src), mask_mode, A.ptr(mask_src), A.ptr(mscale), A.ptr(mshift),
                     stats_mode, A.ptr(stats), A.ptr(xhat_src), A.ptr(mean), A.ptr(invstd), 0)
    rc = A.load().nsg_c1_conv_tc(C.byref(geom), A.ptr(x1), A.ptr(packed), C.byref(epi), A.ptr(out), A.stream_ptr())
    A.check(rc, 'nsg_c1_conv_tc')
    launch_count += 1
    return out


def c1_conv_bn_eval_tc(geom, x1, packed, out, gamma, beta, running_mean, running_var, eps=1e-5, bias=None, relu_out=True):
    """nsg_c1_conv_bn_tc: Conv2d(1 -> 64, 4, 2, 1) + eval-mode BatchNorm [+ ReLU] in one launch (models.py:165-167)."""
    global launch_count
    A.require_device()
    _cuda(x1, packed, out, bias, gamma, beta, running_mean, running_var)
    _bf16(x1, out)
    epi = A.Epilogue(A.ptr(bias), None, MASK_NONE, None, None, None, STATS_NONE, None, None, None, None,
                     2 if relu_out else 0)
    bn = A.BnFuse(BNF_EVAL, 0, A.ptr(gamma), A.ptr(beta), eps, 0.0, 1.0, A.ptr(running_mean), A.ptr(running_var), None,
                  None, None, None, None, None, None, None)
    rc = A.load().nsg_c1_conv_bn_tc(C.byref(geom), A.ptr(x1), A.ptr(packed), C.byref(epi), C.byref(bn), A.ptr(out),
                                    A.stream_ptr())
    A.check(rc, 'nsg_c1_conv_bn_tc')
    launch_count += 1
    return out


def c1_convT_tc(geom, x, packed, out1, bias=None, tanh_out=False):
    global launch_count
    A.require_device()
    _cuda(x, packed, out1, bias)
    _bf16(x, out1)
    rc = A.load().nsg_c1_convT_tc(C.byref(geom), A.ptr(x), A.ptr(packed), A.ptr(bias), int(bool(tanh_out)),
                                  A.ptr(out1), A.stream_ptr())
    A.check(rc, 'nsg_c1_convT_tc')
    launch_count += 1
    return out1


def c1_wgrad_tc(geom, p1, q, dw, q_colsum=None):
    global launch_count
    A.require_device()
    _cuda(p1, q, dw, q_colsum)
    _bf16(p1, q)
    rc = A.load().nsg_c1_wgrad_tc(C.byref(geom), A.ptr(p1), A.ptr(q), A.ptr(dw), A.ptr(q_colsum), A.stream_ptr())
    A.check(rc, 'nsg_c1_wgrad_tc')
    launch_count += 1


def mel_collate(frames, offsets, n_mels, t_out, out_f32=None, out_bf16=None):
    """nsg_mel_collate: frames fp32 (sum_len, n_mels) device tensor, offsets int32 (B + 1) device tensor (in frames)."""
    global launch_count
    A.require_device()
    _cuda(frames, offsets, out_f32, out_bf16)
    if frames.dtype != torch.float32 or offsets.dtype != torch.int32:
        raise TypeError('mel_collate: frames must be fp32 and offsets int32')
    if out_bf16 is not None:
        _bf16(out_bf16)
    _contig(frames, 'frames')
    rc = A.load().nsg_mel_collate(A.ptr(frames), A.ptr(offsets), offsets.numel() - 1, n_mels, t_out, A.ptr(out_f32),
                                  A.ptr(out_bf16), A.stream_ptr())
    A.check(rc, 'nsg_mel_collate')
    launch_count += 1


def peer_sync(world, rank, slot, area_ptrs):
    """nsg_peer_sync for one exchange: ``area_ptrs`` = every rank's exchange area as mapped on this GPU (ints)."""
    ps = A.PeerSync()
    ps.world, ps.rank, ps.slot = world, rank, slot
    for i, a in enumerate(area_ptrs):
        ps.areas[i] = a
    return ps


def peer_allreduce_stats(sums, peer):
    """sums (128 fp64, device) := sum over the ranks of the node through peer memory (no NCCL launch)."""
    global launch_count
    A.require_device()
    _cuda(sums)
    if sums.dtype != torch.float64 or sums.numel() != 128:
        raise TypeError('peer_allreduce_stats: 128 fp64 sums expected')
    rc = A.load().nsg_peer_allreduce_stats(A.ptr(sums), C.byref(peer), A.stream_ptr())
    A.check(rc, 'nsg_peer_allreduce_stats')
    launch_count += 1


# ---------------------------------------------------------------------------------------------- row N4 (VAE / PixelCNN)
def _f32(*ts):
    for t in ts:
        if t is not None and t.dtype != torch.float32:
            raise A.NsgError('this entry point takes float32 tensors (got %s)' % t.dtype)


def conv_rect(x, w, out, kh, kw, pad_h, pad_w, gather, w_cin_stride, w_cout_stride, bias=None, add_src=None, act=0,
              stats=None):
    """nsg_conv_rect: stride-1 convolution with a (kh, kw) kernel over NHWC fp32 ``x`` into ``out`` (whose height / width
    may be the cropped leading part of the full result).  gather 0 = Conv2d-like, 1 = ConvTranspose2d-like."""
    global launch_count
    A.require_device()
    _cuda(x, w, out, bias, add_src, stats)
    _f32(x, w, out, bias, add_src)
    for t, what in ((x, 'conv input'), (w, 'conv weight'), (out, 'conv output'), (add_src, 'add_src')):
        _contig(t, what)
    b, ih, iw, cin = x.shape
    _, oh, ow, cout = out.shape
    g = A.RectGeom(b, ih, iw, cin, oh, ow, cout, kh, kw, pad_h, pad_w, gather, w_cin_stride, w_cout_stride)
    rc = A.load().nsg_conv_rect(C.byref(g), A.ptr(x), A.ptr(w), A.ptr(bias), A.ptr(add_src), act, A.ptr(stats), A.ptr(out),
                                A.stream_ptr())
    A.check(rc, 'nsg_conv_rect')
    launch_count += 1
    return out


def conv_rect_wgrad(p, q, dw, kh, kw, pad_h, pad_w, w_cin_stride, w_cout_stride, q_colsum=None):
    """dw[tap + a*w_cin_stride + b*w_cout_stride] += sum over q's pixels of p[tap-shifted pixel, a] * q[pixel, b]."""
    global launch_count
    A.require_device()
    _cuda(p, q, dw, q_colsum)
    _f32(p, q, dw, q_colsum)
    _contig(p, 'wgrad P')
    _contig(q, 'wgrad Q')
    b, ih, iw, ca = p.shape
    _, oh, ow, cb = q.shape
    g = A.RectGeom(b, ih, iw, ca, oh, ow, cb, kh, kw, pad_h, pad_w, 0, w_cin_stride, w_cout_stride)
    rc = A.load().nsg_conv_rect_wgrad(C.byref(g), A.ptr(p), A.ptr(q), A.ptr(dw), A.ptr(q_colsum), A.stream_ptr())
    A.check(rc, 'nsg_conv_rect_wgrad')
    launch_count += 1


def gated_act(x, out, add2=None, cond=None):
    """out[..., c] = tanh(a) * sigmoid(b) with (a | b) = halves of x + add2 + cond[item]; x (B, H, W, 2C), cond (B, 2C)."""
    global launch_count
    A.require_device()
    _cuda(x, out, add2, cond)
    _f32(x, out, add2, cond)
    for t, what in ((x, 'gate input'), (out, 'gate output'), (add2, 'add2'), (cond, 'cond')):
        _contig(t, what)
    c = out.shape[-1]
    pixels = out.numel() // c
    rc = A.load().nsg_gated_act(A.ptr(x), A.ptr(add2), A.ptr(cond), pixels, c, pixels // x.shape[0], A.ptr(out),
                                A.stream_ptr())
    A.check(rc, 'nsg_gated_act')
    launch_count += 1
    return out


def vae_reparam_kl(stats, eps, z, kl_sum):
    global launch_count
    A.require_device()
    _cuda(stats, eps, z, kl_sum)
    _f32(stats, eps, z)
    zc = z.shape[-1]
    rc = A.load().nsg_vae_reparam_kl(A.ptr(stats), A.ptr(eps), z.numel() // zc, zc, A.ptr(z), A.ptr(kl_sum), A.stream_ptr())
    A.check(rc, 'nsg_vae_reparam_kl')
    launch_count += 1


def vae_reparam_kl_bwd(stats, eps, dz, kl_scale, dstats):
    global launch_count
    A.require_device()
    _cuda(stats, eps, dz, dstats)
    _f32(stats, eps, dz, dstats)
    zc = eps.shape[-1]
    rc = A.load().nsg_vae_reparam_kl_bwd(A.ptr(stats), A.ptr(eps), A.ptr(dz), float(kl_scale), eps.numel() // zc, zc,
                                         A.ptr(dstats), A.stream_ptr())
    A.check(rc, 'nsg_vae_reparam_kl_bwd')
    launch_count += 1


def sqerr_sum(a, b, total=None, grad=None, grad_scale=1.0, through_tanh=False):
    global launch_count
    A.require_device()
    _cuda(a, b, total, grad)
    _f32(a, b, grad)
    _contig(a, 'sqerr a')
    _contig(b, 'sqerr b')
    rc = A.load().nsg_sqerr_sum(A.ptr(a), A.ptr(b), a.numel(), A.ptr(total), float(grad_scale), int(through_tanh),
                                A.ptr(grad), A.stream_ptr())
    A.check(rc, 'nsg_sqerr_sum')
    launch_count += 1


def softmax_sample(logits, row_stride, offset, rows, classes, u, out, out_stride):
    global launch_count
    A.require_device()
    _cuda(logits, u, out)
    _f32(logits, u)
    if out.dtype != torch.int64:
        raise A.NsgError('softmax_sample: int64 output expected')
    rc = A.load().nsg_softmax_sample(A.ptr(logits), rows, classes, row_stride, offset, A.ptr(u), A.ptr(out), out_stride,
                                     A.stream_ptr())
    A.check(rc, 'nsg_softmax_sample')
    launch_count += 1
