"""Explicit forward / backward schedule of the VQ-VAE step over the C-ABI kernels.

This is the host-side mirror of what autograd does for the reference
``VQVAE`` (/root/reference/src/models.py:161-216): a fixed sequence of kernel
launches over NHWC activation buffers, with BatchNorm statistics produced in
the convolution epilogues and BatchNorm-apply / ReLU consumed in the next
convolution's prologue.  Parameters and gradients are addressed by the
reference's own ``state_dict`` keys.

Nothing here computes in PyTorch; torch only owns the memory.
"""
import torch

from . import ops
from .ops import (PRO_NONE, PRO_RELU, PRO_AFFINE_RELU, MASK_POS, MASK_AFFINE, STATS_MOMENTS, STATS_BNBWD)

BN_EPS = 1e-5
BN_MOMENTUM = 0.1

RES_BLOCKS = ('encoder.4', 'encoder.5', 'decoder.0', 'decoder.1')
BN_LAYERS = (['encoder.1'] + [b + s for b in RES_BLOCKS for s in ('.block.2', '.block.5')] + ['decoder.4'])
BN_INDEX = {n: i for i, n in enumerate(BN_LAYERS)}


class Ctx(object):
    """Activations saved by ``forward`` for ``backward``."""
    pass


def _out_hw(h, w, k, s, p, transposed):
    if transposed:
        return (h - 1) * s - 2 * p + k, (w - 1) * s - 2 * p + k
    return (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1


class Engine(object):
    def __init__(self, input_dim, dim, z_dim, act_dtype=torch.float32, conv_impl=0, vq_mode=0, passes=None,
                 wgrad_impl=None):
        self.cin0, self.dim, self.k = input_dim, dim, z_dim
        self.act_dtype = act_dtype
        self.conv_impl = conv_impl      # 0 = fp32 CUDA-core kernels, 1 = tcgen05 kernels where shapes allow
        self.wgrad_impl = conv_impl if wgrad_impl is None else wgrad_impl
        self.vq_mode = vq_mode
        self.dt = ops.F32 if act_dtype == torch.float32 else ops.BF16
        # operand precision of the tcgen05 convolutions (include/nsg_b200.h): fp32 activations default to the
        # split-bf16 "fp32-accuracy" mode, bf16 activations to plain bf16 x bf16
        self.passes = passes if passes is not None else (3 if act_dtype == torch.float32 else 1)

    # ------------------------------------------------------------------ helpers
    def _impl(self, cin, cout):
        return 1 if (self.conv_impl == 1 and cin == 64 and cout == 64) else 0

    def _wimpl(self, cin, cout):
        return 1 if (self.wgrad_impl == 1 and cin == 64 and cout == 64) else 0

    def _new(self, *shape, dtype=None, like=None):
        return torch.empty(shape, dtype=dtype or self.act_dtype, device=like.device)

    def _bn(self, ctx, name):
        i = BN_INDEX[name]
        co = ctx.coef[i]
        return dict(mean=co[0], invstd=co[1], scale=co[2], shift=co[3], fstats=ctx.stats[i, 0], bstats=ctx.stats[i, 1])

    def _finalize(self, ctx, P, Bufs, name, count):
        bn = self._bn(ctx, name)
        if ctx.training:
            ops.bn_finalize(bn['fstats'], count, P[name + '.weight'], P[name + '.bias'], BN_EPS, BN_MOMENTUM,
                            Bufs[name + '.running_mean'], Bufs[name + '.running_var'],
                            Bufs[name + '.num_batches_tracked'], bn['mean'], bn['invstd'], bn['scale'], bn['shift'])
        else:
            ops.bn_eval_coeffs(P[name + '.weight'], P[name + '.bias'], Bufs[name + '.running_mean'],
                               Bufs[name + '.running_var'], BN_EPS, bn['scale'], bn['shift'])
        return bn

    def _conv_fwd(self, x, w, bias, k, s, p, transposed, pro, stats=None, tanh_out=False):
        """Conv2d (OIHW weight) or ConvTranspose2d ((Cin,Cout,KH,KW) weight) forward on NHWC x."""
        b, h, wd, cin = x.shape
        if transposed:
            cout = w.shape[1]
            wci, wco = cout * k * k, k * k
        else:
            cout = w.shape[0]
            wci, wco = k * k, cin * k * k
        oh, ow = _out_hw(h, wd, k, s, p, transposed)
        out = torch.empty((b, oh, ow, cout), dtype=x.dtype, device=x.device)
        g = ops.conv_geom(b, h, wd, cin, oh, ow, cout, k, s, p, 1 if transposed else 0, wci, wco, self._dt(x))
        ops.conv_gemm(g, x, w, out, pro=pro, bias=bias, stats_mode=STATS_MOMENTS if stats is not None else 0,
                      stats=stats, tanh_out=tanh_out, impl=self._impl(cin, cout), passes=self.passes)
        return out

    def _dt(self, t):
        return ops.F32 if t.dtype == torch.float32 else ops.BF16

    def _conv_dgrad(self, dy, w, k, s, p, transposed, in_hw, **epi):
        """Gradient wrt the input of a Conv2d / ConvTranspose2d forward whose input was (in_hw)."""
        b, oh, ow, cdy = dy.shape
        ih, iw = in_hw
        if transposed:      # forward was ConvTranspose2d(Cin->Cout): weight (Cin, Cout, k, k); dgrad gathers like a conv
            cin_f = w.shape[0]
            wci, wco = k * k, w.shape[1] * k * k          # GEMM in = co_f, GEMM out = ci_f
            gather = 0
        else:               # forward was Conv2d: weight (Cout, Cin, k, k); dgrad gathers like a transposed conv
            cin_f = w.shape[1]
            wci, wco = cin_f * k * k, k * k
            gather = 1
        out = torch.empty((b, ih, iw, cin_f), dtype=dy.dtype, device=dy.device)
        g = ops.conv_geom(b, oh, ow, cdy, ih, iw, cin_f, k, s, p, gather, wci, wco, self._dt(dy))
        ops.conv_gemm(g, dy, w, out, impl=self._impl(cdy, cin_f), passes=self.passes, **epi)
        return out

    def _conv_wgrad(self, x, dy, dw, k, s, p, transposed, pro_x, dbias=None):
        """dw += weight gradient; for Conv2d dbias (+= colsum dy) is fused, for ConvTranspose2d it is separate."""
        b, h, wd, cin = x.shape
        _, oh, ow, cout = dy.shape
        if transposed:      # dW[ci][co][t]: P = dy (gathered over x pixels), Q = x
            g = ops.conv_geom(b, oh, ow, cout, h, wd, cin, k, s, p, 0, k * k, cout * k * k, self._dt(x))
            ops.conv_wgrad(g, dy, x, dw, pro_p=None, pro_q=pro_x, q_colsum=None, impl=self._wimpl(cin, cout), passes=self.passes)
            if dbias is not None:
                ops.colsum(dy, dbias)
        else:               # dW[co][ci][t]: P = x (gathered over dy pixels), Q = dy
            g = ops.conv_geom(b, h, wd, cin, oh, ow, cout, k, s, p, 0, k * k, cin * k * k, self._dt(x))
            ops.conv_wgrad(g, x, dy, dw, pro_p=pro_x, pro_q=None, q_colsum=dbias, impl=self._wimpl(cin, cout), passes=self.passes)

    # ------------------------------------------------------------------ ResBlock (models.py:145-158)
    def resblock_fwd(self, ctx, P, Bufs, name, xin):
        """y = relu(xin) + BN5(Conv1x1(relu(BN2(Conv3x3(relu(xin))))))  (in-place leading ReLU: SURVEY D7)."""
        n = xin.numel() // xin.shape[-1]
        bn2, bn5 = self._bn(ctx, name + '.block.2'), self._bn(ctx, name + '.block.5')
        h1 = self._conv_fwd(xin, P[name + '.block.1.weight'], P[name + '.block.1.bias'], 3, 1, 1, False,
                            ops.prologue(PRO_RELU), stats=bn2['fstats'] if ctx.training else None)
        self._finalize(ctx, P, Bufs, name + '.block.2', n)
        h2 = self._conv_fwd(h1, P[name + '.block.4.weight'], P[name + '.block.4.bias'], 1, 1, 0, False,
                            ops.prologue(PRO_AFFINE_RELU, bn2['scale'], bn2['shift']),
                            stats=bn5['fstats'] if ctx.training else None)
        self._finalize(ctx, P, Bufs, name + '.block.5', n)
        y = torch.empty_like(xin)
        ops.bn_apply(h2, bn5['scale'], bn5['shift'], y, res=xin, res_relu=True)
        ctx.saved[name] = (xin, h1, h2)
        return y

    def resblock_bwd(self, ctx, P, G, name, g_y):
        xin, h1, h2 = ctx.saved[name]
        n = xin.numel() // xin.shape[-1]
        d = xin.shape[-1]
        bn2, bn5 = self._bn(ctx, name + '.block.2'), self._bn(ctx, name + '.block.5')
        # BN5 backward (no ReLU after it)
        ops.bn_bwd_reduce(g_y, h2, bn5['mean'], bn5['invstd'], bn5['bstats'])
        g_h2 = torch.empty_like(h2)
        ops.bn_bwd_apply(g_y, h2, bn5['mean'], bn5['invstd'], P[name + '.block.5.weight'], bn5['bstats'], n, g_h2,
                         dgamma=G[name + '.block.5.weight'], dbeta=G[name + '.block.5.bias'])
        # Conv1x1: input v = relu(BN2(h1))
        self._conv_wgrad(h1, g_h2, G[name + '.block.4.weight'], 1, 1, 0, False,
                         ops.prologue(PRO_AFFINE_RELU, bn2['scale'], bn2['shift']), dbias=G[name + '.block.4.bias'])
        g_v = self._conv_dgrad(g_h2, P[name + '.block.4.weight'], 1, 1, 0, False, xin.shape[1:3],
                               mask_mode=MASK_AFFINE, mask_src=h1, mscale=bn2['scale'], mshift=bn2['shift'],
                               stats_mode=STATS_BNBWD, stats=bn2['bstats'], xhat_src=h1, mean=bn2['mean'],
                               invstd=bn2['invstd'])
        g_h1 = torch.empty_like(h1)
        ops.bn_bwd_apply(g_v, h1, bn2['mean'], bn2['invstd'], P[name + '.block.2.weight'], bn2['bstats'], n, g_h1,
                         dgamma=G[name + '.block.2.weight'], dbeta=G[name + '.block.2.bias'])
        # Conv3x3: input r = relu(xin); total input gradient = (g_y + dgrad) masked by xin > 0
        self._conv_wgrad(xin, g_h1, G[name + '.block.1.weight'], 3, 1, 1, False, ops.prologue(PRO_RELU),
                         dbias=G[name + '.block.1.bias'])
        g_x = self._conv_dgrad(g_h1, P[name + '.block.1.weight'], 3, 1, 1, False, xin.shape[1:3],
                               add_src=g_y, mask_mode=MASK_POS, mask_src=xin)
        return g_x

    # ------------------------------------------------------------------ encoder / decoder (models.py:164-184)
    def begin(self, x_nhwc, training):
        ctx = Ctx()
        ctx.training = training
        ctx.saved = {}
        dev = x_nhwc.device
        ctx.coef = torch.empty((len(BN_LAYERS), 4, self.dim), dtype=torch.float32, device=dev)
        ctx.stats = torch.zeros((len(BN_LAYERS), 2, 2 * self.dim), dtype=torch.float64, device=dev)
        return ctx

    def encoder_fwd(self, ctx, P, Bufs, x):
        """x (B,H,W,Cin0) fp32 -> z_e (B,H/4,W/4,D)."""
        bn1 = self._bn(ctx, 'encoder.1')
        a0 = self._conv_fwd(x, P['encoder.0.weight'], P['encoder.0.bias'], 4, 2, 1, False, None,
                            stats=bn1['fstats'] if ctx.training else None)
        self._finalize(ctx, P, Bufs, 'encoder.1', a0.numel() // self.dim)
        a1 = self._conv_fwd(a0, P['encoder.3.weight'], P['encoder.3.bias'], 4, 2, 1, False,
                            ops.prologue(PRO_AFFINE_RELU, bn1['scale'], bn1['shift']))
        y = self.resblock_fwd(ctx, P, Bufs, 'encoder.4', a1)
        z_e = self.resblock_fwd(ctx, P, Bufs, 'encoder.5', y)
        ctx.x, ctx.a0 = x, a0
        return z_e

    def decoder_fwd(self, ctx, P, Bufs, z_q):
        """z_q (B,h,w,D) -> x_tilde (B,4h,4w,Cin0) fp32 (tanh applied)."""
        d0 = self.resblock_fwd(ctx, P, Bufs, 'decoder.0', z_q)
        d1 = self.resblock_fwd(ctx, P, Bufs, 'decoder.1', d0)
        bn4 = self._bn(ctx, 'decoder.4')
        u0 = self._conv_fwd(d1, P['decoder.3.weight'], P['decoder.3.bias'], 4, 2, 1, True, ops.prologue(PRO_RELU),
                            stats=bn4['fstats'] if ctx.training else None)
        self._finalize(ctx, P, Bufs, 'decoder.4', u0.numel() // self.dim)
        xt = self._conv_fwd(u0, P['decoder.6.weight'], P['decoder.6.bias'], 4, 2, 1, True,
                            ops.prologue(PRO_AFFINE_RELU, bn4['scale'], bn4['shift']), tanh_out=True)
        ctx.d1, ctx.u0 = d1, u0
        return xt

    def decoder_bwd(self, ctx, P, G, dpre):
        """dpre (B,H',W',Cin0): gradient wrt the pre-tanh output.  Returns grad wrt decoder input z_q."""
        u0, d1 = ctx.u0, ctx.d1
        bn4 = self._bn(ctx, 'decoder.4')
        n4 = u0.numel() // self.dim
        if dpre.dtype != u0.dtype:
            dpre = dpre.to(u0.dtype)
        # decoder.6: ConvTranspose2d(D -> Cin0), input relu(BN4(u0))
        self._conv_wgrad(u0, dpre, G['decoder.6.weight'], 4, 2, 1, True,
                         ops.prologue(PRO_AFFINE_RELU, bn4['scale'], bn4['shift']), dbias=G['decoder.6.bias'])
        g_v0 = self._conv_dgrad(dpre, P['decoder.6.weight'], 4, 2, 1, True, u0.shape[1:3],
                                mask_mode=MASK_AFFINE, mask_src=u0, mscale=bn4['scale'], mshift=bn4['shift'],
                                stats_mode=STATS_BNBWD, stats=bn4['bstats'], xhat_src=u0, mean=bn4['mean'],
                                invstd=bn4['invstd'])
        g_u0 = torch.empty_like(u0)
        ops.bn_bwd_apply(g_v0, u0, bn4['mean'], bn4['invstd'], P['decoder.4.weight'], bn4['bstats'], n4, g_u0,
                         dgamma=G['decoder.4.weight'], dbeta=G['decoder.4.bias'])
        # decoder.3: ConvTranspose2d(D -> D), input relu(d1)
        self._conv_wgrad(d1, g_u0, G['decoder.3.weight'], 4, 2, 1, True, ops.prologue(PRO_RELU),
                         dbias=G['decoder.3.bias'])
        g_d1 = self._conv_dgrad(g_u0, P['decoder.3.weight'], 4, 2, 1, True, d1.shape[1:3],
                                mask_mode=MASK_POS, mask_src=d1)
        g_d0 = self.resblock_bwd(ctx, P, G, 'decoder.1', g_d1)
        return self.resblock_bwd(ctx, P, G, 'decoder.0', g_d0)

    def encoder_bwd(self, ctx, P, G, g_ze, need_dx=False):
        x, a0 = ctx.x, ctx.a0
        bn1 = self._bn(ctx, 'encoder.1')
        g_y4 = self.resblock_bwd(ctx, P, G, 'encoder.5', g_ze)
        g_a1 = self.resblock_bwd(ctx, P, G, 'encoder.4', g_y4)
        # encoder.3: Conv2d(D -> D, 4, 2, 1), input relu(BN1(a0))
        self._conv_wgrad(a0, g_a1, G['encoder.3.weight'], 4, 2, 1, False,
                         ops.prologue(PRO_AFFINE_RELU, bn1['scale'], bn1['shift']), dbias=G['encoder.3.bias'])
        g_v = self._conv_dgrad(g_a1, P['encoder.3.weight'], 4, 2, 1, False, a0.shape[1:3],
                               mask_mode=MASK_AFFINE, mask_src=a0, mscale=bn1['scale'], mshift=bn1['shift'],
                               stats_mode=STATS_BNBWD, stats=bn1['bstats'], xhat_src=a0, mean=bn1['mean'],
                               invstd=bn1['invstd'])
        g_a0 = torch.empty_like(a0)
        ops.bn_bwd_apply(g_v, a0, bn1['mean'], bn1['invstd'], P['encoder.1.weight'], bn1['bstats'],
                         a0.numel() // self.dim, g_a0, dgamma=G['encoder.1.weight'], dbeta=G['encoder.1.bias'])
        # encoder.0: Conv2d(Cin0 -> D)
        self._conv_wgrad(x, g_a0, G['encoder.0.weight'], 4, 2, 1, False, None, dbias=G['encoder.0.bias'])
        if need_dx:
            return self._conv_dgrad(g_a0, P['encoder.0.weight'], 4, 2, 1, False, x.shape[1:3])
        return None

    def bottleneck_bwd(self, ctx, g_st, z_e, z_q, coef):
        """dz = g_st + coef*(z_e - z_q): straight-through clone + commitment term (vector_quantization.py:50-52)."""
        dz = torch.empty_like(z_e)
        ops.vq_backward_inputs(g_st, z_e, z_q, coef, dz)
        return dz

    def bottleneck_bwd_sum(self, ctx, g_st, g_ze):
        if g_ze is None:
            return g_st
        dz = torch.empty_like(g_st)
        ops.vq_backward_inputs(g_st, g_ze.to(g_st.dtype), None, 1.0, dz)
        return dz

    def finish_backward(self, ctx, G):
        pass

    # ------------------------------------------------------------------ bottleneck (models.py:127-142)
    def quantize(self, z_e, codebook, want_codes=True, sqerr=None):
        n = z_e.numel() // z_e.shape[-1]
        idx = torch.empty((n,), dtype=torch.int64, device=z_e.device)
        codes = torch.empty_like(z_e) if want_codes else None
        mode = self.vq_mode if ops.vq_mode_serves(codebook.shape[0], codebook.shape[1], self.vq_mode) else 0
        ops.vq_forward(z_e, codebook, idx, codes, sqerr, mode=mode)
        return idx, codes
