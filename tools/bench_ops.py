"""Per-kernel timings of the bf16 TMA path at the config-2 layer sizes (CUDA events, L2 flushed between launches).

    python tools/bench_ops.py            # prints one line per op: us, TFLOP/s or GB/s
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import neural_sound_generation_b200 as nsg  # noqa: E402,F401
from neural_sound_generation_b200 import ops  # noqa: E402

DEV = 'cuda'
BF = torch.bfloat16
C = 64


def timed(fn, iters=int(os.environ.get('BENCH_ITERS', '20')), flush=None):
    for _ in range(3):
        fn()
    if flush is None:       # warm: 20 launches captured in one CUDA graph (no host launch overhead in the figure)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(iters):
                fn()
        g.replay()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) * 1e3 / iters
    evs = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        evs.append((s, e))
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) * 1e3 for s, e in evs)
    return ts[len(ts) // 2]


def geoms(B, H, W, k, s, p, transposed):
    if transposed:
        OH, OW = (H - 1) * s - 2 * p + k, (W - 1) * s - 2 * p + k
        fwd = ops.conv_geom(B, H, W, C, OH, OW, C, k, s, p, 1, C * k * k, k * k, ops.BF16)
        dgr = ops.conv_geom(B, OH, OW, C, H, W, C, k, s, p, 0, k * k, C * k * k, ops.BF16)
        wgr = ops.conv_geom(B, OH, OW, C, H, W, C, k, s, p, 0, k * k, C * k * k, ops.BF16)
    else:
        OH, OW = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        fwd = ops.conv_geom(B, H, W, C, OH, OW, C, k, s, p, 0, k * k, C * k * k, ops.BF16)
        dgr = ops.conv_geom(B, OH, OW, C, H, W, C, k, s, p, 1, C * k * k, k * k, ops.BF16)
        wgr = ops.conv_geom(B, H, W, C, OH, OW, C, k, s, p, 0, k * k, C * k * k, ops.BF16)
    return fwd, dgr, wgr, OH, OW


def main():
    torch.manual_seed(0)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    B = 32
    layers = [('conv3x3 s1 (ResBlock)', 20, 64, 3, 1, 1, False), ('conv1x1 (ResBlock)', 20, 64, 1, 1, 0, False),
              ('conv4x4 s2 (encoder.3)', 40, 128, 4, 2, 1, False), ('convT4x4 s2 (decoder.3)', 20, 64, 4, 2, 1, True)]
    only = os.environ.get('BENCH_LAYER')
    if only is not None:
        layers = [layers[int(only)]] if int(only) >= 0 else []
    for name, H, W, k, s, p, tr in layers:
        fwd, dgr, wgr, OH, OW = geoms(B, H, W, k, s, p, tr)
        x = torch.randn(B, H, W, C, device=DEV).to(BF)
        gy = torch.randn(B, OH, OW, C, device=DEV).to(BF)
        w = torch.randn(C, C, k, k, device=DEV) * 0.1
        bias = torch.zeros(C, device=DEV)
        y = torch.empty(B, OH, OW, C, dtype=BF, device=DEV)
        dx = torch.empty_like(x)
        stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
        pf, pd = ops.conv_pack_tma(fwd, w), ops.conv_pack_tma(dgr, w)
        dwp = torch.zeros(k * k, C, C, device=DEV)
        col = torch.zeros(C, device=DEV)
        macs = B * (OH * OW if not tr else H * W) * C * C * k * k
        flop = 2.0 * macs
        for fl, tag in (((flush, 'cold'),) if only is not None else ((flush, 'cold'), (None, 'warm'))):
            t = timed(lambda: ops.conv_gemm_tma(fwd, x, pf, y, bias=bias, stats_mode=ops.STATS_MOMENTS, stats=stats), flush=fl)
            print('%-26s fwd+stats  %s %8.1f us  %7.1f TFLOP/s' % (name, tag, t, flop / t / 1e6))
            t = timed(lambda: ops.conv_gemm_tma(dgr, gy, pd, dx, mask_mode=ops.MASK_POS, mask_src=x), flush=fl)
            print('%-26s dgrad+mask %s %8.1f us  %7.1f TFLOP/s' % (name, tag, t, flop / t / 1e6))
            if tr:
                t = timed(lambda: ops.conv_wgrad_tma(wgr, gy, x, dwp, q_colsum=col), flush=fl)
            else:
                t = timed(lambda: ops.conv_wgrad_tma(wgr, x, gy, dwp, q_colsum=col), flush=fl)
            print('%-26s wgrad      %s %8.1f us  %7.1f TFLOP/s' % (name, tag, t, flop / t / 1e6))
            ns = ops.conv_wgrad_slices(wgr)
            sl = torch.empty(ns, k * k, C, C, device=DEV)
            if tr:
                t = timed(lambda: ops.conv_wgrad_tma(wgr, gy, x, sl, q_colsum=col, sliced=True), flush=fl)
            else:
                t = timed(lambda: ops.conv_wgrad_tma(wgr, x, gy, sl, q_colsum=col, sliced=True), flush=fl)
            print('%-26s wgrad/%-3d  %s %8.1f us  %7.1f TFLOP/s' % (name, ns, tag, t, flop / t / 1e6))
    if only is not None and int(only) >= 0:
        return
    # single-channel ends (conv_c1_fast.cu): encoder.0 fwd / wgrad, decoder.6 fwd / dgrad / wgrad
    x1 = torch.rand(B, 80, 256, 1, device=DEV).to(BF)
    a0 = torch.empty(B, 40, 128, C, dtype=BF, device=DEV)
    v4 = torch.randn(B, 40, 128, C, device=DEV).to(BF)
    g64 = torch.randn(B, 40, 128, C, device=DEV).to(BF)
    w0 = torch.randn(C, 1, 4, 4, device=DEV) * 0.1
    w6 = torch.randn(C, 1, 4, 4, device=DEV) * 0.1
    b0, b6 = torch.zeros(C, device=DEV), torch.zeros(1, device=DEV)
    xt = torch.empty(B, 80, 256, 1, dtype=BF, device=DEV)
    st = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    mean, invstd = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    dw = torch.zeros(C, 1, 4, 4, device=DEV)
    gf = ops.conv_geom(B, 80, 256, 1, 40, 128, C, 4, 2, 1, 0, 16, 16, ops.BF16)
    gt = ops.conv_geom(B, 40, 128, C, 80, 256, 1, 4, 2, 1, 1, 16, 16, ops.BF16)
    nb = a0.numel() * 2
    for fl, tag in (((flush, 'cold'),) if only is not None else ((flush, 'cold'), (None, 'warm'))):
        t = timed(lambda: ops.conv_gemm(gf, x1, w0, a0, bias=b0, stats_mode=ops.STATS_MOMENTS, stats=st), flush=fl)
        print('c1 conv 1->64 fwd+stats     %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
        t = timed(lambda: ops.conv_gemm(gf, x1, w6, a0, mask_mode=ops.MASK_POS, mask_src=v4, stats_mode=ops.STATS_BNBWD,
                                        stats=st, xhat_src=g64, mean=mean, invstd=invstd), flush=fl)
        print('c1 conv 1->64 dgrad+mask+bn %s %8.1f us  %7.1f GB/s' % (tag, t, 3 * nb / t / 1e3))
        t = timed(lambda: ops.conv_gemm(gt, v4, w6, xt, bias=b6, tanh_out=True), flush=fl)
        print('c1 convT 64->1 fwd+tanh     %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
        t = timed(lambda: ops.conv_wgrad(gf, x1, g64, dw), flush=fl)
        print('c1 wgrad                    %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
        pk1, pk2 = ops.c1_pack(w0, -1), ops.c1_pack(w6, -2)
        t = timed(lambda: ops.c1_conv_tc(gf, x1, pk1, a0, bias=b0, stats_mode=ops.STATS_MOMENTS, stats=st), flush=fl)
        print('c1t conv 1->64 fwd+stats    %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
        t = timed(lambda: ops.c1_conv_tc(gf, x1, pk1, a0, mask_mode=ops.MASK_POS, mask_src=v4, stats_mode=ops.STATS_BNBWD,
                                         stats=st, xhat_src=g64, mean=mean, invstd=invstd), flush=fl)
        print('c1t conv 1->64 dgrad+mask+bn %s %8.1f us  %7.1f GB/s' % (tag, t, 3 * nb / t / 1e3))
        t = timed(lambda: ops.c1_convT_tc(gt, v4, pk2, xt, bias=b6, tanh_out=True), flush=fl)
        print('c1t convT 64->1 fwd+tanh    %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
        t = timed(lambda: ops.c1_wgrad_tc(gf, x1, g64, dw), flush=fl)
        print('c1t wgrad                   %s %8.1f us  %7.1f GB/s' % (tag, t, nb / t / 1e3))
    if only is not None:
        return
    for name, H, W in (('20x64', 20, 64), ('40x128', 40, 128)):
        h = torch.randn(B, H, W, C, device=DEV).to(BF)
        g = torch.randn(B, H, W, C, device=DEV).to(BF)
        out = torch.empty_like(h)
        gamma, beta = torch.ones(C, device=DEV), torch.zeros(C, device=DEV)
        rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
        coef = [torch.zeros(C, device=DEV) for _ in range(4)]
        st = torch.rand(2 * C, dtype=torch.float64, device=DEV) * h.numel() / C
        st[C:] += st[:C] ** 2 / (h.numel() / C)
        nb = h.numel() * 2
        for fl, tag in ((flush, 'cold'), (None, 'warm')):
            t = timed(lambda: ops.bn_act64(h, out, gamma, beta, stats=st, count=h.numel() // C, running_mean=rm,
                                           running_var=rv, coef=coef), flush=fl)
            print('bn_act64 %-8s %s %8.1f us  %7.1f GB/s' % (name, tag, t, 2 * nb / t / 1e3))
            t = timed(lambda: ops.bn_bwd64(g, h, out, st, h.numel() // C, coef[0], coef[1], gamma, coef[2], coef[3], rm),
                      flush=fl)
            print('bn_bwd64 %-8s %s %8.1f us  %7.1f GB/s' % (name, tag, t, 3 * nb / t / 1e3))


if __name__ == '__main__':
    main()
