"""Op-level parity of the TMA-fed tcgen05 convolution kernels and their streaming companions (bf16, 64 channels),
called through the C ABI, against CPU fp64 restatements fed the SAME bf16-rounded operands (SURVEY.md H7 form (ii)).

Run on the B200 box: python -m pytest tests -m gpu
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import relerr, nhwc, nchw

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg  # noqa: F401
    from neural_sound_generation_b200 import ops

DEV = 'cuda'
BF = torch.bfloat16
OUT_TOL = 4e-3       # bf16 storage of the result: 2^-9 relative per element
ACC_TOL = 2e-4       # fp32-accumulated quantities (statistics, weight gradients)


def _rand(rng, *shape):
    return torch.tensor(rng.standard_normal(shape).astype(np.float32))


def _r16(t):
    return t.bfloat16().float()


def _geoms(B, H, W, k, s, p, transposed):
    """(fwd geom, dgrad geom, wgrad geom, OH, OW) for a 64 -> 64 layer, as engine.Engine builds them."""
    C = 64
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


TMA_CASES = [
    # (k, s, p, transposed, H, W, B)
    (3, 1, 1, False, 7, 9, 3),
    (3, 1, 1, False, 20, 64, 2),
    (1, 1, 0, False, 5, 6, 2),
    (1, 1, 0, False, 20, 64, 1),
    (4, 2, 1, False, 12, 10, 2),
    (4, 2, 1, False, 40, 128, 1),
    (4, 2, 1, False, 9, 7, 1),         # odd input extents: ragged parity views
    (4, 2, 1, True, 5, 7, 2),
    (4, 2, 1, True, 20, 64, 1),
    (3, 1, 1, False, 20, 215, 1),      # inference width (config 5)
]


@pytest.mark.parametrize('case', TMA_CASES)
def test_conv_tma_forward_dgrad_wgrad(case):
    k, s, p, transposed, H, W, B = case
    C = 64
    rng = np.random.RandomState(abs(hash(case)) % 2 ** 31)
    x = _r16(_rand(rng, B, C, H, W))
    w = _r16(_rand(rng, C, C, k, k) * 0.2)
    bias = _rand(rng, C)
    fwd, dgr, wgr, OH, OW = _geoms(B, H, W, k, s, p, transposed)
    xd = x.double().requires_grad_(True)
    wd = w.double().requires_grad_(True)
    yref = F.conv_transpose2d(xd, wd, bias.double(), s, p) if transposed else F.conv2d(xd, wd, bias.double(), s, p)
    assert tuple(yref.shape[2:]) == (OH, OW)
    xg = nhwc(x).to(DEV).to(BF)
    wg, bg = w.to(DEV), bias.to(DEV)
    # forward (+ BatchNorm moment statistics)
    y = torch.empty((B, OH, OW, C), dtype=BF, device=DEV)
    stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    ops.conv_gemm_tma(fwd, xg, ops.conv_pack_tma(fwd, wg), y, bias=bg, stats_mode=ops.STATS_MOMENTS, stats=stats)
    torch.cuda.synchronize()
    assert relerr(nchw(y.float()), yref) < OUT_TOL, 'fwd'
    sref = torch.cat([yref.sum((0, 2, 3)), (yref ** 2).sum((0, 2, 3))])
    assert relerr(stats, sref) < ACC_TOL, 'stats'
    # data and weight gradients
    gy = _r16(_rand(rng, *yref.shape))
    yref.backward(gy.double())
    gyg = nhwc(gy).to(DEV).to(BF)
    dx = torch.empty((B, H, W, C), dtype=BF, device=DEV)
    ops.conv_gemm_tma(dgr, gyg, ops.conv_pack_tma(dgr, wg), dx)
    torch.cuda.synchronize()
    assert relerr(nchw(dx.float()), xd.grad) < OUT_TOL, 'dgrad'
    dwp = torch.zeros((k * k, C, C), dtype=torch.float32, device=DEV)
    colsum = torch.zeros(C, dtype=torch.float32, device=DEV)
    if transposed:
        ops.conv_wgrad_tma(wgr, gyg, xg, dwp, q_colsum=colsum)       # P = dy, Q = x
        cref = x.double().sum((0, 2, 3))
    else:
        ops.conv_wgrad_tma(wgr, xg, gyg, dwp, q_colsum=colsum)       # P = x, Q = dy
        cref = gy.double().sum((0, 2, 3))
    dw = torch.zeros_like(wg)
    ops.wgrad_unpack_many([(dwp, dw, k * k, wgr.w_cin_stride, wgr.w_cout_stride)])
    torch.cuda.synchronize()
    assert relerr(dw, wd.grad) < ACC_TOL, 'wgrad'
    assert relerr(colsum, cref) < ACC_TOL, 'colsum'
    # accumulation semantics: a second call doubles the packed image
    if transposed:
        ops.conv_wgrad_tma(wgr, gyg, xg, dwp)
    else:
        ops.conv_wgrad_tma(wgr, xg, gyg, dwp)
    dw2 = torch.zeros_like(wg)
    ops.wgrad_unpack_many([(dwp, dw2, k * k, wgr.w_cin_stride, wgr.w_cout_stride)])
    torch.cuda.synchronize()
    assert relerr(dw2, 2 * wd.grad) < ACC_TOL, 'wgrad accumulate'
    # sliced form (the training step's): per-CTA partial images, overwritten, summed by the unpack; stale contents
    # of the slices must not matter and two runs must agree bit for bit (no atomics on the weight gradient)
    ns = ops.conv_wgrad_slices(wgr)
    assert ns >= 1
    outs = []
    for fill in (float('nan'), 7.0):
        sl = torch.full((ns, k * k, C, C), fill, dtype=torch.float32, device=DEV)
        if transposed:
            ops.conv_wgrad_tma(wgr, gyg, xg, sl, sliced=True)
        else:
            ops.conv_wgrad_tma(wgr, xg, gyg, sl, sliced=True)
        dw3 = torch.zeros_like(wg)
        ops.wgrad_unpack_many([(sl, dw3, k * k, wgr.w_cin_stride, wgr.w_cout_stride, ns)])
        torch.cuda.synchronize()
        assert relerr(dw3, wd.grad) < ACC_TOL, 'wgrad sliced'
        outs.append(dw3)
    assert torch.equal(outs[0], outs[1]), 'sliced wgrad is not run-to-run deterministic'
    with pytest.raises(ValueError):
        ops.conv_wgrad_tma(wgr, xg, gyg, torch.zeros(16, device=DEV), sliced=True)


def test_conv_tma_epilogue_mask_add_bnbwd_stats():
    rng = np.random.RandomState(7)
    B, C, H, W = 2, 64, 6, 5
    gy = _r16(_rand(rng, B, C, H, W))
    w = _r16(_rand(rng, C, C, 3, 3) * 0.1)
    add = _r16(_rand(rng, B, C, H, W))
    h = _r16(_rand(rng, B, C, H, W))
    scale, shift = 1 + 0.2 * _rand(rng, C), 0.3 * _rand(rng, C)
    mean, invstd = 0.1 * _rand(rng, C), 1 + 0.1 * _rand(rng, C).abs()
    xin = torch.zeros(B, C, H, W, dtype=torch.float64, requires_grad=True)
    F.conv2d(xin, w.double(), None, 1, 1).backward(gy.double())
    g = xin.grad + add.double()
    mask = (h.double() * scale.double()[None, :, None, None] + shift.double()[None, :, None, None]) > 0
    g = g * mask
    xhat = (h.double() - mean.double()[None, :, None, None]) * invstd.double()[None, :, None, None]
    sref = torch.cat([g.sum((0, 2, 3)), (g * xhat).sum((0, 2, 3))])
    _, dgr, _, _, _ = _geoms(B, H, W, 3, 1, 1, False)
    packed = ops.conv_pack_tma(dgr, w.to(DEV))
    stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    hg = nhwc(h).to(DEV).to(BF)
    addg = nhwc(add).to(DEV).to(BF)
    keep = (scale.to(DEV), shift.to(DEV), mean.to(DEV), invstd.to(DEV))
    out = torch.empty((B, H, W, C), dtype=BF, device=DEV)
    ops.conv_gemm_tma(dgr, nhwc(gy).to(DEV).to(BF), packed, out, add_src=addg, mask_mode=ops.MASK_AFFINE, mask_src=hg,
                      mscale=keep[0], mshift=keep[1], stats_mode=ops.STATS_BNBWD, stats=stats, xhat_src=hg,
                      mean=keep[2], invstd=keep[3])
    torch.cuda.synchronize()
    assert relerr(nchw(out.float()), g) < OUT_TOL
    assert relerr(stats, sref) < 5e-4
    out2 = torch.empty_like(out)
    ops.conv_gemm_tma(dgr, nhwc(gy).to(DEV).to(BF), packed, out2, mask_mode=ops.MASK_POS, mask_src=hg)
    torch.cuda.synchronize()
    assert relerr(nchw(out2.float()), xin.grad * (h.double() > 0)) < OUT_TOL


def test_conv_tma_rejects_other_shapes():
    g = ops.conv_geom(1, 4, 4, 32, 4, 4, 64, 3, 1, 1, 0, 9, 288, ops.BF16)
    x = torch.zeros((1, 4, 4, 32), dtype=BF, device=DEV)
    out = torch.zeros((1, 4, 4, 64), dtype=BF, device=DEV)
    with pytest.raises(nsg._cabi.NsgError):
        ops.conv_gemm_tma(g, x, torch.zeros(9 * 8192, dtype=torch.uint8, device=DEV), out)


@pytest.mark.parametrize('train', [True, False])
@pytest.mark.parametrize('res', [False, True])
def test_bn_act64(train, res):
    rng = np.random.RandomState(11)
    M, C = 1000, 64
    h = _r16(_rand(rng, M, C) * 1.5 + 0.3)
    r = _r16(_rand(rng, M, C).abs())
    gamma, beta = 1 + 0.1 * _rand(rng, C), 0.1 * _rand(rng, C)
    rm, rv = 0.1 * _rand(rng, C), 1 + 0.1 * _rand(rng, C).abs()
    hd = h.double()
    if train:
        mu, var = hd.mean(0), hd.var(0, unbiased=False)
    else:
        mu, var = rm.double(), rv.double()
    yref = (hd - mu) / torch.sqrt(var + 1e-5) * gamma.double() + beta.double()
    if res:
        yref = yref + r.double()
    yref = F.relu(yref)
    stats = torch.cat([hd.sum(0), (hd ** 2).sum(0)]).to(DEV) if train else None
    rmg, rvg = rm.to(DEV), rv.to(DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    coef = [torch.zeros(C, device=DEV) for _ in range(4)]
    out = torch.empty((M, C), dtype=BF, device=DEV)
    ops.bn_act64(h.to(DEV).to(BF), out, gamma.to(DEV), beta.to(DEV), stats=stats, count=M,
                 res=r.to(DEV).to(BF) if res else None, relu_out=True, running_mean=rmg, running_var=rvg, nbt=nbt,
                 coef=coef)
    torch.cuda.synchronize()
    assert relerr(out.float(), yref) < OUT_TOL
    assert relerr(coef[0], mu) < 1e-5 and relerr(coef[1], 1 / torch.sqrt(var + 1e-5)) < 1e-5
    if train:
        assert nbt.item() == 1
        assert relerr(rmg, 0.9 * rm.double() + 0.1 * mu) < 1e-5
        assert relerr(rvg, 0.9 * rv.double() + 0.1 * hd.var(0, unbiased=True)) < 1e-5
    else:
        assert nbt.item() == 0 and torch.equal(rmg.cpu(), rm)


def test_bn_bwd64_relu64_vq_bwd_stats64():
    rng = np.random.RandomState(12)
    M, C = 777, 64
    h = _r16(_rand(rng, M, C))
    g = _r16(_rand(rng, M, C))
    gamma = 1 + 0.1 * _rand(rng, C)
    hd = h.double().requires_grad_(True)
    mu, var = hd.mean(0), hd.var(0, unbiased=False)
    y = (hd - mu) / torch.sqrt(var + 1e-5) * gamma.double()
    y.backward(g.double())
    mu_d, is_d = mu.detach(), (1 / torch.sqrt(var + 1e-5)).detach()
    xhat = (h.double() - mu_d) * is_d
    stats = torch.cat([g.double().sum(0), (g.double() * xhat).sum(0)]).to(DEV)
    dh = torch.empty((M, C), dtype=BF, device=DEV)
    dgamma, dbeta, dbias = (torch.zeros(C, device=DEV) for _ in range(3))
    ops.bn_bwd64(g.to(DEV).to(BF), h.to(DEV).to(BF), dh, stats, M, mu_d.float().to(DEV), is_d.float().to(DEV),
                 gamma.to(DEV), dgamma, dbeta, dbias)
    torch.cuda.synchronize()
    assert relerr(dh.float(), hd.grad) < OUT_TOL
    assert relerr(dgamma, (g.double() * xhat).sum(0)) < 1e-5 and relerr(dbeta, g.double().sum(0)) < 1e-5
    assert dbias.abs().max().item() < 0.05 * g.abs().sum(0).max().item()      # analytically zero
    out = torch.empty((M, C), dtype=BF, device=DEV)
    ops.relu64(h.to(DEV).to(BF), out)
    assert torch.equal(out.float().cpu(), F.relu(h))
    z, zq = _r16(_rand(rng, M, C)), _r16(_rand(rng, M, C))
    dz = torch.empty((M, C), dtype=BF, device=DEV)
    st = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    ops.vq_bwd_stats64(g.to(DEV).to(BF), z.to(DEV).to(BF), zq.to(DEV).to(BF), 0.25, dz, h=h.to(DEV).to(BF),
                       mean=mu_d.float().to(DEV), invstd=is_d.float().to(DEV), stats=st)
    torch.cuda.synchronize()
    dref = g.double() + 0.25 * (z.double() - zq.double())
    assert relerr(dz.float(), dref) < OUT_TOL
    assert relerr(st, torch.cat([dref.sum(0), (dref * xhat).sum(0)])) < ACC_TOL


C1_CASES = [(2, 16, 12), (1, 80, 256), (3, 10, 34), (1, 8, 300)]      # (B, H, W) of the single-channel grid


@pytest.mark.parametrize('case', C1_CASES)
def test_c1_tensor_core_kernels(case):
    """encoder.0 / decoder.6 on the tensor cores vs fp64 on the same bf16-rounded operands."""
    B, H, W = case
    C, OH, OW = 64, H // 2, W // 2
    rng = np.random.RandomState(B * 1000 + W)
    x1 = _r16(_rand(rng, B, 1, H, W))
    w0 = _r16(_rand(rng, C, 1, 4, 4) * 0.3)                 # Conv2d(1 -> 64) weight; also ConvT(64 -> 1) weight shape
    bias = _rand(rng, C)
    # ---- Conv2d(1 -> 64, 4, 2, 1) forward + moment statistics
    xd = x1.double()
    wd = w0.double().requires_grad_(True)
    yref = F.conv2d(xd, wd, bias.double(), 2, 1)
    gf = ops.conv_geom(B, H, W, 1, OH, OW, C, 4, 2, 1, 0, 16, 16, ops.BF16)
    x1g = nhwc(x1).to(DEV).to(BF)
    y = torch.empty((B, OH, OW, C), dtype=BF, device=DEV)
    stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    w0g = w0.to(DEV)
    ops.c1_conv_tc(gf, x1g, ops.c1_pack(w0g, -1), y, bias=bias.to(DEV), stats_mode=ops.STATS_MOMENTS, stats=stats)
    torch.cuda.synchronize()
    assert relerr(nchw(y.float()), yref) < OUT_TOL, 'c1 conv fwd'
    assert relerr(stats, torch.cat([yref.sum((0, 2, 3)), (yref ** 2).sum((0, 2, 3))])) < ACC_TOL
    # ---- its weight gradient (+ column sums of dy)
    gy = _r16(_rand(rng, B, C, OH, OW))
    yref.backward(gy.double())
    gyg = nhwc(gy).to(DEV).to(BF)
    dw = torch.zeros_like(w0g)
    col = torch.zeros(C, device=DEV)
    ops.c1_wgrad_tc(gf, x1g, gyg, dw, q_colsum=col)
    torch.cuda.synchronize()
    assert relerr(dw, wd.grad) < ACC_TOL, 'c1 wgrad'
    assert relerr(col, gy.double().sum((0, 2, 3))) < ACC_TOL, 'c1 colsum'
    # ---- ConvTranspose2d(64 -> 1, 4, 2, 1) forward with tanh, on a (B, OH, OW, 64) input -> (B, H, W, 1)
    v = _r16(_rand(rng, B, C, OH, OW))
    b6 = _rand(rng, 1)
    tref = torch.tanh(F.conv_transpose2d(v.double(), w0.double(), b6.double(), 2, 1))
    gt = ops.conv_geom(B, OH, OW, C, 2 * OH, 2 * OW, 1, 4, 2, 1, 1, 16, 16, ops.BF16)
    xt = torch.empty((B, 2 * OH, 2 * OW, 1), dtype=BF, device=DEV)
    ops.c1_convT_tc(gt, nhwc(v).to(DEV).to(BF), ops.c1_pack(w0g, -2), xt, bias=b6.to(DEV), tanh_out=True)
    torch.cuda.synchronize()
    assert relerr(nchw(xt.float()), tref) < OUT_TOL, 'c1 convT fwd'
    # ---- its data gradient: conv-like gather of the 1-channel gradient with the same weight, masked, BN-bwd sums
    dpre = _r16(_rand(rng, B, 1, 2 * OH, 2 * OW))
    vin = v.double().requires_grad_(True)
    F.conv_transpose2d(vin, w0.double(), None, 2, 1).backward(dpre.double())
    h = _r16(_rand(rng, B, C, OH, OW))
    mean, invstd = 0.1 * _rand(rng, C), 1 + 0.1 * _rand(rng, C).abs()
    gref = vin.grad * (v.double() > 0)
    xhat = (h.double() - mean.double()[None, :, None, None]) * invstd.double()[None, :, None, None]
    gd = ops.conv_geom(B, 2 * OH, 2 * OW, 1, OH, OW, C, 4, 2, 1, 0, 16, 16, ops.BF16)
    g = torch.empty((B, OH, OW, C), dtype=BF, device=DEV)
    st = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    keep = (mean.to(DEV), invstd.to(DEV))
    ops.c1_conv_tc(gd, nhwc(dpre).to(DEV).to(BF), ops.c1_pack(w0g, -1), g, mask_mode=ops.MASK_POS,
                   mask_src=nhwc(v).to(DEV).to(BF), stats_mode=ops.STATS_BNBWD, stats=st,
                   xhat_src=nhwc(h).to(DEV).to(BF), mean=keep[0], invstd=keep[1])
    torch.cuda.synchronize()
    assert relerr(nchw(g.float()), gref) < OUT_TOL, 'c1 dgrad'
    assert relerr(st, torch.cat([gref.sum((0, 2, 3)), (gref * xhat).sum((0, 2, 3))])) < 5e-4


def test_programmatic_dependent_launch_keeps_results_and_orders_repacked_weights():
    """nsg_pdl (common.cuh): kernels may start their prologue under the predecessor's tail.  Results must be identical
    with the attribute on and off, including when a layer's weight image is re-packed by the kernel right before the
    convolution that consumes it (the one case where the early weight load has to be suppressed), and when the
    activations are produced by the kernel right before."""
    from neural_sound_generation_b200 import _cabi
    lib = _cabi.load()
    B, C, H, W, k = 4, 64, 20, 64, 3
    fwd, dgr, wgr, OH, OW = _geoms(B, H, W, k, 1, 1, False)
    g = torch.Generator(device='cpu').manual_seed(3)
    ws = [(torch.randn(C, C, k, k, generator=g) * 0.04).to(DEV) for _ in range(6)]
    x0 = torch.randn(B, H, W, C, generator=g).to(DEV).to(BF)
    img = torch.empty(k * k * 8192, dtype=torch.uint8, device=DEV)
    results = {}
    prev = lib.nsg_pdl(-1)
    try:
        for mode in (0, 1):
            lib.nsg_pdl(mode)
            outs = []
            a, b = x0.clone(), torch.empty_like(x0)
            for it in range(24):
                w = ws[it % len(ws)]
                # same image buffer re-packed every iteration, consumed by the very next launch
                ops.conv_pack_many([(w, img, k * k, fwd.w_cin_stride, fwd.w_cout_stride)])
                ops.conv_gemm_tma(fwd, a, img, b, relu_out=True)
                ops.conv_gemm_tma(fwd, b, img, a)                    # chained: input written by the previous launch
                if it % 8 == 7:
                    outs.append(a.clone())
            torch.cuda.synchronize()
            results[mode] = outs
    finally:
        lib.nsg_pdl(prev)
    for o0, o1 in zip(results[0], results[1]):
        assert torch.isfinite(o0.float()).all()
        assert torch.equal(o0, o1)


# ---------------------------------------------------------------------------------------------------------------------
# BatchNorm fused into the producing convolution across a grid barrier (nsg_conv_bn_tma): train forward, eval fold and
# backward, against fp64 autograd fed the same bf16-rounded operands.  Sizes cover one tile per CTA, several tiles per
# CTA of a full 148-CTA grid (the 4- and 8-accumulator instantiations) and a shape that does not fit.
# ---------------------------------------------------------------------------------------------------------------------
BN_CASES = [
    # (k, H, W, B, with residual, relu)
    (3, 7, 9, 3, False, True),
    (1, 5, 6, 2, True, False),
    (3, 20, 64, 32, False, True),      # configs[1] latent grid: 320 tiles on 148 CTAs (3 per CTA, 256 TMEM columns)
    (1, 20, 64, 32, True, True),
    (3, 20, 64, 70, True, True),       # 700 tiles: 5 per CTA (512 TMEM columns)
]


def _bf(t):
    return nhwc(t).to(DEV).to(BF)


@pytest.mark.parametrize('case', BN_CASES)
def test_conv_bn_tma_train_eval_backward(case):
    k, H, W, B, with_res, relu = case
    C, p = 64, k // 2
    rng = np.random.RandomState(abs(hash(case)) % 2 ** 31)
    x = _r16(_rand(rng, B, C, H, W))
    w = _r16(_rand(rng, C, C, k, k) * 0.2)
    bias, gamma, beta = _rand(rng, C), _rand(rng, C) * 0.5 + 1.0, _rand(rng, C) * 0.3
    res = _r16(_rand(rng, B, C, H, W)) if with_res else None
    rm0, rv0 = _rand(rng, C) * 0.1, torch.rand(C) + 0.5
    fwd, dgr, _, OH, OW = _geoms(B, H, W, k, 1, p, False)
    assert ops.conv_bn_tma_fits(fwd)
    n = B * OH * OW
    # ---- fp64 reference (torch semantics: biased variance normalises, unbiased updates the running statistics)
    xd, wd = x.double(), w.double()
    h_ref = F.conv2d(xd, wd, bias.double(), 1, p)
    mean, var = h_ref.mean((0, 2, 3)), h_ref.var((0, 2, 3), unbiased=False)
    invstd = 1.0 / torch.sqrt(var + 1e-5)
    y_ref = (h_ref - mean[None, :, None, None]) * (invstd * gamma.double())[None, :, None, None] + beta.double()[None, :, None, None]
    if with_res:
        y_ref = y_ref + res.double()
    v_ref = y_ref.clamp_min(0) if relu else y_ref
    # ---- fused train forward
    xg, wg = _bf(x), w.to(DEV)
    packed = ops.conv_pack_tma(fwd, wg)
    h = torch.empty((B, OH, OW, C), dtype=BF, device=DEV)
    v = torch.empty_like(h)
    stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    bars = torch.zeros(4, dtype=torch.int32, device=DEV)
    coef = torch.zeros((4, C), dtype=torch.float32, device=DEV)
    rm, rv = rm0.to(DEV), rv0.to(DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    ops.conv_bn_tma(fwd, xg, packed, h, ops.BNF_TRAIN, gamma.to(DEV), beta=beta.to(DEV), eps=1e-5, momentum=0.1, count=n,
                    running_mean=rm, running_var=rv, nbt=nbt, coef=coef, res=_bf(res) if with_res else None,
                    barrier=bars[0:1], out2=v, bn_relu=relu, bias=bias.to(DEV), stats=stats)
    torch.cuda.synchronize()
    assert relerr(nchw(h.float()), h_ref) < OUT_TOL, 'h'
    assert relerr(nchw(v.float()), v_ref) < OUT_TOL, 'v'
    assert relerr(coef[0].cpu(), mean) < ACC_TOL and relerr(coef[1].cpu(), invstd) < ACC_TOL
    assert relerr(rm.cpu(), 0.9 * rm0.double() + 0.1 * mean) < ACC_TOL
    assert relerr(rv.cpu(), 0.9 * rv0.double() + 0.1 * var * n / (n - 1)) < ACC_TOL
    assert nbt.item() == 1 and 0 < bars[0].item() <= 148        # every CTA of the grid arrived once
    # the fused result equals the two-launch form (convolution + bn_act) up to the bf16 rounding of h
    h2 = torch.empty_like(h)
    v2 = torch.empty_like(h)
    stats2 = torch.zeros_like(stats)
    ops.conv_gemm_tma(fwd, xg, packed, h2, bias=bias.to(DEV), stats_mode=ops.STATS_MOMENTS, stats=stats2)
    ops.bn_act64(h2, v2, gamma.to(DEV), beta.to(DEV), stats=stats2, count=n, res=_bf(res) if with_res else None,
                 relu_out=relu, eps=1e-5, momentum=0.1)
    torch.cuda.synchronize()
    assert torch.equal(h, h2)
    assert relerr(stats, stats2) < 1e-5
    assert relerr(v.float(), v2.float()) < OUT_TOL
    # ---- eval fold
    ve = torch.empty_like(h)
    ops.conv_bn_tma(fwd, xg, packed, ve, ops.BNF_EVAL, gamma.to(DEV), beta=beta.to(DEV), eps=1e-5,
                    running_mean=rm0.to(DEV), running_var=rv0.to(DEV), bias=bias.to(DEV),
                    add_src=_bf(res) if with_res else None, relu_out=relu)
    torch.cuda.synchronize()
    ye = (h_ref - rm0.double()[None, :, None, None]) * (gamma.double() / torch.sqrt(rv0.double() + 1e-5))[None, :, None, None] \
        + beta.double()[None, :, None, None]
    if with_res:
        ye = ye + res.double()
    assert relerr(nchw(ve.float()), ye.clamp_min(0) if relu else ye) < OUT_TOL, 'eval'
    # ---- backward: g = mask(dgrad(dy) [+ skip]) flows into the backward of a BatchNorm whose input was hb
    dy = _r16(_rand(rng, B, C, OH, OW))
    hb = _r16(_rand(rng, B, C, H, W) * 1.5 + 0.2)                   # pre-normalisation tensor of the BN being differentiated
    maskt = _r16(_rand(rng, B, C, H, W))
    skip = _r16(_rand(rng, B, C, H, W)) if with_res else None
    mu_b = hb.double().mean((0, 2, 3))
    is_b = 1.0 / torch.sqrt(hb.double().var((0, 2, 3), unbiased=False) + 1e-5)
    g_ref = F.conv_transpose2d(dy.double(), wd, None, 1, p)
    if with_res:
        g_ref = g_ref + skip.double()
    g_ref = g_ref * (maskt.double() > 0)
    xhat = (hb.double() - mu_b[None, :, None, None]) * is_b[None, :, None, None]
    nb = B * H * W
    sg, sgx = g_ref.sum((0, 2, 3)), (g_ref * xhat).sum((0, 2, 3))
    dh_ref = (gamma.double() * is_b)[None, :, None, None] * (g_ref - sg[None, :, None, None] / nb
                                                             - xhat * sgx[None, :, None, None] / nb)
    gout = torch.empty((B, H, W, C), dtype=BF, device=DEV)
    dh = torch.empty_like(gout)
    bst = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    dgamma = torch.full((C,), 0.5, dtype=torch.float32, device=DEV)
    dbeta = torch.full((C,), -0.25, dtype=torch.float32, device=DEV)
    ops.conv_bn_tma(dgr, _bf(dy), ops.conv_pack_tma(dgr, wg), gout, ops.BNF_BWD, gamma.to(DEV), count=nb,
                    barrier=bars[1:2], dgamma=dgamma, dbeta=dbeta, out2=dh, stats=bst, xhat_src=_bf(hb),
                    mean=mu_b.float().to(DEV), invstd=is_b.float().to(DEV), add_src=_bf(skip) if with_res else None,
                    mask_mode=ops.MASK_POS, mask_src=_bf(maskt))
    torch.cuda.synchronize()
    assert relerr(nchw(gout.float()), g_ref) < OUT_TOL, 'g'
    assert relerr(nchw(dh.float()), dh_ref) < 2 * OUT_TOL, 'dh'
    assert relerr(dgamma.cpu() - 0.5, sgx) < ACC_TOL * 5 and relerr(dbeta.cpu() + 0.25, sg) < ACC_TOL * 5
    # g not wanted: same dh
    dh2 = torch.empty_like(dh)
    bst.zero_()
    ops.conv_bn_tma(dgr, _bf(dy), ops.conv_pack_tma(dgr, wg), None, ops.BNF_BWD, gamma.to(DEV), count=nb,
                    barrier=bars[2:3], out2=dh2, stats=bst, xhat_src=_bf(hb), mean=mu_b.float().to(DEV),
                    invstd=is_b.float().to(DEV), add_src=_bf(skip) if with_res else None, mask_mode=ops.MASK_POS,
                    mask_src=_bf(maskt))
    torch.cuda.synchronize()
    assert relerr(dh2.float(), dh.float()) < 1e-3


def test_conv_bn_tma_more_tiles_than_accumulators():
    """The 40 x 128 grid at batch 32 is 1280 tiles = 9 per CTA, one more than the 8 TMEM accumulators: the first tile
    of each CTA is read back in pass 2 from the h / g that pass 1 stored.  Fused launches against the two-launch form
    (convolution + bn_act64 / bn_bwd64) on decoder.3's forward and encoder.3's data gradient."""
    rng = np.random.RandomState(11)
    B, C = 32, 64
    gamma, beta = (_rand(rng, C) * 0.5 + 1.0).to(DEV), (_rand(rng, C) * 0.3).to(DEV)
    bias = _rand(rng, C).to(DEV)
    # ---- decoder.3 forward: ConvTranspose2d(64, 64, 4, 2, 1) 20 x 64 -> 40 x 128, BatchNorm, ReLU
    fwd, _, _, OH, OW = _geoms(B, 20, 64, 4, 2, 1, True)
    assert ops.conv_bn_tma_fits(fwd)
    x = _rand(rng, B, 20, 64, C).to(DEV).to(BF)
    w = (_rand(rng, C, C, 4, 4) * 0.1).to(DEV)
    packed = ops.conv_pack_tma(fwd, w)
    n = B * OH * OW
    h, v = (torch.empty((B, OH, OW, C), dtype=BF, device=DEV) for _ in range(2))
    st = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    bars = torch.zeros(4, dtype=torch.int32, device=DEV)
    ops.conv_bn_tma(fwd, x, packed, h, ops.BNF_TRAIN, gamma, beta=beta, count=n, barrier=bars[0:1], out2=v, bn_relu=True,
                    bias=bias, stats=st)
    h2, v2 = torch.empty_like(h), torch.empty_like(h)
    st2 = torch.zeros_like(st)
    ops.conv_gemm_tma(fwd, x, packed, h2, bias=bias, stats_mode=ops.STATS_MOMENTS, stats=st2)
    ops.bn_act64(h2, v2, gamma, beta, stats=st2, count=n, relu_out=True)
    torch.cuda.synchronize()
    assert torch.equal(h, h2) and relerr(st, st2) < 1e-5
    assert relerr(v.float(), v2.float()) < OUT_TOL
    # ---- encoder.3 data gradient: Conv2d(64, 64, 4, 2, 1) 40 x 128 -> 20 x 64, gradient back on the 40 x 128 grid
    _, dgr, _, _, _ = _geoms(B, 40, 128, 4, 2, 1, False)
    dy = _rand(rng, B, 20, 64, C).to(DEV).to(BF)
    hb = (_rand(rng, B, 40, 128, C) * 1.5 + 0.2).to(DEV).to(BF)
    maskt = _rand(rng, B, 40, 128, C).to(DEV).to(BF)
    mu = hb.float().mean((0, 1, 2))
    istd = 1.0 / torch.sqrt(hb.float().var((0, 1, 2), unbiased=False) + 1e-5)
    pk = ops.conv_pack_tma(dgr, w)
    nb = B * 40 * 128
    g, dh = torch.empty_like(hb), torch.empty_like(hb)
    bst = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    dgam, dbet = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.conv_bn_tma(dgr, dy, pk, g, ops.BNF_BWD, gamma, count=nb, barrier=bars[1:2], dgamma=dgam, dbeta=dbet, out2=dh,
                    stats=bst, xhat_src=hb, mean=mu, invstd=istd, mask_mode=ops.MASK_POS, mask_src=maskt)
    g2, dh2 = torch.empty_like(hb), torch.empty_like(hb)
    bst2 = torch.zeros_like(bst)
    dgam2, dbet2 = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.conv_gemm_tma(dgr, dy, pk, g2, mask_mode=ops.MASK_POS, mask_src=maskt, stats_mode=ops.STATS_BNBWD, stats=bst2,
                      xhat_src=hb, mean=mu, invstd=istd)
    ops.bn_bwd64(g2, hb, dh2, bst2, nb, mu, istd, gamma, dgamma=dgam2, dbeta=dbet2)
    torch.cuda.synchronize()
    assert torch.equal(g, g2) and relerr(bst, bst2) < 1e-4
    assert relerr(dh.float(), dh2.float()) < OUT_TOL
    assert relerr(dgam, dgam2) < 1e-4 and relerr(dbet, dbet2) < 1e-4
    with pytest.raises(Exception, match='first output'):
        ops.conv_bn_tma(fwd, x, packed, None, ops.BNF_TRAIN, gamma, beta=beta, count=n, barrier=bars[2:3], out2=v,
                        bias=bias, stats=st)
