"""Op-level parity of the C-ABI kernels (called through ctypes) against CPU fp64 restatements.

Run on the B200 box: python -m pytest tests -m gpu
"""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import relerr, nhwc, nchw

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import ops
    from oracle import vq_ref


def _skip_if_tc_missing(fn):
    """mode=1 (tcgen05) paths raise NSG_E_UNSUPPORTED for shapes/builds they do not cover."""
    try:
        return fn()
    except nsg._cabi.NsgError as e:
        if '(-3)' in str(e):
            pytest.skip(str(e))
        raise

DEV = 'cuda'
TOL = {torch.float32: 2e-5, torch.bfloat16: 1.5e-2}


def _rand(rng, *shape):
    return torch.tensor(rng.standard_normal(shape).astype(np.float32))


def _pro_ref(x, mode, scale, shift):
    if mode == ops.PRO_RELU:
        return F.relu(x)
    if mode == ops.PRO_AFFINE_RELU:
        return F.relu(x * scale[None, :, None, None] + shift[None, :, None, None])
    if mode == ops.PRO_AFFINE:
        return x * scale[None, :, None, None] + shift[None, :, None, None]
    return x


CONV_CASES = [
    # (cin, cout, k, s, p, transposed, H, W, B, pro)
    (1, 64, 4, 2, 1, False, 16, 12, 2, 0),
    (64, 64, 4, 2, 1, False, 12, 10, 2, 2),
    (64, 64, 3, 1, 1, False, 7, 9, 3, 1),
    (64, 64, 1, 1, 0, False, 5, 6, 2, 2),
    (64, 64, 4, 2, 1, True, 5, 7, 2, 1),
    (64, 1, 4, 2, 1, True, 6, 5, 2, 2),
    (3, 20, 4, 2, 1, False, 9, 11, 1, 0),      # generic (non-multiple-of-4) channel counts
    (20, 3, 4, 2, 1, True, 4, 5, 1, 3),
    (32, 48, 3, 1, 1, False, 6, 6, 1, 0),
]


TC_TOL = {torch.float32: 5e-5, torch.bfloat16: 1.5e-2}     # tcgen05: split-bf16 (3 passes) / plain bf16 operands


@pytest.mark.parametrize('case', CONV_CASES)
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('impl', [0, 1])
def test_conv_forward_dgrad_wgrad(case, dtype, impl):
    cin, cout, k, s, p, transposed, H, W, B, pro = case
    if impl == 1 and not (cin == 64 and cout == 64):
        pytest.skip('tcgen05 convolutions serve the 64 -> 64 channel layers')
    TOL = TC_TOL if impl == 1 else globals()['TOL']
    rng = np.random.RandomState(hash(case) % 2 ** 31)
    x = _rand(rng, B, cin, H, W)
    w = _rand(rng, *((cin, cout, k, k) if transposed else (cout, cin, k, k))) * 0.2
    bias = _rand(rng, cout)
    scale = 1.0 + 0.2 * _rand(rng, cin)
    shift = 0.3 * _rand(rng, cin)
    if dtype == torch.bfloat16:      # both sides see the same (rounded) activations
        x = x.bfloat16().float()
    xd = x.double().requires_grad_(True)
    wd = w.double().requires_grad_(True)
    xin = _pro_ref(xd, pro, scale.double(), shift.double())
    if transposed:
        yref = F.conv_transpose2d(xin, wd, bias.double(), s, p)
    else:
        yref = F.conv2d(xin, wd, bias.double(), s, p)
    OH, OW = yref.shape[2:]
    eng = nsg.engine.Engine(cin, 64, 8, act_dtype=dtype, conv_impl=impl)
    xg = nhwc(x).to(DEV).to(dtype)
    wg, bg = w.to(DEV), bias.to(DEV)
    scale_g, shift_g = scale.to(DEV), shift.to(DEV)      # keep alive: the prologue struct holds raw pointers
    prol = ops.prologue(pro, scale_g, shift_g) if pro else None
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    y = eng._conv_fwd(xg, wg, bg, k, s, p, transposed, prol, stats=stats)
    torch.cuda.synchronize()
    assert tuple(y.shape) == (B, OH, OW, cout)
    err = relerr(nchw(y.float()), yref)
    assert err < TOL[dtype], ('fwd', err)
    # statistics epilogue: sum and sum of squares per output channel (computed before rounding to storage)
    sref = torch.cat([yref.sum((0, 2, 3)), (yref ** 2).sum((0, 2, 3))])
    assert relerr(stats, sref) < (1e-4 if dtype == torch.float32 else 2e-2)
    # dgrad + wgrad against autograd of the fp64 restatement
    gy = _rand(rng, *yref.shape)
    if dtype == torch.bfloat16:
        gy = gy.bfloat16().float()
    yref.backward(gy.double())
    # gradient wrt the (post-prologue) conv input
    xin2 = _pro_ref(x.double(), pro, scale.double(), shift.double()).detach().requires_grad_(True)
    y2 = (F.conv_transpose2d(xin2, w.double(), None, s, p) if transposed else F.conv2d(xin2, w.double(), None, s, p))
    y2.backward(gy.double())
    gyg = nhwc(gy).to(DEV).to(dtype)
    dx = eng._conv_dgrad(gyg, wg, k, s, p, transposed, (H, W))
    torch.cuda.synchronize()
    assert relerr(nchw(dx.float()), xin2.grad) < TOL[dtype], 'dgrad'
    dw = torch.zeros_like(wg)
    db = torch.zeros(cout, device=DEV)
    eng._conv_wgrad(xg, gyg, dw, k, s, p, transposed, prol, dbias=db)
    torch.cuda.synchronize()
    assert relerr(dw, wd.grad) < TOL[dtype], 'wgrad'
    assert relerr(db, gy.double().sum((0, 2, 3))) < TOL[dtype], 'dbias'


@pytest.mark.parametrize('impl', [0, 1])
def test_conv_epilogue_mask_add_bnbwd_stats(impl):
    rng = np.random.RandomState(7)
    B, C, H, W = 2, 64, 6, 5
    gy = _rand(rng, B, C, H, W)
    w = _rand(rng, C, C, 3, 3) * 0.1
    add = _rand(rng, B, C, H, W)
    h = _rand(rng, B, C, H, W)
    scale, shift = 1 + 0.2 * _rand(rng, C), 0.3 * _rand(rng, C)
    mean, invstd = 0.1 * _rand(rng, C), 1 + 0.1 * _rand(rng, C).abs()
    # reference: dgrad of conv3x3, + add, masked by affine(h) > 0, sums of g and g*xhat
    xin = torch.zeros(B, C, H, W, dtype=torch.float64, requires_grad=True)
    F.conv2d(xin, w.double(), None, 1, 1).backward(gy.double())
    g = xin.grad + add.double()
    mask = (h.double() * scale.double()[None, :, None, None] + shift.double()[None, :, None, None]) > 0
    g = g * mask
    xhat = (h.double() - mean.double()[None, :, None, None]) * invstd.double()[None, :, None, None]
    sref = torch.cat([g.sum((0, 2, 3)), (g * xhat).sum((0, 2, 3))])
    eng = nsg.engine.Engine(1, C, 8, conv_impl=impl)
    stats = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    hg = nhwc(h).to(DEV)
    out = eng._conv_dgrad(nhwc(gy).to(DEV), w.to(DEV), 3, 1, 1, False, (H, W), add_src=nhwc(add).to(DEV),
                          mask_mode=ops.MASK_AFFINE, mask_src=hg, mscale=scale.to(DEV), mshift=shift.to(DEV),
                          stats_mode=ops.STATS_BNBWD, stats=stats, xhat_src=hg, mean=mean.to(DEV),
                          invstd=invstd.to(DEV))
    torch.cuda.synchronize()
    assert relerr(nchw(out), g) < (5e-5 if impl else 2e-5)
    assert relerr(stats, sref) < 1e-4
    # MASK_POS
    out2 = eng._conv_dgrad(nhwc(gy).to(DEV), w.to(DEV), 3, 1, 1, False, (H, W), mask_mode=ops.MASK_POS, mask_src=hg)
    assert relerr(nchw(out2), xin.grad * (h.double() > 0)) < (5e-5 if impl else 2e-5)


def test_tanh_epilogue_and_recon_loss():
    rng = np.random.RandomState(8)
    B, C, H, W = 2, 64, 5, 4
    x = _rand(rng, B, C, H, W)
    w = _rand(rng, C, 1, 4, 4) * 0.1
    b = _rand(rng, 1)
    eng = nsg.engine.Engine(1, C, 8)
    y = eng._conv_fwd(nhwc(x).to(DEV), w.to(DEV), b.to(DEV), 4, 2, 1, True, None, tanh_out=True)
    ref = torch.tanh(F.conv_transpose2d(x.double(), w.double(), b.double(), 2, 1))
    assert relerr(nchw(y), ref) < 2e-5
    # reconstruction loss with right zero padding (train.py:118-129) and its gradient through tanh
    Wt, Wc = 2 * W, 2 * W + 3
    xt = ref[:, 0].float().contiguous()                       # (B, 2H, 2W)
    c = torch.tensor(rng.uniform(0, 1, (B, 2 * H, Wc)).astype(np.float32))
    xt64 = xt.double().requires_grad_(True)
    pre = torch.atanh(xt64.detach().clamp(-0.999999, 0.999999)).requires_grad_(True)
    loss = F.mse_loss(F.pad(torch.tanh(pre), (0, Wc - Wt)), c.double())
    loss.backward()
    sums = torch.zeros(1, dtype=torch.float64, device=DEV)
    dpre = torch.empty(B, 2 * H, Wt, device=DEV)
    ops.recon_loss(torch.tanh(pre).detach().float().to(DEV), c.to(DEV), sums, dpre)
    torch.cuda.synchronize()
    assert abs(sums.item() / c.numel() - loss.item()) < 1e-6 * max(1.0, loss.item())
    assert relerr(dpre, pre.grad) < 1e-5
    g = _rand(rng, B, 2 * H, Wt)
    out = torch.empty_like(dpre)
    ops.tanh_bwd(g.to(DEV), xt.to(DEV), out)
    assert relerr(out, g.double() * (1 - xt.double() ** 2)) < 1e-6


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('C', [64, 20, 1])
def test_batchnorm_kernels(dtype, C):
    rng = np.random.RandomState(9 + C)
    B, H, W = 3, 7, 5
    h = _rand(rng, B, C, H, W) * 1.5 + 0.3
    res = _rand(rng, B, C, H, W)
    g = _rand(rng, B, C, H, W)
    if dtype == torch.bfloat16:
        h, res, g = h.bfloat16().float(), res.bfloat16().float(), g.bfloat16().float()
    gamma, beta = 1 + 0.2 * _rand(rng, C), 0.1 * _rand(rng, C)
    rm, rv = 0.1 * _rand(rng, C), 1 + 0.1 * _rand(rng, C).abs()
    n = B * H * W
    hd = h.double().requires_grad_(True)
    gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    rm_ref, rv_ref = rm.double().clone(), rv.double().clone()
    yref = F.batch_norm(hd, rm_ref, rv_ref, gd, bd, True, 0.1, 1e-5)
    out_ref = F.relu(res.double()) + yref
    out_ref.backward(g.double())
    # forward: stats -> finalize -> apply
    hg = nhwc(h).to(DEV).to(dtype)
    stats = torch.cat([h.double().sum((0, 2, 3)), (h.double() ** 2).sum((0, 2, 3))]).to(DEV)
    coef = torch.empty(4, C, device=DEV)
    rmg, rvg = rm.to(DEV), rv.to(DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    ops.bn_finalize(stats, n, gamma.to(DEV), beta.to(DEV), 1e-5, 0.1, rmg, rvg, nbt, coef[0], coef[1], coef[2], coef[3])
    y = torch.empty_like(hg)
    ops.bn_apply(hg, coef[2], coef[3], y, res=nhwc(res).to(DEV).to(dtype), res_relu=True)
    torch.cuda.synchronize()
    assert int(nbt) == 1
    assert relerr(rmg, rm_ref) < 1e-5 and relerr(rvg, rv_ref) < 1e-5
    assert relerr(nchw(y.float()), out_ref) < TOL[dtype]
    # backward: reduce -> apply
    bst = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    gg = nhwc(g).to(DEV).to(dtype)
    ops.bn_bwd_reduce(gg, hg, coef[0], coef[1], bst)
    dh = torch.empty_like(hg)
    dgam, dbet = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    ops.bn_bwd_apply(gg, hg, coef[0], coef[1], gamma.to(DEV), bst, n, dh, dgamma=dgam, dbeta=dbet)
    torch.cuda.synchronize()
    assert relerr(nchw(dh.float()), hd.grad) < (2e-4 if dtype == torch.float32 else 3e-2)
    assert relerr(dgam, gd.grad) < 1e-4 and relerr(dbet, bd.grad) < 1e-4
    # eval coefficients, relu mask variant, colsum, relu_bwd_add
    sc, sh = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    ops.bn_eval_coeffs(gamma.to(DEV), beta.to(DEV), rm.to(DEV), rv.to(DEV), 1e-5, sc, sh)
    isd = 1 / torch.sqrt(rv.double() + 1e-5)
    assert relerr(sc, gamma.double() * isd) < 1e-6 and relerr(sh, beta.double() - rm.double() * gamma.double() * isd) < 1e-5
    cs = torch.zeros(C, device=DEV)
    ops.colsum(gg, cs)
    assert relerr(cs, g.double().sum((0, 2, 3))) < 1e-5
    dx = torch.empty_like(hg)
    ops.relu_bwd_add(gg, hg, nhwc(res).to(DEV).to(dtype), dx)
    assert relerr(nchw(dx.float()), (g.double() + h.double()) * (res.double() > 0)) < TOL[dtype]
    bst2 = torch.zeros(2 * C, dtype=torch.float64, device=DEV)
    ops.bn_bwd_reduce(gg, hg, coef[0], coef[1], bst2, scale=coef[2], shift=coef[3], relu_mask=True)
    m = (yref.detach() > 0).double()
    xhat = (h.double() - h.double().mean((0, 2, 3), keepdim=True)) / torch.sqrt(
        h.double().var((0, 2, 3), unbiased=False, keepdim=True) + 1e-5)
    ref2 = torch.cat([(g.double() * m).sum((0, 2, 3)), (g.double() * m * xhat).sum((0, 2, 3))])
    assert relerr(bst2, ref2) < (1e-4 if dtype == torch.float32 else 5e-2)


# ------------------------------------------------------------------------------------ vector quantisation
def _vq_case(tag, golden_dir):
    g = np.load(os.path.join(golden_dir, 'vq_c1.npz'))
    rng = np.random.RandomState(int(g[tag + '/seed']))
    z = rng.standard_normal((16, 128, 64)).astype(np.float32)
    cb = (rng.uniform(-1 / 512, 1 / 512, (512, 64)) if tag == 'init' else rng.standard_normal((512, 64))).astype(np.float32)
    gout = rng.standard_normal((16, 128, 64)).astype(np.float32)
    return g, z, cb, gout


@pytest.mark.parametrize('tag', ['init', 'spread'])
@pytest.mark.parametrize('mode', [0, 1])
def test_vq_c1_golden(golden_dir, tag, mode):
    """Config C1 against the reference's own outputs (tests/golden/vq_c1.npz)."""
    g, z, cb, gout = _vq_case(tag, golden_dir)
    nsg.vector_quantization.DEFAULT_MODE = mode
    try:
        zt = torch.tensor(z, device=DEV, requires_grad=True)
        cbt = torch.tensor(cb, device=DEV, requires_grad=True)
        idx = _skip_if_tc_missing(lambda: nsg.vq(zt, cbt))
        assert idx.dtype == torch.int64 and tuple(idx.shape) == (16, 128) and not idx.requires_grad
        codes, flat = nsg.vq_st(zt, cbt)
        assert tuple(flat.shape) == (2048,) and codes.shape == zt.shape and torch.equal(idx.view(-1), flat)
        ref_idx = g[tag + '/indices'].astype(np.int64)
        n_mis, n_viol = vq_ref.index_parity(idx.cpu().numpy(), ref_idx, z, cb)
        assert n_viol == 0, (n_mis, n_viol)
        assert n_mis <= 8
        assert torch.equal(codes.detach().view(-1, 64), cbt.detach()[flat])       # bit-exact gather
        codes.backward(torch.tensor(gout, device=DEV))
        assert torch.equal(zt.grad.cpu(), torch.tensor(gout))                     # straight-through clone
        if n_mis == 0:
            ref_gc = g[tag + '/grad_codebook']
            assert np.abs(cbt.grad.cpu().numpy() - ref_gc).max() <= 1e-5 * np.abs(ref_gc).max()
        gi, gc = vq_ref.vq_st_backward(gout, flat.cpu().numpy(), cb.shape)
        assert np.abs(cbt.grad.cpu().numpy() - gc).max() <= 1e-5 * np.abs(gc).max()
    finally:
        nsg.vector_quantization.DEFAULT_MODE = None


@pytest.mark.parametrize('mode', [0, 1])
def test_vq_ties_and_errors(golden_dir, mode):
    g = np.load(os.path.join(golden_dir, 'vq_c1.npz'))
    rng = np.random.RandomState(int(g['ties/seed']))
    cb = rng.standard_normal((8, 4)).astype(np.float32)
    cb[5] = cb[2]
    cb[7] = cb[0]
    z = np.concatenate([cb, rng.standard_normal((8, 4)).astype(np.float32)], 0)
    nsg.vector_quantization.DEFAULT_MODE = mode
    try:
        idx = _skip_if_tc_missing(lambda: nsg.vq(torch.tensor(z, device=DEV), torch.tensor(cb, device=DEV)))
    finally:
        nsg.vector_quantization.DEFAULT_MODE = None
    assert np.array_equal(idx.cpu().numpy(), g['ties/indices'].astype(np.int64))
    with pytest.raises(RuntimeError, match='VectorQuantization'):
        nsg.vector_quantization.VectorQuantization.backward(None, None)
    assert str(g['vq_backward_message']).startswith('Trying to call `.grad()`')
    with pytest.raises(RuntimeError):
        nsg.vq(torch.zeros(4, 6, device=DEV).t(), torch.zeros(3, 4, device=DEV))   # non-contiguous like .view


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('k,dups', [
    (512, [(3, 300, 511)]),                      # three equal codewords, both sides of the 256-column MMA boundary
    (512, [(0, 255), (256, 257), (100, 400)]),   # pairs inside one N=256 MMA, adjacent, and straddling the boundary
    (1024, [(7, 519), (300, 800, 1023)]),        # duplicates in different 512-codeword chunks (streaming kernel)
    (8192, [(1, 4097, 8191), (512, 513)]),
])
def test_vq_tensor_core_lowest_index_ties(k, dups, dtype):
    """Exact duplicate codewords through the tcgen05 search (vq_search_res_kernel for K <= 512, vq_search_tc_kernel
    above): rows equal to a duplicated codeword must come back with the LOWEST index, as torch.min returns on CPU
    (vector_quantization.py:19) -- the top-2 trackers and the fp32 re-score are where a later duplicate
    could win."""
    d = 64
    rng = np.random.RandomState(k + len(dups))
    cb = rng.standard_normal((k, d)).astype(np.float32)
    if dtype == torch.bfloat16:
        cb = torch.tensor(cb).bfloat16().float().numpy()          # rows of z equal to codewords stay equal in bf16
    rows, want = [], []
    for group in dups:
        for j in group[1:]:
            cb[j] = cb[group[0]]
        for _ in range(40):                                        # spread over several 128-row tiles
            rows.append(cb[group[0]].copy())
            want.append(group[0])
    filler = rng.standard_normal((700, d)).astype(np.float32)
    order = rng.permutation(len(rows) + len(filler))
    z = np.concatenate([np.stack(rows), filler], 0)[order]
    want_all = np.concatenate([np.array(want), -np.ones(len(filler), dtype=np.int64)])[order]
    zt = torch.tensor(z, device=DEV).to(dtype)
    idx = torch.empty(z.shape[0], dtype=torch.int64, device=DEV)
    assert ops.vq_mode_serves(k, d, 1), 'the tcgen05 search must serve D = 64, K <= 8192'
    ops.vq_forward(zt, torch.tensor(cb, device=DEV), idx, None, None, mode=1)
    torch.cuda.synchronize()
    got = idx.cpu().numpy()
    sel = want_all >= 0
    assert np.array_equal(got[sel], want_all[sel]), (got[sel][:16], want_all[sel][:16])
    ref = vq_ref.vq_indices(zt.float().cpu().numpy(), cb)
    n_mis, n_viol = vq_ref.index_parity(got, ref, zt.float().cpu().numpy(), cb)
    assert n_viol == 0, (n_mis, n_viol)
    # near-duplicates: the later copy is closer by one ulp in one coordinate -> it must win for z equal to it
    cb2 = cb.copy()
    g0 = dups[0]
    cb2[g0[1]] = cb2[g0[0]]
    cb2[g0[1], 5] = np.nextafter(cb2[g0[1], 5], np.float32(10.0))
    z2 = np.stack([cb2[g0[1]]] * 130).astype(np.float32)
    if dtype == torch.float32:                                    # (bf16 inputs cannot represent the one-ulp neighbour)
        idx2 = torch.empty(130, dtype=torch.int64, device=DEV)
        ops.vq_forward(torch.tensor(z2, device=DEV), torch.tensor(cb2, device=DEV), idx2, None, None, mode=1)
        # the exact match (fp64 distance 0) must win over its one-ulp neighbour with the LOWER index.  (The reference's
        # fp32 expansion |e|^2 + |z|^2 - 2 z.e rounds both to the same value and returns the lower index: the farther
        # codeword by (1 ulp)^2.  Ours is the true argmin; the symmetric 1e-6 rule relative to the reference's
        # distance cannot express "closer than the reference", so the winner itself is asserted.)
        got2 = idx2.cpu().numpy()
        assert (got2 == g0[1]).all(), got2[:8]


@pytest.mark.parametrize('n,k,d', [(1, 1, 1), (5, 3, 1), (100, 17, 7), (333, 512, 64), (4096, 1024, 128), (257, 640, 256)])
@pytest.mark.parametrize('mode', [0, 1])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_vq_shapes_vs_oracle(n, k, d, mode, dtype):
    rng = np.random.RandomState(n + k + d)
    z = rng.standard_normal((n, d)).astype(np.float32)
    cb = rng.standard_normal((k, d)).astype(np.float32)
    zt = torch.tensor(z, device=DEV).to(dtype)
    z_used = zt.float().cpu().numpy()
    idx = torch.empty(n, dtype=torch.int64, device=DEV)
    codes = torch.empty_like(zt)
    sq = torch.zeros(1, dtype=torch.float64, device=DEV)
    _skip_if_tc_missing(lambda: ops.vq_forward(zt, torch.tensor(cb, device=DEV), idx, codes, sq, mode=mode))
    torch.cuda.synchronize()
    ref = vq_ref.vq_indices(z_used, cb)
    n_mis, n_viol = vq_ref.index_parity(idx.cpu().numpy(), ref, z_used, cb)
    assert n_viol == 0, (n_mis, n_viol)
    got = idx.cpu().numpy()
    assert np.array_equal(codes.float().cpu().numpy(), torch.tensor(cb[got]).to(dtype).float().numpy())
    want = ((z_used.astype(np.float64) - cb[got].astype(np.float64)) ** 2).sum()
    assert abs(sq.item() - want) <= (1e-5 if dtype == torch.float32 else 1e-5) * want + 1e-9


def test_scatter_add_gather_and_bottleneck_grad():
    rng = np.random.RandomState(21)
    n, k, d = 1000, 37, 64
    rows, sub = _rand(rng, n, d), _rand(rng, n, d)
    idx = torch.tensor(rng.randint(0, k, n))
    table = torch.zeros(k, d, device=DEV)
    counts = torch.zeros(k, device=DEV)
    ops.scatter_add(rows.to(DEV), idx.to(DEV), table, alpha=0.5, sub=sub.to(DEV), counts=counts)
    ref = torch.zeros(k, d, dtype=torch.float64).index_add_(0, idx, 0.5 * (rows.double() - sub.double()))
    assert relerr(table, ref) < 1e-5
    assert torch.equal(counts.cpu(), torch.bincount(idx, minlength=k).float())
    # EMA extension statistics agree with the (unpinned) oracle restatement
    _, _, _, c_ref, s_ref = vq_ref.ema_update(rows.numpy(), idx.numpy(), np.ones(k, np.float32), np.zeros((k, d), np.float32))
    t2 = torch.zeros(k, d, device=DEV)
    ops.scatter_add(rows.to(DEV), idx.to(DEV), t2)
    assert relerr(t2, s_ref) < 1e-5 and np.array_equal(counts.cpu().numpy(), c_ref)
    out = torch.empty(n, d, device=DEV)
    ops.gather_rows(table, idx.to(DEV), out)
    assert torch.equal(out, table[idx.to(DEV)])
    dz = torch.empty(n, d, device=DEV)
    ops.vq_backward_inputs(rows.to(DEV), sub.to(DEV), out, 0.25, dz)
    assert relerr(dz, rows.double() + 0.25 * (sub.double() - out.double().cpu())) < 1e-6
    # odd widths take the scalar path
    t3 = torch.zeros(5, 3, device=DEV)
    r3 = _rand(rng, 11, 3)
    i3 = torch.tensor(rng.randint(0, 5, 11))
    ops.scatter_add(r3.to(DEV), i3.to(DEV), t3)
    assert relerr(t3, torch.zeros(5, 3, dtype=torch.float64).index_add_(0, i3, r3.double())) < 1e-6


@pytest.mark.parametrize('n,k,d,dtype,dist', [
    (40960, 512, 64, torch.bfloat16, 'uniform'),      # config 2: 10 partitions x 30 codeword groups
    (40960, 512, 64, torch.float32, 'skew'),          # most rows on a handful of codewords (run aggregation)
    (5000, 37, 64, torch.float32, 'uniform'),         # ragged last partition, K not a multiple of G
    (9001, 1000, 128, torch.bfloat16, 'uniform'),     # 32 lanes per row
    (8192, 64, 256, torch.float32, 'one'),            # two vector groups per lane; every row on one codeword
    (20000, 8192, 12, torch.float32, 'uniform'),      # D/4 not a power of two; mostly untouched codewords
    (1 << 18, 512, 64, torch.bfloat16, 'uniform'),    # large-N regime: few groups, big shared table
])
def test_scatter_add_aggregated_kernel(n, k, d, dtype, dist):
    """Block-aggregated scatter-add (scatter_ema.cu) == index_add_ (vector_quantization.py:58-61) == the
    per-element kernel; counts == bincount; out-of-range indices are skipped."""
    rng = np.random.RandomState(n % 1000 + d)
    rows, sub = _rand(rng, n, d).to(dtype), _rand(rng, n, d).to(dtype)
    if dist == 'uniform':
        idx = rng.randint(0, k, n)
    elif dist == 'skew':
        idx = np.minimum((rng.exponential(3.0, n)).astype(np.int64), k - 1)
    else:
        idx = np.full(n, k // 2, np.int64)
    idx = torch.tensor(idx)
    idx[::997] = -1                 # ignored, like the per-element kernel
    idx[5::1013] = k + 3
    ok = (idx >= 0) & (idx < k)
    ref = torch.zeros(k, d, dtype=torch.float64).index_add_(0, idx[ok], 0.25 * (rows.double() - sub.double())[ok])
    cref = torch.bincount(idx[ok], minlength=k).float()
    lib = nsg._cabi.load()
    out = {}
    for mode in (0, 1):
        prev = lib.nsg_scatter_mode(mode)
        try:
            table = torch.zeros(k, d, device=DEV)
            counts = torch.zeros(k, device=DEV)
            ops.scatter_add(rows.to(DEV), idx.to(DEV), table, alpha=0.25, sub=sub.to(DEV), counts=counts)
            # accumulate semantics: a second call without sub / counts adds on top
            ops.scatter_add(rows.to(DEV), idx.to(DEV), table, alpha=-0.25)
            ops.scatter_add(sub.to(DEV), idx.to(DEV), table, alpha=0.25)
            torch.cuda.synchronize()
        finally:
            lib.nsg_scatter_mode(prev)
        out[mode] = table
        assert torch.equal(counts.cpu(), cref), mode
    # three passes cancel exactly up to fp32 rounding: compare the first pass through a fresh call instead
    for mode in (0, 1):
        prev = lib.nsg_scatter_mode(mode)
        try:
            table = torch.zeros(k, d, device=DEV)
            ops.scatter_add(rows.to(DEV), idx.to(DEV), table, alpha=0.25, sub=sub.to(DEV))
        finally:
            lib.nsg_scatter_mode(prev)
        assert relerr(table, ref) < 2e-6, (mode, relerr(table, ref))
        assert float(out[mode].abs().max()) < 1e-3 * float(ref.abs().max() + 1)


def test_ema_update_vs_oracle_restatement():
    """EMA codebook extension (not in the reference: parity unpinned) against oracle/vq_ref.ema_update."""
    rng = np.random.RandomState(3)
    n, k, d = 6000, 128, 64
    z = _rand(rng, n, d)
    idx = torch.tensor(rng.randint(0, k - 9, n))          # nine dead codewords
    ema_n = torch.tensor(rng.rand(k).astype(np.float32) * 5)
    ema_m = _rand(rng, k, d)
    n_ref, m_ref, e_ref, c_ref, s_ref = vq_ref.ema_update(z.numpy(), idx.numpy(), ema_n.numpy(), ema_m.numpy(), 0.97, 1e-5)
    emb = nsg.VQEmbedding(k, d).to(DEV)
    gn, gm = ema_n.to(DEV), ema_m.to(DEV).contiguous()
    z_nchw = z.view(n // 100, 10, 10, d).permute(0, 3, 1, 2).to(DEV)
    counts, sums = nsg.models.ema_codebook_update(emb, z_nchw, idx.to(DEV), gn, gm, decay=0.97, eps=1e-5)
    assert np.array_equal(counts.cpu().numpy(), c_ref) and relerr(sums, s_ref) < 1e-5
    assert relerr(gn, n_ref) < 1e-6 and relerr(gm, m_ref) < 1e-5
    assert relerr(emb.embedding.weight.detach(), e_ref) < 1e-5


def test_adam_known_answer(golden_dir):
    """torch.optim.Adam KAT recorded from the reference's optimizer (tests/golden/adam_kat.npz)."""
    g = np.load(os.path.join(golden_dir, 'adam_kat.npz'))
    rng = np.random.RandomState(41)
    shapes = [(64,), (64, 16), (5,)]
    ps = [rng.standard_normal(s).astype(np.float32) for s in shapes]
    sizes = [p.size for p in ps]
    flat = torch.tensor(np.concatenate([p.reshape(-1) for p in ps]), device=DEV)
    m, v = torch.zeros_like(flat), torch.zeros_like(flat)
    step = torch.zeros((), dtype=torch.int64, device=DEV)
    for _ in range(3):
        gs = []
        for i in range(3):
            scale = [1.0, 1e-4, 1e-8][i]
            gs.append((scale * rng.standard_normal(shapes[i])).astype(np.float32).reshape(-1))
        ops.adam_step(flat, torch.tensor(np.concatenate(gs), device=DEV), m, v, step, 1e-3)
    assert int(step) == 3
    out = flat.cpu().numpy()
    o = 0
    for i, s in enumerate(sizes):
        assert np.allclose(out[o:o + s].reshape(shapes[i]), g['p%d' % i], rtol=2e-6, atol=2e-7), i
        o += s


def test_recon_loss_bf16_fused_and_loss_finalize():
    """nsg_recon_loss_bf16 = nsg_recon_loss on the bf16-rounded x_tilde, plus the two conversions around it
    (train.py:118-129); nsg_loss_finalize = the three loss scalars of train.py:129-134 from the two sums."""
    rng = np.random.RandomState(18)
    B, H, Wt, Wc = 3, 10, 8, 11
    xt16 = torch.tensor(np.tanh(rng.standard_normal((B, H, Wt))).astype(np.float32)).to(torch.bfloat16)
    c = torch.tensor(rng.uniform(0, 1, (B, H, Wc)).astype(np.float32))
    x64 = xt16.double()
    loss = F.mse_loss(F.pad(x64, (0, Wc - Wt)), c.double())
    dref = 2.0 * (x64 - c.double()[..., :Wt]) / c.numel() * (1.0 - x64 * x64)
    sums = torch.zeros(2, dtype=torch.float64, device=DEV)
    dpre = torch.empty(B, H, Wt, dtype=torch.bfloat16, device=DEV)
    x32 = torch.full((B, H, Wt), 7.0, device=DEV)
    dsum = torch.full((1,), 0.125, device=DEV)
    ops.recon_loss_bf16(xt16.to(DEV), c.to(DEV), sums[0:1], dpre, x32, dpre_sum=dsum)
    torch.cuda.synchronize()
    assert abs(sums[0].item() / c.numel() - loss.item()) < 1e-6 * max(1.0, loss.item())
    # the output layer's bias gradient: += the sum of the ROUNDED dpre values (what ops.colsum(dpre) would return)
    assert abs(dsum.item() - 0.125 - dpre.double().sum().item()) < 1e-5 * dpre.double().abs().sum().item()
    assert torch.equal(x32.cpu(), xt16.float())
    assert torch.equal(dpre.cpu(), dref.float().to(torch.bfloat16)) or relerr(dpre, dref) < 4e-3
    # same numbers as the fp32 kernel fed the widened tensor
    sums2 = torch.zeros(1, dtype=torch.float64, device=DEV)
    dpre2 = torch.empty(B, H, Wt, device=DEV)
    ops.recon_loss(xt16.float().to(DEV), c.to(DEV), sums2, dpre2)
    assert sums2.item() == sums[0].item() or abs(sums2.item() - sums[0].item()) < 1e-9 * abs(sums2.item())
    assert torch.equal(dpre2.to(torch.bfloat16), dpre)
    # optional outputs may be omitted
    s3 = torch.zeros(1, dtype=torch.float64, device=DEV)
    ops.recon_loss_bf16(xt16.to(DEV), c.to(DEV), s3)
    assert abs(s3.item() - sums[0].item()) < 1e-9 * abs(s3.item())
    with pytest.raises(Exception):
        ops.recon_loss_bf16(xt16.float().to(DEV), c.to(DEV), s3)
    sums[1] = 5.0
    out = ops.loss_finalize(sums, 1.0 / c.numel(), 0.25)
    torch.cuda.synchronize()
    assert out.tolist() == [sums[0].item() * (1.0 / c.numel()), 1.25, 1.25]


def test_adam_vector_path_and_gradient_clearing():
    """16-byte path (n % 4 == 0) == scalar path bit for bit; zero_grads clears exactly the gradients, after use."""
    rng = np.random.RandomState(5)
    n = 4096 + 8
    p0 = torch.tensor(rng.standard_normal(n).astype(np.float32), device=DEV)
    res = {}
    for tag, off in (('vec', 0), ('scalar', 1)):          # a 4-byte offset forces the scalar kernel on the same data
        buf = [torch.zeros(n + 4, device=DEV) for _ in range(4)]
        p, g, m, v = [b[off:off + n] for b in buf]
        p.copy_(p0)
        step = torch.zeros((), dtype=torch.int64, device=DEV)
        rg = np.random.RandomState(6)
        for it in range(3):
            g.copy_(torch.tensor(rg.standard_normal(n).astype(np.float32)))
            ops.adam_step(p, g, m, v, step, 1e-3, grad_scale=0.5, zero_grads=(it == 2))
            assert (float(g.abs().max()) == 0.0) == (it == 2)
        assert int(step) == 3
        res[tag] = (p.clone(), m.clone(), v.clone())
        # one-launch form: the last block advances the step counter (ticket left at zero) -- same numbers
        p2, g2, m2, v2 = [torch.zeros(n + 4, device=DEV)[off:off + n] for _ in range(4)]
        p2.copy_(p0)
        step2, ticket = torch.zeros((), dtype=torch.int64, device=DEV), torch.zeros((), dtype=torch.int32, device=DEV)
        rg = np.random.RandomState(6)
        for it in range(3):
            g2.copy_(torch.tensor(rg.standard_normal(n).astype(np.float32)))
            l0 = ops.launch_count
            ops.adam_step(p2, g2, m2, v2, step2, 1e-3, grad_scale=0.5, zero_grads=(it == 2), ticket=ticket)
            assert ops.launch_count - l0 == 1 and int(step2) == it + 1 and int(ticket) == 0
        assert torch.equal(p2, res[tag][0]) and torch.equal(m2, res[tag][1]) and torch.equal(v2, res[tag][2])
    for a, b in zip(res['vec'], res['scalar']):
        assert torch.equal(a, b)
    # against torch.optim.Adam
    q = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([q], lr=1e-3)
    rg = np.random.RandomState(6)
    for it in range(3):
        q.grad = torch.tensor(rg.standard_normal(n).astype(np.float32), device=DEV) * 0.5
        opt.step()
    assert relerr(res['vec'][0], q.detach()) < 1e-6
