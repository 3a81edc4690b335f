"""GPU tests of the reference's other model families on the B200 kernels: the convolutional VAE and
the GatedPixelCNN prior against fixtures produced by the UNMODIFIED reference (tests/golden/n4_*.npz, made by
oracle/gen_golden_n4.py) and against the oracle restatement (oracle/n4_ref.py).  fp32 kernels: 1e-5 on tensors and
losses, 2e-4 on gradients (sums over up to 10^4 pixels of products of fp32 values; the reference's own run-to-run
noise between thread counts is ~1e-6)."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import relerr

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import ops, models_n4
    from oracle import n4_ref as R

DEV = 'cuda'
PRE_BN_BIAS = ['encoder.%d.bias' % i for i in (0, 3, 6, 9)] + ['decoder.%d.bias' % i for i in (0, 3, 6)]


def _load(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name))
    return g, {k[6:]: torch.tensor(g[k]) for k in g.files if k.startswith('state/')}


@pytest.mark.parametrize('kh,kw,ph,pw,gather,cin,cout,H,W,B,crop', [
    (5, 5, 0, 0, 0, 16, 24, 12, 14, 2, False),      # VAE encoder.6
    (5, 5, 0, 0, 1, 16, 24, 8, 10, 2, False),       # VAE decoder.3 (transposed-like gather)
    (4, 7, 3, 3, 0, 8, 16, 9, 9, 3, True),          # PixelCNN vertical stack, k = 7, cropped rows
    (1, 4, 0, 3, 0, 8, 16, 9, 11, 3, True),         # PixelCNN horizontal stack, k = 7, cropped columns
    (2, 3, 1, 1, 0, 64, 128, 16, 16, 1, True),
    (1, 1, 0, 0, 0, 70, 33, 5, 7, 2, False),        # channel counts that are not multiples of the tile
])
def test_conv_rect_forward_and_wgrad_vs_torch(kh, kw, ph, pw, gather, cin, cout, H, W, B, crop):
    rng = np.random.RandomState(kh * 100 + kw * 10 + cin)
    x = torch.tensor(rng.standard_normal((B, cin, H, W)).astype(np.float32))
    bias = torch.tensor(rng.standard_normal(cout).astype(np.float32))
    if gather == 0:
        w = torch.tensor((rng.standard_normal((cout, cin, kh, kw)) * 0.1).astype(np.float32))
        full = F.conv2d(x.double(), w.double(), bias.double(), 1, (ph, pw))
        wci, wco = kh * kw, cin * kh * kw
    else:
        w = torch.tensor((rng.standard_normal((cin, cout, kh, kw)) * 0.1).astype(np.float32))
        full = F.conv_transpose2d(x.double(), w.double(), bias.double(), 1, (ph, pw))
        wci, wco = cout * kh * kw, kh * kw
    oh, ow = (H, W) if crop else tuple(full.shape[2:])
    want = full[:, :, :oh, :ow]
    xg = x.permute(0, 2, 3, 1).contiguous().to(DEV)
    out = torch.empty((B, oh, ow, cout), device=DEV)
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    ops.conv_rect(xg, w.to(DEV), out, kh, kw, ph, pw, gather, wci, wco, bias=bias.to(DEV), stats=stats)
    torch.cuda.synchronize()
    assert relerr(out.permute(0, 3, 1, 2), want) < 1e-5
    assert relerr(stats[:cout], want.sum((0, 2, 3))) < 1e-5 and relerr(stats[cout:], (want ** 2).sum((0, 2, 3))) < 1e-5
    # add_src + ReLU epilogue
    add = torch.tensor(rng.standard_normal((B, oh, ow, cout)).astype(np.float32)).to(DEV)
    out2 = torch.empty_like(out)
    ops.conv_rect(xg, w.to(DEV), out2, kh, kw, ph, pw, gather, wci, wco, bias=bias.to(DEV), add_src=add, act=2)
    assert relerr(out2.permute(0, 3, 1, 2), (want + add.cpu().double().permute(0, 3, 1, 2)).clamp_min(0)) < 1e-5
    if gather == 0 and not crop:
        # weight gradient of the same convolution: P = x, Q = dy
        dy = torch.tensor(rng.standard_normal(tuple(want.shape)).astype(np.float32))
        wd = w.double().requires_grad_(True)
        F.conv2d(x.double(), wd, None, 1, (ph, pw)).backward(dy.double())
        dw = torch.zeros_like(w).to(DEV)
        colsum = torch.zeros(cout, device=DEV)
        ops.conv_rect_wgrad(xg, dy.permute(0, 2, 3, 1).contiguous().to(DEV), dw, kh, kw, ph, pw, wci, wco, q_colsum=colsum)
        torch.cuda.synchronize()
        assert relerr(dw, wd.grad) < 1e-5
        assert relerr(colsum, dy.double().sum((0, 2, 3))) < 1e-5
    with pytest.raises(Exception):
        ops.conv_rect(xg, w.to(DEV), torch.empty((B, full.shape[2] + 1, ow, cout), device=DEV), kh, kw, ph, pw, gather, wci, wco)


@pytest.mark.parametrize('tag', ['mel', 'img'])
def test_vae_step_vs_reference_fixture(golden_dir, tag):
    g, p = _load(golden_dir, 'n4_vae_%s.npz' % tag)
    input_dim, dim, z_dim = [int(v) for v in g['meta'][:3]]
    m = nsg.VAE(input_dim, dim, z_dim)
    m.load_state_dict(p, strict=True)
    m = m.to(DEV).train()
    x, eps = torch.tensor(g['x']).to(DEV), torch.tensor(g['eps']).to(DEV)
    x_tilde, kl = m(x, eps=eps)
    assert tuple(x_tilde.shape) == g['x_tilde'].shape and kl.dim() == 0
    target = F.pad(x_tilde, (0, x.shape[3] - x_tilde.shape[3]))            # train.py:58-60
    loss = models_n4.mse_loss(target, x, kl)                               # loss.py:22-24
    loss.backward()
    torch.cuda.synchronize()
    assert relerr(x_tilde, g['x_tilde']) < 1e-5
    assert abs(kl.item() - float(g['kl'])) < 1e-5 * abs(float(g['kl']))
    assert abs(loss.item() - float(g['loss'])) < 1e-5 * abs(float(g['loss']))
    worst = ('', 0.0)
    for k, v in m.named_parameters():
        if k in PRE_BN_BIAS:
            continue
        e = relerr(v.grad, g['grad/' + k])
        worst = max(worst, (k, e), key=lambda t: t[1])
        assert e < 2e-4, (k, e)
    print('VAE %s: x_tilde %.2e, worst gradient %s %.2e' % (tag, relerr(x_tilde, g['x_tilde']), worst[0], worst[1]))
    sd = m.state_dict()
    for k in g.files:
        if k.startswith('after/') and 'num_batches' not in k:
            assert relerr(sd[k[6:]], g[k]) < 1e-4, k
        elif k.startswith('after/'):
            assert int(sd[k[6:]].item()) == int(g[k])
    # its own noise: a second call draws eps on the device and still returns finite values of the same shapes
    xt2, kl2 = m(x)
    assert tuple(xt2.shape) == tuple(x_tilde.shape) and torch.isfinite(xt2).all() and torch.isfinite(kl2)
    with pytest.raises(Exception):
        m(x.cpu())


def test_vae_oracle_at_another_size_and_eval_mode():
    torch.manual_seed(5)
    m = nsg.VAE(1, 64, 16).to(DEV).train()
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    x = torch.rand(4, 1, 80, 64)
    with torch.no_grad():
        zshape = (4, 16, 14, 10)
    eps = torch.randn(*zshape)
    r = R.vae_step_grads(p, x, eps)
    xt, kl = m(x.to(DEV), eps=eps.to(DEV))
    loss = models_n4.mse_loss(F.pad(xt, (0, 0)), x.to(DEV), kl)
    loss.backward()
    assert relerr(xt, r['x_tilde']) < 1e-5 and abs(loss.item() - r['loss'].item()) < 1e-5 * abs(r['loss'].item())
    for k, v in m.named_parameters():
        if k not in PRE_BN_BIAS:
            assert relerr(v.grad, r['grads'][k]) < 2e-4, k
    m.eval()
    with torch.no_grad():
        xe, _ = m(x.to(DEV), eps=eps.to(DEV))
    assert tuple(xe.shape) == tuple(xt.shape) and torch.isfinite(xe).all()


@pytest.mark.parametrize('tag', ['k32', 'latents'])
def test_pixelcnn_forward_vs_reference_fixture(golden_dir, tag):
    g, p = _load(golden_dir, 'n4_pixelcnn_%s.npz' % tag)
    input_dim, dim, n_layers, n_classes = [int(v) for v in g['meta'][:4]]
    m = nsg.GatedPixelCNN(input_dim, dim, n_layers, n_classes)
    m.load_state_dict(p, strict=True)
    m = m.to(DEV).eval()
    logits = m(torch.tensor(g['x']).to(DEV), torch.tensor(g['label']).to(DEV))
    torch.cuda.synchronize()
    assert tuple(logits.shape) == g['logits'].shape
    assert relerr(logits, g['logits']) < 1e-5
    sd = m.state_dict()
    for k in ('layers.0.vert_stack.weight', 'layers.0.horiz_stack.weight'):     # make_causal masked them in place
        assert np.array_equal(sd[k].cpu().numpy(), g['after/' + k])


def test_pixelcnn_non_square_and_generate():
    torch.manual_seed(9)
    m = nsg.GatedPixelCNN(24, 16, 3, 5).to(DEV).eval()
    p = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    # non-square input (what the VQ-VAE's 20 x W latent grids are): the reference's crop cannot run there, the oracle's
    # per-axis crop is the evident intent
    x = torch.randint(0, 24, (2, 5, 9))
    label = torch.tensor([1, 4])
    got = m(x.to(DEV), label.to(DEV))

    def oracle_logits(xx):
        q = {k: v.clone() for k, v in p.items()}
        emb = q['embedding.weight'][xx.reshape(-1)].reshape(xx.shape + (-1,)).permute(0, 3, 1, 2)
        xv = xh = emb
        for i in range(3):
            pre, k = 'layers.%d.' % i, (7 if i == 0 else 3)
            wv, wh = q[pre + 'vert_stack.weight'], q[pre + 'horiz_stack.weight']
            if i == 0:
                wv[:, :, -1] = 0
                wh[:, :, :, -1] = 0
            h = q[pre + 'class_cond_embedding.weight'][label]
            hv = F.conv2d(xv, wv, q[pre + 'vert_stack.bias'], 1, (k // 2, k // 2))[:, :, :xv.shape[2], :]
            gate = lambda t: torch.tanh(t.chunk(2, 1)[0]) * torch.sigmoid(t.chunk(2, 1)[1])      # noqa: E731
            out_v = gate(hv + h[:, :, None, None])
            hh = F.conv2d(xh, wh, q[pre + 'horiz_stack.bias'], 1, (0, k // 2))[:, :, :, :xh.shape[3]]
            out = gate(F.conv2d(hv, q[pre + 'vert_to_horiz.weight'], q[pre + 'vert_to_horiz.bias']) + hh + h[:, :, None, None])
            res = F.conv2d(out, q[pre + 'horiz_resid.weight'], q[pre + 'horiz_resid.bias'])
            xv, xh = out_v, (res + xh if i > 0 else res)
        hid = F.relu(F.conv2d(xh, q['output_conv.0.weight'], q['output_conv.0.bias']))
        return F.conv2d(hid, q['output_conv.2.weight'], q['output_conv.2.bias'])
    assert relerr(got, oracle_logits(x)) < 1e-5
    # generate: raster-scan sampling with supplied uniforms == the oracle's inverse-CDF draws on the oracle's logits
    u = torch.rand(4, 4, 2, generator=torch.Generator().manual_seed(3))
    xs = m.generate(label.to(DEV), shape=(4, 4), batch_size=2, uniforms=u).cpu()
    assert xs.shape == (2, 4, 4) and xs.dtype == torch.int64 and int(xs.min()) >= 0 and int(xs.max()) < 24
    ref = torch.zeros(2, 4, 4, dtype=torch.int64)
    agree = 0
    for i in range(4):
        for j in range(4):
            ref[:, i, j] = R.sample_inverse_cdf(oracle_logits(ref)[:, :, i, j], u[i, j])
            agree += int((ref[:, i, j] == xs[:, i, j]).sum())
            ref[:, i, j] = xs[:, i, j]          # teacher-force so one boundary-case draw cannot derail the rest
    assert agree >= 31, agree                    # (a uniform within 1e-6 of a CDF step may fall either side in fp32)
    assert torch.equal(xs, m.generate(label.to(DEV), shape=(4, 4), batch_size=2, uniforms=u).cpu())     # deterministic
    free = m.generate(label.to(DEV), shape=(3, 3), batch_size=2)
    assert free.shape == (2, 3, 3)


def test_softmax_sample_distribution_and_gated_act():
    logits = torch.tensor([[-5.0, 0.0, 5.0, 0.0, -1.0] + [-30.0] * 60], device=DEV).repeat(20000, 1).contiguous()
    u = torch.rand(20000, device=DEV, generator=torch.Generator(device=DEV).manual_seed(0))
    out = torch.empty(20000, dtype=torch.int64, device=DEV)
    ops.softmax_sample(logits, 65, 0, 20000, 65, u, out, 1)
    freq = torch.bincount(out.cpu(), minlength=65).double() / 20000
    assert (freq - torch.softmax(logits[0].double().cpu(), -1)).abs().max() < 0.01
    assert torch.equal(out.cpu(), R.sample_inverse_cdf(logits.cpu(), u.cpu())) or \
        (out.cpu() != R.sample_inverse_cdf(logits.cpu(), u.cpu())).sum() <= 3
    x = torch.randn(2, 6, 3, 5, device=DEV)
    got = nsg.GatedActivation()(x)
    a, b = x.chunk(2, dim=1)
    assert relerr(got, torch.tanh(a.double()) * torch.sigmoid(b.double())) < 1e-6


def test_main_trains_the_vae(tmp_path, monkeypatch):
    """``main.py --model vae`` (main.py:117-118, 130-132) end to end on synthetic mels: train_vae / test_vae loops, the
    reference's checkpoint layout with torch.optim.Adam's state, the per-epoch sample dump and its audio."""
    from neural_sound_generation_b200 import main as M
    monkeypatch.chdir(tmp_path)
    argv = ['--model', 'vae', '--dataset', 'ljspeech', '--dim', '32', '--z-dim', '8', '--batch-size', '2',
            '--synthetic-batches', '2', '--synthetic-frames', '64', '--log-interval', '1', '--epochs', '1',
            '--sampledir', str(tmp_path / 'results')]
    assert M.main(argv) == 0
    ck = torch.load(str(tmp_path / 'models' / 'vae' / 'checkpoint_ljspeech_32_8.pth.tar'), map_location='cpu')
    assert ck['arch'] == 'vae' and ck['epoch'] == 1
    assert set(ck['state_dict'].keys()) == set(nsg.VAE(1, 32, 8).state_dict().keys())
    rec = np.load(str(tmp_path / 'results' / 'ljspeech' / 'reconstruction_vae_data_ljspeech_dim_32_z_dim_8_epoch_1.npy'))
    assert rec.shape == (2, 80, 64) and np.isfinite(rec).all()
