"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU.

Run in the authoring container only (needs /root/reference):
    python -m oracle.gen_golden

Imports /root/reference/src/{vector_quantization,models,train,test}.py through
sys.path, stubbing ``util`` (its real module needs librosa/lws/umap; only the
one-line ``unsqueeze_to_device`` is imported by train.py:15).  Inputs and
weights come from numpy's MT19937 via oracle.vqvae_ref.make_state/make_mel, so
tests regenerate them bit-identically without the reference and the fixtures
hold only reference OUTPUTS.  Large gradient / weight tensors are stored as
(sum, L2 norm, 32 strided samples); everything <= 8192 elements is stored whole.
"""
import argparse
import contextlib
import io
import os
import sys
import types

import numpy as np
import torch

from . import vqvae_ref as R

REF_SRC = '/root/reference/src'
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def import_reference():
    util = types.ModuleType('util')
    util.unsqueeze_to_device = lambda x, device: x.unsqueeze(1).contiguous().to(device)
    sys.modules['util'] = util
    sys.path.insert(0, REF_SRC)
    import vector_quantization as ref_vq   # noqa: E402
    import models as ref_models            # noqa: E402
    import train as ref_train              # noqa: E402
    import test as ref_test                # noqa: E402
    return ref_vq, ref_models, ref_train, ref_test


def summarize(t):
    """Whole tensor if small, else (sum, norm, strided samples)."""
    a = t.detach().cpu().numpy().astype(np.float32)
    if a.size <= 8192:
        return {'full': a}
    flat = a.reshape(-1).astype(np.float64)
    stride = max(1, flat.size // 32)
    return {'sum': np.float64(flat.sum()), 'norm': np.float64(np.sqrt((flat ** 2).sum())),
            'samples': a.reshape(-1)[::stride][:32].copy()}


def put(d, prefix, t):
    for k, v in summarize(t).items():
        d[prefix + '/' + k] = v


class _Loader(list):
    """In-memory stand-in for the LJSpeech DataLoader: yields (x, y, c, g, input_lengths)
    as dataloader.collate_fn does (dataloader.py:410-434) and exposes ``.dataset``."""
    def __init__(self, items):
        super().__init__(items)
        self.dataset = list(range(sum(int(i[2].shape[0]) for i in items)))


def loader_from_mels(mels):
    items = []
    for c in mels:
        bsz, _, t = c.shape
        x = torch.zeros(bsz, 1, t * 4)
        y = torch.zeros(bsz, t * 4, 1)
        items.append((x, y, c, None, torch.full((bsz,), t, dtype=torch.int64)))
    return _Loader(items)


def gen_vq(ref_vq):
    """Config C1: vq / vq_st forward + backward, N = 16*128, K = 512, D = 64."""
    out = {}
    for tag, seed in (('init', 11), ('spread', 12)):
        rng = np.random.RandomState(seed)
        z = torch.tensor(rng.standard_normal((16, 128, 64)).astype(np.float32), requires_grad=True)
        if tag == 'init':
            cb = rng.uniform(-1 / 512, 1 / 512, (512, 64))
        else:
            cb = rng.standard_normal((512, 64))
        cb = torch.tensor(cb.astype(np.float32), requires_grad=True)
        gout = torch.tensor(rng.standard_normal((16, 128, 64)).astype(np.float32))
        idx = ref_vq.vq(z, cb)
        codes, idx_flat = ref_vq.vq_st(z, cb)
        codes.backward(gout)
        assert torch.equal(idx.view(-1), idx_flat)
        out[tag + '/seed'] = np.int64(seed)
        out[tag + '/indices'] = idx.numpy().astype(np.int16)
        out[tag + '/codes_sum'] = np.float64(codes.detach().double().sum())
        out[tag + '/grad_inputs_equal_gout'] = np.bool_(torch.equal(z.grad, gout))
        out[tag + '/grad_codebook'] = cb.grad.numpy()
    # ties: duplicated codewords must resolve to the lowest index (torch.min)
    rng = np.random.RandomState(13)
    cb = rng.standard_normal((8, 4)).astype(np.float32)
    cb[5] = cb[2]
    cb[7] = cb[0]
    z = np.concatenate([cb, rng.standard_normal((8, 4)).astype(np.float32)], 0)
    idx = ref_vq.vq(torch.tensor(z), torch.tensor(cb))
    out['ties/seed'] = np.int64(13)
    out['ties/indices'] = idx.numpy().astype(np.int16)
    # backward of plain vq must raise
    try:
        zz = torch.tensor(z, requires_grad=True)
        ref_vq.VectorQuantization.backward(None, None)
        out['vq_backward_raises'] = np.bool_(False)
    except RuntimeError as e:
        out['vq_backward_raises'] = np.bool_(True)
        out['vq_backward_message'] = np.array(str(e))
    np.savez_compressed(os.path.join(OUT, 'vq_c1.npz'), **out)
    print('vq_c1.npz', {k: getattr(v, 'shape', None) for k, v in out.items()})


def load_into(model, params, buffers):
    sd = {k: v.clone() for k, v in params.items()}
    sd.update({k: v.clone() for k, v in buffers.items()})
    missing = model.load_state_dict(sd, strict=True)
    return missing


def gen_step(ref_models, ref_train, ref_test, tag, codebook, frames, seed, batch=2, steps=2):
    """Train-mode forward/backward + ``steps`` verbatim train_vqvae iterations + test_vqvae."""
    out = {}
    dim, k = 64, 512
    params, buffers = R.make_state(1, dim, k, seed, codebook=codebook)
    mels = [R.make_mel(batch, 80, frames, seed + 100 + i) for i in range(steps)]
    out['meta'] = np.array([1, dim, k, batch, frames, seed, steps], dtype=np.int64)
    out['relu_margin'] = np.float64(R.relu_margins(params, buffers, mels[0]))
    print(tag, 'relu margin', out['relu_margin'])
    # ---- single fwd/bwd (train.py:115-135 evaluated by the reference modules themselves), twice:
    # fp32 as the reference runs, and the same unmodified modules cast to fp64.  The fp64 run is the
    # fixture of record for tensors and gradients: on this tiny batch the reference's own fp32
    # gradients sit ~1e-3 (relative) away from its fp64 gradients, so they cannot pin anything to
    # 1e-5.  That self-noise is stored per tensor under 'noise32/'.
    import torch.nn.functional as F
    grads = {}
    for dt in (torch.float32, torch.float64):
        model = ref_models.VQVAE(1, dim, k)
        load_into(model, params, buffers)
        model = model.to(dt)
        model.train()
        c = mels[0].to(dt).unsqueeze(1)
        x_tilde, z_e_x, z_q_x = model(c)
        with torch.no_grad():
            latents = ref_models.vq(z_e_x.permute(0, 2, 3, 1).contiguous(), model.codebook.embedding.weight)
        target = torch.zeros(c.size(0), c.size(1), c.size(2), c.size(3), dtype=dt)
        target[:, :, :, :x_tilde.size(3)] = x_tilde
        lr_ = F.mse_loss(target, c)
        lv = F.mse_loss(z_q_x, z_e_x.detach())
        lc = F.mse_loss(z_e_x, z_q_x.detach())
        loss = lr_ + lv + 1.0 * lc
        model.zero_grad()
        loss.backward()
        grads[dt] = {n: q.grad.double().clone() for n, q in model.named_parameters()}
        sfx = '' if dt == torch.float64 else '32'
        out['x_tilde' + sfx] = x_tilde.detach().float().numpy()
        out['z_e_x' + sfx] = z_e_x.detach().float().numpy()
        out['indices' + sfx] = latents.numpy().astype(np.int16)
        out['losses' + sfx] = np.array([lr_.item(), lv.item(), lc.item(), loss.item()], dtype=np.float64)
        # fp64 gradients are the fixture of record ('grad/'); the reference's own fp32 gradients are kept too
        # ('grad32/'): a pre-activation that sits within rounding of a ReLU threshold makes the two differ by a
        # discrete O(1/pixels) step, and an fp32 implementation may land on either side (a tie, like equal distances).
        for name, prm in model.named_parameters():
            put(out, ('grad/' if dt == torch.float64 else 'grad32/') + name, prm.grad)
        if dt == torch.float64:
            for name, buf in model.named_buffers():
                if 'num_batches' not in name:
                    out['buf1/' + name] = buf.detach().float().numpy().copy()
    for name in grads[torch.float64]:
        a, b_ = grads[torch.float32][name], grads[torch.float64][name]
        out['noise32/' + name] = np.float64((a - b_).norm() / max(b_.norm().item(), 1e-300))
    # ---- verbatim train loop from a fresh copy of the same state
    model = ref_models.VQVAE(1, dim, k)
    load_into(model, params, buffers)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)           # main.py:124
    args = argparse.Namespace(dataset='ljspeech', beta=1.0, log_interval=1)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(io.StringIO()):
        ref_train.train_vqvae(args, model, opt, loader_from_mels(mels), torch.device('cpu'), 1)
    out['train_stdout'] = np.array(buf.getvalue())
    for name, prm in model.named_parameters():
        put(out, 'after/' + name, prm)
    for name, b in model.named_buffers():
        out['bufN/' + name] = b.detach().numpy().copy()
    # ---- eval loop on the trained model (test.py:73-106)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(io.StringIO()):
        ref_test.test_vqvae(args, model, loader_from_mels(mels[:1]), torch.device('cpu'), 1)
    out['test_stdout'] = np.array(buf.getvalue())
    np.savez_compressed(os.path.join(OUT, 'vqvae_step_%s.npz' % tag), **out)
    print('vqvae_step_%s.npz' % tag, 'losses', out['losses'], buf.getvalue().strip())


class _ImageLoader(list):
    """In-memory stand-in for the torchvision DataLoader of main.py:84-99: yields (data, label), exposes ``.dataset``."""
    def __init__(self, items):
        super().__init__(items)
        self.dataset = list(range(sum(int(i[0].shape[0]) for i in items)))


def gen_step_images(ref_models, ref_train, tag, input_dim, hw, seed, batch=4, steps=2):
    """The torchvision branch (main.py:100-116 picks input_dim 1 or 3; train.py:150-179 is the loop): one fp64
    forward/backward of the reference modules on an image batch + ``steps`` verbatim train_vqvae iterations."""
    import torch.nn.functional as F
    out = {}
    dim, k = 64, 512
    params, buffers = R.make_state(input_dim, dim, k, seed, codebook='spread')
    imgs = [R.make_images(batch, input_dim, hw, hw, seed + 100 + i) for i in range(steps)]
    out['meta'] = np.array([input_dim, dim, k, batch, hw, seed, steps], dtype=np.int64)
    model = ref_models.VQVAE(input_dim, dim, k)
    load_into(model, params, buffers)
    model = model.double().train()
    data = imgs[0].double()
    x_tilde, z_e_x, z_q_x = model(data)
    with torch.no_grad():
        latents = ref_models.vq(z_e_x.permute(0, 2, 3, 1).contiguous(), model.codebook.embedding.weight)
    lr_ = F.mse_loss(x_tilde, data)                                  # train.py:160
    lv = F.mse_loss(z_q_x, z_e_x.detach())
    lc = F.mse_loss(z_e_x, z_q_x.detach())
    loss = lr_ + lv + 1.0 * lc
    model.zero_grad()
    loss.backward()
    out['x_tilde'] = x_tilde.detach().float().numpy()
    out['z_e_x'] = z_e_x.detach().float().numpy()
    out['indices'] = latents.numpy().astype(np.int16)
    out['losses'] = np.array([lr_.item(), lv.item(), lc.item(), loss.item()], dtype=np.float64)
    for name, prm in model.named_parameters():
        put(out, 'grad/' + name, prm.grad)
    # ---- verbatim torchvision-branch loop from a fresh copy of the same state
    model = ref_models.VQVAE(input_dim, dim, k)
    load_into(model, params, buffers)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)              # main.py:124
    args = argparse.Namespace(dataset='CIFAR10' if input_dim == 3 else 'MNIST', beta=1.0, log_interval=1)
    buf = io.StringIO()
    loader = _ImageLoader([(im, torch.zeros(batch, dtype=torch.long)) for im in imgs])
    with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(io.StringIO()):
        ref_train.train_vqvae(args, model, opt, loader, torch.device('cpu'), 1)
    out['train_stdout'] = np.array(buf.getvalue())
    for name, prm in model.named_parameters():
        put(out, 'after/' + name, prm)
    np.savez_compressed(os.path.join(OUT, 'vqvae_step_%s.npz' % tag), **out)
    print('vqvae_step_%s.npz' % tag, 'losses', out['losses'], buf.getvalue().strip().splitlines()[-1])


def gen_eval(ref_models):
    """Config C5-shaped (small): eval-mode encode -> decode with non-trivial running stats."""
    out = {}
    dim, k, seed, frames = 64, 512, 31, 37
    params, buffers = R.make_state(1, dim, k, seed, codebook='spread')
    rng = np.random.RandomState(seed + 1)
    for n in R.bn_names():
        buffers[n + '.running_mean'] = torch.tensor(0.1 * rng.standard_normal(dim).astype(np.float32))
        buffers[n + '.running_var'] = torch.tensor(rng.uniform(0.5, 1.5, dim).astype(np.float32))
    model = ref_models.VQVAE(1, dim, k)
    load_into(model, params, buffers)
    model.eval()
    x = R.make_mel(1, 80, frames, seed + 2).unsqueeze(1)
    with torch.no_grad():
        lat = model.encode(x)
        xt = model.decode(lat)
        xt2, z_e, z_q = model(x)
    print("eval encode-decode vs forward max diff", (xt - xt2).abs().max().item())
    assert torch.allclose(xt, xt2, atol=1e-5)
    out['meta'] = np.array([1, dim, k, 1, frames, seed], dtype=np.int64)
    out['latents'] = lat.numpy().astype(np.int16)
    out['x_tilde'] = xt.numpy()
    out['z_e_x'] = z_e.numpy()
    np.savez_compressed(os.path.join(OUT, 'vqvae_eval.npz'), **out)
    print('vqvae_eval.npz', lat.shape, xt.shape)


def gen_adam():
    """Known-answer vectors for torch.optim.Adam as the reference constructs it (main.py:124)."""
    out = {}
    rng = np.random.RandomState(41)
    shapes = [(64,), (64, 16), (5,)]
    ps = [torch.nn.Parameter(torch.tensor(rng.standard_normal(s).astype(np.float32))) for s in shapes]
    opt = torch.optim.Adam(ps, lr=1e-3)
    for step in range(3):
        for i, prm in enumerate(ps):
            scale = [1.0, 1e-4, 1e-8][i]          # includes the eps-dominated regime
            prm.grad = torch.tensor((scale * rng.standard_normal(prm.shape)).astype(np.float32))
        opt.step()
    for i, prm in enumerate(ps):
        out['p%d' % i] = prm.detach().numpy().copy()
    np.savez_compressed(os.path.join(OUT, 'adam_kat.npz'), **out)
    print('adam_kat.npz')


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--only', default=None, help="'images': just the torchvision-branch fixtures")
    only = ap.parse_args().only
    torch.manual_seed(1)
    torch.set_num_threads(1)          # fixed summation order for the fixtures
    os.makedirs(OUT, exist_ok=True)
    ref_vq, ref_models, ref_train, ref_test = import_reference()
    gen_step_images(ref_models, ref_train, 'img3_32', 3, 32, 71)     # CIFAR10-shaped (main.py:109-116: input_dim = 3)
    gen_step_images(ref_models, ref_train, 'img1_28', 1, 28, 72)     # MNIST / FashionMNIST-shaped (28 x 28, one channel)
    if only == 'images':
        return
    gen_vq(ref_vq)
    # seeds chosen (search over 200 seeds each) for the largest ReLU margin of the first batch, see
    # oracle.vqvae_ref.relu_margins: >= 1.8e-5 relative, i.e. ~20x fp32 rounding of a pre-activation
    gen_step(ref_models, ref_train, ref_test, 'spread_t32', 'spread', 32, 163)
    gen_step(ref_models, ref_train, ref_test, 'init_t30', 'init', 30, 38)
    gen_eval(ref_models)
    gen_adam()


if __name__ == '__main__':
    main()
