"""torch-CPU functional restatement of the reference VQ-VAE step (test oracle only).

Restates /root/reference/src/models.py (VQEmbedding :121-142, ResBlock :145-158,
VQVAE :161-216), the inline loss terms of /root/reference/src/train.py:116-136
and torch.optim.Adam as constructed at /root/reference/src/main.py:124.  It is
written against a flat ``{state_dict key: tensor}`` dict -- no nn.Module -- so
that it shares nothing with the product modules it checks.  Floating point:
fp32 by default, fp64 when the params passed in are fp64.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import vq_ref

BN_EPS = 1e-5        # nn.BatchNorm2d default (models.py:151 etc. pass only num_features)
BN_MOMENTUM = 0.1


# ----------------------------------------------------------------------------
# deterministic, torch-RNG-independent state (shared by gen_golden.py and tests)
# ----------------------------------------------------------------------------
def param_shapes(input_dim, dim, z_dim):
    """state_dict keys/shapes of reference ``VQVAE(input_dim, dim, z_dim)`` (models.py:162-184)."""
    s = {}

    def conv(name, co, ci, k):
        s[name + '.weight'] = (co, ci, k, k)
        s[name + '.bias'] = (co,)

    def bn(name):
        s[name + '.weight'] = (dim,)
        s[name + '.bias'] = (dim,)

    def res(name):
        conv(name + '.block.1', dim, dim, 3)
        bn(name + '.block.2')
        conv(name + '.block.4', dim, dim, 1)
        bn(name + '.block.5')

    conv('encoder.0', dim, input_dim, 4)
    bn('encoder.1')
    conv('encoder.3', dim, dim, 4)
    res('encoder.4')
    res('encoder.5')
    s['codebook.embedding.weight'] = (z_dim, dim)
    res('decoder.0')
    res('decoder.1')
    # ConvTranspose2d weight is (Cin, Cout, KH, KW)
    s['decoder.3.weight'] = (dim, dim, 4, 4)
    s['decoder.3.bias'] = (dim,)
    bn('decoder.4')
    s['decoder.6.weight'] = (dim, input_dim, 4, 4)
    s['decoder.6.bias'] = (input_dim,)
    return s


def bn_names(dim=None):
    names = ['encoder.1']
    for blk in ('encoder.4', 'encoder.5', 'decoder.0', 'decoder.1'):
        names += [blk + '.block.2', blk + '.block.5']
    names.append('decoder.4')
    return names


def make_state(input_dim, dim, z_dim, seed, codebook='init', dtype=torch.float32):
    """Deterministic parameters from numpy's MT19937 (stable across machines / torch versions).

    Conv weights: Xavier-uniform bound as in ``weights_init`` (models.py:25-32);
    biases / BN affine get small non-trivial values so every gradient path is
    exercised (the reference initialises them to 0 / (1, 0)).  ``codebook``:
    'init' = U(-1/K, 1/K) (models.py:125, adversarial near-tie case) or
    'spread' = N(0, 1) (trained-like).
    """
    rng = np.random.RandomState(seed)
    params = {}
    for name, shape in param_shapes(input_dim, dim, z_dim).items():
        if name == 'codebook.embedding.weight':
            if codebook == 'init':
                v = rng.uniform(-1.0 / z_dim, 1.0 / z_dim, size=shape)
            else:
                v = rng.standard_normal(size=shape)
        elif len(shape) == 4:
            rf = shape[2] * shape[3]
            bound = math.sqrt(6.0 / (shape[1] * rf + shape[0] * rf))
            v = rng.uniform(-bound, bound, size=shape)
        elif name.endswith('.weight'):      # BN gamma
            v = 1.0 + 0.1 * rng.standard_normal(size=shape)
        else:                               # conv bias / BN beta
            v = 0.05 * rng.standard_normal(size=shape)
        params[name] = torch.tensor(v.astype(np.float32)).to(dtype)
    buffers = {}
    for n in bn_names():
        buffers[n + '.running_mean'] = torch.zeros(dim, dtype=dtype)
        buffers[n + '.running_var'] = torch.ones(dim, dtype=dtype)
        buffers[n + '.num_batches_tracked'] = torch.zeros((), dtype=torch.int64)
    return params, buffers


def make_mel(batch, n_mels, frames, seed, dtype=torch.float32):
    """Synthetic mel batch ``c`` (B, n_mels, T) uniform in [0, 1]."""
    rng = np.random.RandomState(seed)
    return torch.tensor(rng.uniform(0.0, 1.0, size=(batch, n_mels, frames)).astype(np.float32)).to(dtype)


def make_images(batch, channels, height, width, seed, dtype=torch.float32):
    """Deterministic image batch in [0, 1] (what transforms.ToTensor() yields, dataloader.py:45-49): (B, C, H, W)."""
    rng = np.random.RandomState(seed)
    return torch.tensor(rng.uniform(0.0, 1.0, size=(batch, channels, height, width)).astype(np.float32)).to(dtype)


# ----------------------------------------------------------------------------
# layers
# ----------------------------------------------------------------------------
_RELU_MARGINS = None     # when a list: every _relu call appends min|x| / rms(x) of its input
_FORCED_MASKS = None     # when an iterator: every _relu call multiplies by the next mask instead of thresholding
_RECORD_MASKS = None     # when a list: every _relu call appends its own (x > 0) decision


def _relu(x):
    if _FORCED_MASKS is not None:
        # teacher-forced activation pattern (step_grads(relu_masks=...)): relu(x) -> x * mask, d/dx = mask.  Removes
        # the ReLU discontinuity from a comparison with an implementation whose pre-activations differ by rounding.
        m = next(_FORCED_MASKS)
        assert m.shape == x.shape, (tuple(m.shape), tuple(x.shape))
        return x * m.to(x.dtype)
    if _RECORD_MASKS is not None:
        _RECORD_MASKS.append((x.detach() > 0))
    if _RELU_MARGINS is not None:
        with torch.no_grad():
            xd = x.detach().double()
            _RELU_MARGINS.append((xd.abs().min() / xd.pow(2).mean().sqrt().clamp_min(1e-300)).item())
    return F.relu(x)


def relu_margins(p, b, c):
    """Smallest |pre-activation| / rms over every ReLU input of one train-mode forward, evaluated in fp64.

    ReLU makes the gradient a discontinuous function of the activations: a pre-activation within fp32
    rounding (~1e-6 relative) of zero lets two correct fp32 implementations differ by a discrete
    O(1/pixels) step in every upstream gradient.  gen_golden.py picks seeds whose margin is >= 1e-5 so
    that the fixtures pin gradients tightly."""
    global _RELU_MARGINS
    _RELU_MARGINS = []
    try:
        with torch.no_grad():
            forward({k: v.double() for k, v in p.items()},
                    {k: (v.double() if v.is_floating_point() else v) for k, v in b.items()},
                    c.double().unsqueeze(1), True, None)
        return min(_RELU_MARGINS)
    finally:
        _RELU_MARGINS = None
def _bn(x, p, b, name, training, new_buffers):
    """nn.BatchNorm2d forward; train mode uses biased batch variance for the output and
    updates running stats with the unbiased one (torch semantics)."""
    w, bias = p[name + '.weight'], p[name + '.bias']
    if x.dtype == torch.bfloat16:
        x = x.float()        # under torch.autocast nn.BatchNorm2d computes in fp32 (autocast's fp32 op list)
    if training:
        mean = x.mean(dim=(0, 2, 3))
        var = x.var(dim=(0, 2, 3), unbiased=False)
        if new_buffers is not None:
            n = x.numel() // x.shape[1]
            with torch.no_grad():
                rm = b[name + '.running_mean']
                rv = b[name + '.running_var']
                new_buffers[name + '.running_mean'] = (1 - BN_MOMENTUM) * rm + BN_MOMENTUM * mean
                new_buffers[name + '.running_var'] = ((1 - BN_MOMENTUM) * rv
                                                      + BN_MOMENTUM * var * (n / max(n - 1, 1)))
                new_buffers[name + '.num_batches_tracked'] = b[name + '.num_batches_tracked'] + 1
    else:
        mean, var = b[name + '.running_mean'], b[name + '.running_var']
    inv = torch.rsqrt(var + BN_EPS)
    return (x - mean[None, :, None, None]) * (inv * w)[None, :, None, None] + bias[None, :, None, None]


def _resblock(x, p, b, name, training, nb):
    """ResBlock (models.py:145-158).  The leading ReLU is in-place, so the skip
    connection adds relu(x), not x."""
    r = _relu(x)                                                             # :149 (inplace)
    h = F.conv2d(r, p[name + '.block.1.weight'], p[name + '.block.1.bias'], 1, 1)   # :150
    h = _relu(_bn(h, p, b, name + '.block.2', training, nb))                 # :151-152
    h = F.conv2d(h, p[name + '.block.4.weight'], p[name + '.block.4.bias'])  # :153
    h = _bn(h, p, b, name + '.block.5', training, nb)                        # :154
    return r + h                                                             # :158


def encoder(x, p, b, training=True, nb=None):
    """VQVAE.encoder (models.py:164-171)."""
    h = F.conv2d(x, p['encoder.0.weight'], p['encoder.0.bias'], 2, 1)
    h = _relu(_bn(h, p, b, 'encoder.1', training, nb))
    h = F.conv2d(h, p['encoder.3.weight'], p['encoder.3.bias'], 2, 1)
    h = _resblock(h, p, b, 'encoder.4', training, nb)
    return _resblock(h, p, b, 'encoder.5', training, nb)


def decoder(z, p, b, training=True, nb=None):
    """VQVAE.decoder (models.py:175-184)."""
    h = _resblock(z, p, b, 'decoder.0', training, nb)
    h = _resblock(h, p, b, 'decoder.1', training, nb)
    h = _relu(h)
    h = F.conv_transpose2d(h, p['decoder.3.weight'], p['decoder.3.bias'], 2, 1)
    h = _relu(_bn(h, p, b, 'decoder.4', training, nb))
    h = F.conv_transpose2d(h, p['decoder.6.weight'], p['decoder.6.bias'], 2, 1)
    return torch.tanh(h)


def quantize(z_e_x, codebook):
    """VQEmbedding.forward (models.py:127-130): (B,D,H,W) -> int64 (B,H,W)."""
    z = z_e_x.detach().permute(0, 2, 3, 1).contiguous()
    if z.dtype == torch.float64:
        d = vq_ref.vq_distances64(z.numpy(), codebook.detach().numpy())
        idx = np.argmin(d, axis=1).reshape(z.shape[:-1])
    else:
        idx = vq_ref.vq_indices(z.numpy(), codebook.detach().float().numpy())
    return torch.from_numpy(idx)


def forward(p, b, x, training=True, nb=None, indices=None):
    """VQVAE.forward (models.py:198-216) -> (x_tilde, z_e_x, z_q_x_bar, indices).

    straight_through (models.py:132-142): the decoder input carries the codebook
    values forward and the identity gradient back to z_e_x (vq_st backward,
    vector_quantization.py:50-52); z_q_x_bar is the same gather with gradient to
    the codebook.  ``indices`` may be teacher-forced.
    """
    z_e_x = encoder(x, p, b, training, nb)
    cb = p['codebook.embedding.weight']
    if indices is None:
        indices = quantize(z_e_x, cb)
    flat = indices.reshape(-1)
    shape = (z_e_x.shape[0], z_e_x.shape[2], z_e_x.shape[3], z_e_x.shape[1])
    z_q = cb.detach().index_select(0, flat).view(shape).permute(0, 3, 1, 2)
    z_q_st = z_q.detach() + (z_e_x - z_e_x.detach())  # exact codebook values forward, grad -> z_e_x
    z_q_bar = cb.index_select(0, flat).view(shape).permute(0, 3, 1, 2)
    x_tilde = decoder(z_q_st, p, b, training, nb)
    return x_tilde, z_e_x, z_q_bar, indices


def encode(p, b, x):
    """VQVAE.encode (models.py:188-191), eval-mode BN."""
    return quantize(encoder(x, p, b, training=False), p['codebook.embedding.weight'])


def decode(p, b, latents):
    """VQVAE.decode (models.py:193-196), eval-mode BN."""
    z = p['codebook.embedding.weight'][latents].permute(0, 3, 1, 2)
    return decoder(z, p, b, training=False)


def losses(x_tilde, z_e_x, z_q_x, c4, beta):
    """train.py:118-134.  x_tilde is zero-padded on the right to the input width."""
    pad = c4.shape[3] - x_tilde.shape[3]
    target = F.pad(x_tilde, (0, pad)) if pad > 0 else x_tilde            # :118-120
    loss_recons = F.mse_loss(target, c4)                                 # :129
    loss_vq = F.mse_loss(z_q_x, z_e_x.detach())                          # :131
    loss_commit = F.mse_loss(z_e_x, z_q_x.detach())                      # :133
    return loss_recons, loss_vq, loss_commit, loss_recons + loss_vq + beta * loss_commit  # :134


RELU_ORDER = (['encoder.1'] + [blk + s for blk in ('encoder.4', 'encoder.5', 'decoder.0', 'decoder.1')
                                for s in ('.in', '.block.3')] + ['decoder.2', 'decoder.5'])
"""Names of the 11 ReLUs of one forward in call order (state_dict-style: 'encoder.4.in' = the ResBlock's in-place
leading ReLU, models.py:149; 'encoder.4.block.3' = the ReLU after its first BatchNorm).  ``relu_masks`` follows it."""


def step_grads(p, b, c, beta=1.0, training=True, indices=None, relu_masks=None, autocast=None):
    """One forward/backward of the train step on mel batch ``c`` (B, n_mels, T).

    ``indices``: teacher-forced codebook indices (B, H, W).  ``relu_masks``: dict RELU_ORDER name -> {0,1} mask (NCHW)
    that replaces each ReLU's own threshold decision (forward x*mask, backward mask).  ``autocast``: a dtype
    (torch.bfloat16) to run forward + losses under ``torch.autocast('cpu', dtype)`` -- what the reference modules do
    under autocast; it is the yardstick for a reduced-precision implementation.

    Returns dict(outputs, losses, grads, new_buffers).
    """
    global _FORCED_MASKS, _RECORD_MASKS
    import contextlib
    p = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    nb = dict(b)
    c4 = c.unsqueeze(1) if c.dim() == 3 else c                           # train.py:115 (mel) / :151 (image batches)
    _FORCED_MASKS = iter([relu_masks[n] for n in RELU_ORDER]) if relu_masks is not None else None
    _RECORD_MASKS = [] if relu_masks is None else None
    try:
        with (torch.autocast('cpu', dtype=autocast) if autocast is not None else contextlib.nullcontext()):
            x_tilde, z_e_x, z_q_x, idx = forward(p, b, c4, training, nb, indices)
            if autocast is not None:     # the reference's F.mse_loss under autocast runs in fp32
                x_tilde, z_e_x, z_q_x = x_tilde.float(), z_e_x.float(), z_q_x.float()
            lr, lv, lc, loss = losses(x_tilde, z_e_x, z_q_x, c4, beta)
    finally:
        _FORCED_MASKS = None
        own_masks, _RECORD_MASKS = _RECORD_MASKS, None
    loss.backward()
    grads = {k: (v.grad.detach() if v.grad is not None else torch.zeros_like(v)) for k, v in p.items()}
    return dict(x_tilde=x_tilde.detach(), z_e_x=z_e_x.detach(), z_q_x=z_q_x.detach(), indices=idx,
                loss_recons=lr.detach(), loss_vq=lv.detach(), loss_commit=lc.detach(), loss=loss.detach(),
                grads=grads, buffers=nb,
                relu_masks=dict(zip(RELU_ORDER, own_masks)) if own_masks is not None else relu_masks)


def adam_init(p):
    return {k: dict(step=0, m=torch.zeros_like(v), v=torch.zeros_like(v)) for k, v in p.items()}


def adam_step(p, grads, state, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
    """torch.optim.Adam defaults (main.py:124): no weight decay, no amsgrad.

    p <- p - lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
    """
    b1, b2 = betas
    out = {}
    for k, w in p.items():
        st = state[k]
        st['step'] += 1
        g = grads[k]
        st['m'] = b1 * st['m'] + (1 - b1) * g
        st['v'] = b2 * st['v'] + (1 - b2) * g * g
        bc1 = 1 - b1 ** st['step']
        bc2 = 1 - b2 ** st['step']
        denom = st['v'].sqrt() / math.sqrt(bc2) + eps
        out[k] = w - (lr / bc1) * st['m'] / denom
    return out


def train_steps(p, b, batches, beta=1.0, lr=1e-3):
    """``len(batches)`` iterations of the train_vqvae loop body (train.py:108-138)."""
    state = adam_init(p)
    log = []
    for c in batches:
        r = step_grads(p, b, c, beta)
        p = adam_step(p, r['grads'], state, lr)
        b = r['buffers']
        log.append(r)
    return p, b, log
