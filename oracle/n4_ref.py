"""TEST INFRASTRUCTURE (oracle): CPU restatement of the reference's other model families.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this package.

Functional restatements over a flat ``{state_dict key: tensor}`` dict (torch CPU ops; fp32 or fp64 by the dtype of the
dict), each citing /root/reference/src/models.py:

* ``vae_forward``          :64-118  conv VAE: encoder (4 convolutions + train-mode BatchNorm), ``mu, logvar = chunk(2)``,
                                    ``kl = kl_divergence(N(mu, exp(logvar / 2)), N(0, 1)).sum(1).mean()``,
                                    ``z = mu + std * eps`` (``rsample``; the noise is an explicit argument), decoder.
* ``vae_loss``             loss.py:22-24  ``mse_loss(x_tilde, x, size_average=False) / x.size(0) + kl_d``; the zero-pad of
                                    train.py:58-60 (``target[..., :x_tilde.size(3)] = x_tilde``) is ``pad_to``.
* ``pixelcnn_forward``     :228-325 GatedPixelCNN: embedding, 'A' / 'B' gated masked convolutions (``make_causal`` zeroes
                                    the last kernel row / column of the first layer IN PLACE, :255-257), class-conditional
                                    bias, vertical -> horizontal link, residual, output head.
* ``sample_inverse_cdf``   :334-339 one categorical draw per row from softmax(logits) -- the reference draws with
                                    ``probs.multinomial(1)`` (torch's RNG stream: not reproducible elsewhere), here the
                                    uniform numbers are an argument and the draw is the inverse CDF: *sampling parity
                                    unpinned* by construction; ``generate`` is checked on its distribution and
                                    determinism instead.  (The reference's ``generate`` itself cannot run as written:
                                    ``torch.zeros((batch_size, shape))`` with a tuple ``shape`` raises, :328-331.)

Pinned against the reference: tests/golden/n4_*.npz hold outputs of the UNMODIFIED ``models.VAE`` / ``models.GatedPixelCNN``
/ ``loss.mse_loss`` (oracle/gen_golden_n4.py; the VAE's noise is injected by replacing
``torch.distributions.normal._standard_normal``, nothing in the reference is edited).
"""
import torch
import torch.nn.functional as F

BN_EPS = 1e-5


def _bn_train(h, p, name):
    """nn.BatchNorm2d in train mode (batch statistics, biased variance)."""
    mean = h.mean((0, 2, 3))
    var = h.var((0, 2, 3), unbiased=False)
    xh = (h - mean[None, :, None, None]) / torch.sqrt(var + BN_EPS)[None, :, None, None]
    return xh * p[name + '.weight'][None, :, None, None] + p[name + '.bias'][None, :, None, None], mean, var


def vae_forward(p, x, eps):
    """Returns (x_tilde, kl_div, aux) with aux = dict(mu=, logvar=, z=, bn={name: (mean, biased var)})."""
    bn = {}

    def cbr(h, conv, norm, stride, pad, relu=True, transposed=False):
        f = F.conv_transpose2d if transposed else F.conv2d
        h = f(h, p[conv + '.weight'], p[conv + '.bias'], stride, pad)
        h, m, v = _bn_train(h, p, norm)
        bn[norm] = (m, v)
        return F.relu(h) if relu else h

    h = cbr(x, 'encoder.0', 'encoder.1', 2, 1)                       # :67-69
    h = cbr(h, 'encoder.3', 'encoder.4', 2, 1)                       # :70-72
    h = cbr(h, 'encoder.6', 'encoder.7', 1, 0)                       # :73-75
    h = cbr(h, 'encoder.9', 'encoder.10', 1, 0, relu=False)          # :76-77
    mu, logvar = h.chunk(2, dim=1)                                   # :103
    std = (logvar * 0.5).exp()
    kl = (0.5 * (std * std + mu * mu - 1.0) - 0.5 * logvar).sum(1).mean()   # kl_divergence(N(mu, std), N(0, 1)), :107-109
    z = mu + std * eps                                               # rsample, :113
    h = cbr(z, 'decoder.0', 'decoder.1', 1, 0, transposed=True)      # :81-83
    h = cbr(h, 'decoder.3', 'decoder.4', 1, 0, transposed=True)
    h = cbr(h, 'decoder.6', 'decoder.7', 2, 1, transposed=True)
    xt = torch.tanh(F.conv_transpose2d(h, p['decoder.9.weight'], p['decoder.9.bias'], 2, 1))     # :90-91
    return xt, kl, dict(mu=mu, logvar=logvar, z=z, bn=bn)


def vae_loss(x_tilde, x, kl, pad_to=None):
    if pad_to is not None and x_tilde.shape[3] < pad_to:             # train.py:58-60
        x_tilde = F.pad(x_tilde, (0, pad_to - x_tilde.shape[3]))
    return ((x_tilde - x) ** 2).sum() / x.shape[0] + kl              # loss.py:22-24


def vae_step_grads(p, x, eps):
    q = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in p.items()}
    xt, kl, aux = vae_forward(q, x, eps)
    loss = vae_loss(xt, x, kl, pad_to=x.shape[3])
    loss.backward()
    return dict(x_tilde=xt.detach(), kl=kl.detach(), loss=loss.detach(), aux=aux,
                grads={k: v.grad for k, v in q.items() if v.requires_grad and v.grad is not None})


def pixelcnn_forward(p, x, label, n_layers, mask_in_place=True):
    """x (B, H, W) int64 in [0, input_dim), label (B,) int64 -> logits (B, input_dim, H, W).  ``p`` is modified as the
    reference modifies its parameters when ``mask_in_place`` (make_causal, :255-257)."""
    emb = p['embedding.weight'][x.reshape(-1)].reshape(x.shape + (-1,)).permute(0, 3, 1, 2)     # :311-313

    def gate(t):
        a, b = t.chunk(2, dim=1)
        return torch.tanh(a) * torch.sigmoid(b)                      # :222-225

    xv = xh = emb
    for i in range(n_layers):
        pre = 'layers.%d.' % i
        k = 7 if i == 0 else 3                                       # :295-297
        wv, wh = p[pre + 'vert_stack.weight'], p[pre + 'horiz_stack.weight']
        if i == 0:                                                   # mask 'A'
            if mask_in_place:
                wv[:, :, -1].zero_()
                wh[:, :, :, -1].zero_()
            else:
                wv, wh = wv.clone(), wh.clone()
                wv[:, :, -1] = 0
                wh[:, :, :, -1] = 0
        h = p[pre + 'class_cond_embedding.weight'][label]            # :262
        hv = F.conv2d(xv, wv, p[pre + 'vert_stack.bias'], 1, (k // 2, k // 2))[:, :, :xv.shape[-1], :]       # :263-264
        out_v = gate(hv + h[:, :, None, None])
        hh = F.conv2d(xh, wh, p[pre + 'horiz_stack.bias'], 1, (0, k // 2))[:, :, :, :xh.shape[-2]]           # :267-268
        v2h = F.conv2d(hv, p[pre + 'vert_to_horiz.weight'], p[pre + 'vert_to_horiz.bias'])
        out = gate(v2h + hh + h[:, :, None, None])
        res = F.conv2d(out, p[pre + 'horiz_resid.weight'], p[pre + 'horiz_resid.bias'])
        xv, xh = out_v, (res + xh if i > 0 else res)                 # :272-275 (no residual on the first layer)
    h = F.relu(F.conv2d(xh, p['output_conv.0.weight'], p['output_conv.0.bias']))
    return F.conv2d(h, p['output_conv.2.weight'], p['output_conv.2.bias'])                       # :303-307


def sample_inverse_cdf(logits, u):
    """logits (B, C), u (B,) in [0, 1) -> (B,) int64: first class whose cumulative softmax mass exceeds u."""
    pr = torch.softmax(logits.double(), -1)
    cdf = pr.cumsum(-1)
    idx = (cdf > u.double()[:, None] * cdf[:, -1:]).float().argmax(-1)
    none = ~(cdf > u.double()[:, None] * cdf[:, -1:]).any(-1)
    idx[none] = logits.shape[1] - 1
    return idx
