"""CPU tests: the oracle restatement of the reference's VAE / GatedPixelCNN (oracle/n4_ref.py) against
fixtures produced by the UNMODIFIED reference (oracle/gen_golden_n4.py -> tests/golden/n4_*.npz)."""
import os

import numpy as np
import pytest
import torch

from oracle import n4_ref as R
from helpers import relerr


PRE_BN_BIAS = ['encoder.%d.bias' % i for i in (0, 3, 6, 9)] + ['decoder.%d.bias' % i for i in (0, 3, 6)]


def _load(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name))
    state = {k[6:]: torch.tensor(g[k]) for k in g.files if k.startswith('state/')}
    return g, state


@pytest.mark.parametrize('tag', ['mel', 'img'])
def test_vae_oracle_vs_reference_fixture(golden_dir, tag):
    g, p = _load(golden_dir, 'n4_vae_%s.npz' % tag)
    x, eps = torch.tensor(g['x']), torch.tensor(g['eps'])
    r = R.vae_step_grads(p, x, eps)
    assert tuple(r['x_tilde'].shape) == g['x_tilde'].shape
    assert relerr(r['x_tilde'], g['x_tilde']) < 1e-5
    assert abs(r['kl'].item() - float(g['kl'])) < 1e-5 * abs(float(g['kl']))
    assert abs(r['loss'].item() - float(g['loss'])) < 1e-5 * abs(float(g['loss']))
    for k, v in r['grads'].items():
        ref = g['grad/' + k]
        if k in PRE_BN_BIAS:
            continue                  # bias before a BatchNorm: analytically zero, rounding noise on both sides
        assert relerr(v, ref) < 2e-4, (k, relerr(v, ref))
    # batch statistics -> running statistics as torch updates them (momentum 0.1, unbiased variance)
    for name, (m, v) in r['aux']['bn'].items():
        assert relerr(0.1 * m.detach(), g['after/%s.running_mean' % name]) < 1e-4, name


@pytest.mark.parametrize('tag', ['k32', 'latents'])
def test_pixelcnn_oracle_vs_reference_fixture(golden_dir, tag):
    g, p = _load(golden_dir, 'n4_pixelcnn_%s.npz' % tag)
    n_layers = int(g['meta'][2])
    logits = R.pixelcnn_forward(p, torch.tensor(g['x']), torch.tensor(g['label']), n_layers)
    assert relerr(logits, g['logits']) < 1e-5
    # make_causal masked the first layer's kernels in place, as the reference does
    for k in ('layers.0.vert_stack.weight', 'layers.0.horiz_stack.weight'):
        assert np.array_equal(p[k].numpy(), g['after/' + k])
    # causality: position (i, j)'s logits do not depend on x at or after (i, j)
    x2 = torch.tensor(g['x']).clone()
    i, j = 3, 4
    x2[:, i, j:] = (x2[:, i, j:] + 1) % int(g['meta'][0])
    x2[:, i + 1:, :] = (x2[:, i + 1:, :] + 1) % int(g['meta'][0])
    l2 = R.pixelcnn_forward(p, x2, torch.tensor(g['label']), n_layers)
    assert torch.allclose(l2[:, :, i, j], logits[:, :, i, j], atol=1e-6)
    assert torch.allclose(l2[:, :, :i, :], logits[:, :, :i, :], atol=1e-6)


def test_inverse_cdf_sampling():
    logits = torch.tensor([[0.0, 0.0, 0.0, 0.0], [10.0, -10.0, -10.0, -10.0], [-5.0, 0.0, 5.0, 0.0]])
    assert R.sample_inverse_cdf(logits, torch.tensor([0.0, 0.999, 0.5])).tolist() == [0, 0, 2]
    assert R.sample_inverse_cdf(logits[:1].repeat(4, 1), torch.tensor([0.1, 0.3, 0.6, 0.99])).tolist() == [0, 1, 2, 3]
    u = torch.rand(20000, generator=torch.Generator().manual_seed(0))
    draws = R.sample_inverse_cdf(logits[2:].repeat(20000, 1), u)
    freq = torch.bincount(draws, minlength=4).double() / 20000
    assert (freq - torch.softmax(logits[2].double(), -1)).abs().max() < 0.01
