"""Generate tests/golden/n4_*.npz by running the UNMODIFIED reference models on CPU.

Run in the authoring container only (needs /root/reference):
    python -m oracle.gen_golden_n4

Imports /root/reference/src/{models,loss}.py through sys.path.  Nothing in the reference is edited: the one random
draw of ``VAE.forward`` (``q_z_x.rsample()``, models.py:113) is made reproducible by replacing
``torch.distributions.normal._standard_normal`` -- the function ``Normal.rsample`` calls for its noise -- with one that
returns the noise tensor stored in the fixture.  Fixtures hold the full state dicts (small models), the inputs, the
noise and the reference's outputs / gradients / updated BatchNorm buffers.
"""
import os
import sys
import warnings

import numpy as np
import torch

REF_SRC = '/root/reference/src'
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def import_reference():
    sys.path.insert(0, REF_SRC)
    import models as ref_models      # noqa: E402
    import loss as ref_loss          # noqa: E402
    return ref_models, ref_loss


def gen_vae(ref_models, ref_loss, tag, input_dim, dim, z_dim, shape, seed, pad_hack):
    torch.manual_seed(seed)
    model = ref_models.VAE(input_dim, dim, z_dim).train()
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.rand(*shape, generator=g)
    with torch.no_grad():
        zshape = model.encoder(x).chunk(2, dim=1)[0].shape
    model.load_state_dict(model.state_dict())        # (no-op; the probe above moved the running statistics)
    torch.manual_seed(seed)
    model = ref_models.VAE(input_dim, dim, z_dim).train()
    eps = torch.randn(*zshape, generator=g)
    out = {'state/' + k: v.detach().clone().numpy() for k, v in model.state_dict().items()}
    import torch.distributions.normal as tdn
    real = tdn._standard_normal
    tdn._standard_normal = lambda shp, dtype, device: eps.to(dtype).reshape(shp)
    try:
        x_tilde, kl_d = model(x)                                          # models.py:97-118
    finally:
        tdn._standard_normal = real
    if pad_hack:                                                          # train.py:58-60
        target = torch.zeros(x.size(0), x.size(1), x.size(2), x.size(3))
        target[:, :, :, :x_tilde.size(3)] = x_tilde
    else:
        target = x_tilde
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        loss = ref_loss.mse_loss(target, x, kl_d)                         # loss.py:22-24
    loss.backward()
    out.update({'x': x.numpy(), 'eps': eps.numpy(), 'x_tilde': x_tilde.detach().numpy(), 'kl': kl_d.detach().numpy(),
                'loss': loss.detach().numpy()})
    for k, v in model.named_parameters():
        out['grad/' + k] = v.grad.numpy()
    for k, v in model.state_dict().items():
        if 'running' in k or 'num_batches' in k:
            out['after/' + k] = v.detach().clone().numpy()
    out['meta'] = np.array([input_dim, dim, z_dim, seed, int(pad_hack)])
    np.savez_compressed(os.path.join(OUT, 'n4_vae_%s.npz' % tag), **out)
    print('n4_vae_%s: x %s -> x_tilde %s, z %s, kl %.5f loss %.5f' % (tag, tuple(x.shape), tuple(x_tilde.shape), tuple(zshape),
                                                                    kl_d.item(), loss.item()))


def gen_pixelcnn(ref_models, tag, input_dim, dim, n_layers, n_classes, shape, seed):
    torch.manual_seed(seed)
    model = ref_models.GatedPixelCNN(input_dim, dim, n_layers, n_classes).eval()
    out = {'state/' + k: v.detach().clone().numpy() for k, v in model.state_dict().items()}    # BEFORE make_causal
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randint(0, input_dim, shape, generator=g)
    label = torch.randint(0, n_classes, (shape[0],), generator=g)
    with torch.no_grad(), warnings.catch_warnings():
        warnings.simplefilter('ignore')
        logits = model(x, label)                                          # models.py:309-325
    out.update({'x': x.numpy(), 'label': label.numpy(), 'logits': logits.numpy()})
    for k in ('layers.0.vert_stack.weight', 'layers.0.horiz_stack.weight'):
        out['after/' + k] = model.state_dict()[k].detach().clone().numpy()                     # masked in place (:255-257)
    out['meta'] = np.array([input_dim, dim, n_layers, n_classes, seed])
    np.savez_compressed(os.path.join(OUT, 'n4_pixelcnn_%s.npz' % tag), **out)
    print('n4_pixelcnn_%s: x %s -> logits %s' % (tag, tuple(x.shape), tuple(logits.shape)))


def main():
    torch.set_num_threads(1)
    ref_models, ref_loss = import_reference()
    gen_vae(ref_models, ref_loss, 'mel', 1, 32, 8, (2, 1, 80, 66), 41, True)       # mel-shaped, ragged width (66 -> 64)
    gen_vae(ref_models, ref_loss, 'img', 3, 16, 4, (3, 3, 32, 32), 42, False)      # CIFAR10-shaped
    gen_pixelcnn(ref_models, 'k32', 32, 16, 4, 10, (2, 8, 8), 43)
    gen_pixelcnn(ref_models, 'latents', 128, 32, 3, 10, (1, 16, 16), 44)            # dim 32 over 128 codes; square: the
    # reference crops rows with size(-1) and columns with size(-2) (models.py:264,268) and fails on anything else


if __name__ == '__main__':
    main()
