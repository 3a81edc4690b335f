"""GPU debugging aid: per-parameter gradient error of the fused and autograd paths vs the fp64 oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import neural_sound_generation_b200 as nsg
from oracle import vqvae_ref as R

seed = 21
p, b = R.make_state(1, 64, 512, seed, codebook='spread')
mel = R.make_mel(2, 80, 32, seed + 100)


def model():
    m = nsg.VQVAE(1, 64, 512)
    sd = dict(p); sd.update(b)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    return m.to('cuda').train()


for beta in (1.0, 0.25):
    r = R.step_grads({k: v.double() for k, v in p.items()},
                     {k: (v.double() if v.is_floating_point() else v) for k, v in b.items()}, mel.double(), beta=beta)
    m1 = model()
    m1.flatten_parameters(); m1._flat['grads'].zero_()
    m1.loss_and_grads(mel.cuda(), beta=beta)
    g1 = {k: v.clone() for k, v in m1.grad_dict().items()}
    m2 = model()
    c = mel.cuda().unsqueeze(1)
    xt, ze, zq = m2(c)
    (F.mse_loss(xt, c) + F.mse_loss(zq, ze.detach()) + beta * F.mse_loss(ze, zq.detach())).backward()
    g2 = {n: q.grad for n, q in m2.named_parameters()}
    print('beta', beta)
    for n in p:
        ref = r['grads'][n]
        e1 = ((g1[n].cpu().double() - ref).norm() / ref.norm().clamp_min(1e-30)).item()
        e2 = ((g2[n].cpu().double() - ref).norm() / ref.norm().clamp_min(1e-30)).item()
        e12 = ((g1[n] - g2[n]).double().norm() / g1[n].double().norm().clamp_min(1e-30)).item()
        print('  %-32s fused %.2e  autograd %.2e  fused-vs-autograd %.2e' % (n, e1, e2, e12))
