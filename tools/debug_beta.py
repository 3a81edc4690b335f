"""GPU debugging aid: is the fused path's decoder gradient beta- or order-dependent?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
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


def run(beta):
    m1 = model()
    m1.flatten_parameters(); m1._flat['grads'].zero_()
    losses, outs = m1.loss_and_grads(mel.cuda(), beta=beta)
    torch.cuda.synchronize()
    return {k: v.clone() for k, v in m1.grad_dict().items()}, [o.clone() for o in outs], losses.clone()


refs = {}
for beta in (1.0, 0.25):
    refs[beta] = R.step_grads({k: v.double() for k, v in p.items()},
                              {k: (v.double() if v.is_floating_point() else v) for k, v in b.items()}, mel.double(), beta=beta)
key = 'decoder.0.block.1.weight'
d = (refs[1.0]['grads'][key] - refs[0.25]['grads'][key]).norm() / refs[1.0]['grads'][key].norm()
print('oracle: decoder grad beta dependence', d.item())
print('oracle idx equal', bool((refs[1.0]['indices'] == refs[0.25]['indices']).all()))
runs = []
for beta in (1.0, 0.25, 1.0, 0.25, 0.5):
    g, outs, losses = run(beta)
    runs.append((beta, g, outs))
    ref = refs[beta if beta in refs else 1.0]
    e = ((g[key].cpu().double() - ref['grads'][key]).norm() / ref['grads'][key].norm()).item()
    e0 = ((g['encoder.0.weight'].cpu().double() - ref['grads']['encoder.0.weight']).norm() / ref['grads']['encoder.0.weight'].norm()).item()
    nm = (outs[3].cpu() != ref['indices'].reshape(-1)).sum().item()
    print('beta %.2f  %s err %.2e  enc0 err %.2e  idx mismatches %d  losses %s' % (beta, key, e, e0, nm, losses.tolist()))
b0, g0, o0 = runs[0]
for beta, g, outs in runs[1:]:
    print('vs run0: beta %.2f  dgrad diff %.2e  xt diff %.2e  idx diff %d' % (
        beta, ((g[key] - g0[key]).norm() / g0[key].norm()).item(), ((outs[0] - o0[0]).norm() / o0[0].norm()).item(),
        (outs[3] != o0[3]).sum().item()))

print('--- poison test')
def poison(val):
    xs = [torch.full((n,), val, device='cuda') for n in (1 << 8, 1 << 10, 1 << 12, 1 << 14, 1 << 16, 1 << 18, 1 << 20, 1 << 22, 1 << 24) for _ in range(6)]
    torch.cuda.synchronize()
    del xs
for val in (0.0, float('nan'), 1e30, 0.0):
    poison(val)
    g, outs, losses = run(1.0)
    ref = refs[1.0]
    e = ((g[key].cpu().double() - ref['grads'][key]).norm() / ref['grads'][key].norm()).item()
    e0 = ((g['encoder.0.weight'].cpu().double() - ref['grads']['encoder.0.weight']).norm() / ref['grads']['encoder.0.weight'].norm()).item()
    e4 = ((g['decoder.4.bias'].cpu().double() - ref['grads']['decoder.4.bias']).norm() / ref['grads']['decoder.4.bias'].norm()).item()
    print('poison %r: %s err %.2e  enc0 err %.2e dec4.bias err %.2e' % (val, key, e, e0, e4))
