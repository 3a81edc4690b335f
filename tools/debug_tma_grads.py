"""Per-parameter gradient error of the bf16 engines against the teacher-forced fp32 oracle."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import neural_sound_generation_b200 as nsg
from oracle import vqvae_ref as R
from helpers import relerr

def run(impl, seed=5, batch=4, frames=64):
    p, b = R.make_state(1, 64, 512, seed, codebook='spread')
    mel = R.make_mel(batch, 80, frames, seed + 1)
    m = nsg.VQVAE(1, 64, 512)
    sd = dict(p); sd.update(b)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m = m.to('cuda')
    m.act_dtype, m.conv_impl, m.vq_mode = torch.bfloat16, impl, 1
    m.train()
    nsg.optim.Adam(m, lr=1e-3).zero_grad()
    losses, (xt, ze, zq, idx) = m.loss_and_grads(mel.to('cuda'), beta=1.0)
    torch.cuda.synchronize()
    free = R.step_grads(p, b, mel, beta=1.0)
    r = R.step_grads(p, b, mel, beta=1.0, indices=idx.cpu().view(free['indices'].shape))
    g = m.grad_dict()
    print('impl', impl, 'x_tilde', relerr(xt.float(), r['x_tilde']), 'z_e', relerr(ze.float(), r['z_e_x']))
    for name, ref in r['grads'].items():
        print('  %-32s %.3e   |ref| %.3e' % (name, relerr(g[name], ref), ref.norm().item()))

for impl in (1, 2):
    run(impl)
