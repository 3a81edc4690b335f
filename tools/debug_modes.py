"""GPU debugging aid: end-to-end error of each execution mode vs the fp64 oracle on the spread_t32 fixture inputs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neural_sound_generation_b200 as nsg
from oracle import vqvae_ref as R

seed = 163
p, b = R.make_state(1, 64, 512, seed, codebook='spread')
mel = R.make_mel(2, 80, 32, seed + 100)
r = R.step_grads({k: v.double() for k, v in p.items()},
                 {k: (v.double() if v.is_floating_point() else v) for k, v in b.items()}, mel.double(), beta=1.0)
PRE = [n for n in p if n.endswith('.bias') and ('block.1' in n or 'block.4' in n or n in ('encoder.0.bias', 'decoder.3.bias'))]
for mode, (dt, ci, vm) in dict(fp32=(torch.float32, 0, 0), tc32=(torch.float32, 1, 1), bf16=(torch.bfloat16, 1, 1)).items():
    m = nsg.VQVAE(1, 64, 512)
    sd = dict(p); sd.update(b)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m = m.to('cuda').train()
    m.act_dtype, m.conv_impl, m.vq_mode = dt, ci, vm
    m.flatten_parameters(); m._flat['grads'].zero_()
    losses, (xt, ze, zq, idx) = m.loss_and_grads(mel.cuda(), beta=1.0)
    torch.cuda.synchronize()
    g = m.grad_dict()
    rel = lambda a, ref: ((a.detach().cpu().double() - ref).norm() / ref.norm().clamp_min(1e-30)).item()
    nm = (idx.cpu() != r['indices'].reshape(-1)).sum().item()
    errs = {n: rel(g[n], r['grads'][n]) for n in p if n not in PRE}
    worst = max(errs, key=errs.get)
    print('%-5s idx mismatches %d/%d  z_e %.2e  x_tilde %.2e  losses %s  worst grad %s %.2e  median %.2e' % (
        mode, nm, idx.numel(), rel(ze.float(), r['z_e_x']), rel(xt.float(), r['x_tilde']),
        ['%.6f' % v for v in losses.tolist()], worst, errs[worst], sorted(errs.values())[len(errs) // 2]))
print('oracle losses', r['loss_recons'].item(), r['loss_vq'].item())
