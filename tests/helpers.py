import numpy as np
import torch

PRE_BN_BIAS = (['encoder.0.bias', 'decoder.3.bias']
               + ['%s.block.%d.bias' % (blk, i) for blk in ('encoder.4', 'encoder.5', 'decoder.0', 'decoder.1')
                  for i in (1, 4)])


def relerr(a, ref):
    a = torch.as_tensor(a).detach().double().cpu()
    ref = torch.as_tensor(ref).detach().double().cpu()
    return ((a - ref).norm() / ref.norm().clamp_min(1e-30)).item()


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


def nchw(t):
    return t.permute(0, 3, 1, 2).contiguous()


def check_summary(g, prefix, t, rtol, atol=0.0):
    """compare against a fixture entry written by oracle/gen_golden.py::summarize"""
    a = t.detach().cpu().double().numpy()
    if prefix + '/full' in g:
        ref = g[prefix + '/full'].astype(np.float64)
        err = np.linalg.norm(a - ref) / max(np.linalg.norm(ref), 1e-30)
        assert err <= rtol or np.abs(a - ref).max() <= atol, (prefix, err)
    else:
        norm = float(g[prefix + '/norm'])
        assert abs(np.sqrt((a ** 2).sum()) - norm) <= rtol * norm, prefix
        stride = max(1, a.size // 32)
        s = a.reshape(-1)[::stride][:32]
        ref = g[prefix + '/samples'].astype(np.float64)
        assert np.abs(s - ref).max() <= rtol * max(np.abs(ref).max(), norm / np.sqrt(a.size)) + atol, prefix
