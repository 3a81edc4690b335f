"""clock64 timeline of CTA (0, 0) of wgrad_tma_kernel at the config-2 layer sizes (development aid)."""
import ctypes as C, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neural_sound_generation_b200 import ops, _cabi
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_ops import geoms, BF, DEV
NAMES = (['entry', 'setup done'] + ['prod t%d' % i for i in range(10)] + ['prod done'] +
         ['land t%d' % i for i in range(10)] + ['mma issued', 'acc ready', 'epi done', 'exit'])
lib = _cabi.load()
for name, H, W, k, s, p, tr in [('3x3', 20, 64, 3, 1, 1, False), ('1x1', 20, 64, 1, 1, 0, False), ('4x4s2', 40, 128, 4, 2, 1, False)]:
    fwd, dgr, wgr, OH, OW = geoms(32, H, W, k, s, p, tr)
    x = torch.randn(32, H, W, 64, device=DEV).to(BF)
    gy = torch.randn(32, OH, OW, 64, device=DEV).to(BF)
    dwp = torch.zeros(k * k, 64, 64, device=DEV)
    col = torch.zeros(64, device=DEV)
    for _ in range(3):
        ops.conv_wgrad_tma(wgr, x, gy, dwp, q_colsum=col)
    torch.cuda.synchronize()
    lib.nsg_debug_timeline(1, None)
    ops.conv_wgrad_tma(wgr, x, gy, dwp, q_colsum=col)
    buf = (C.c_ulonglong * 32)()
    lib.nsg_debug_timeline(0, buf)
    t0 = buf[0]
    print(name, ' '.join('%s=%d' % (NAMES[i], buf[i] - t0) for i in range(27) if buf[i]))
