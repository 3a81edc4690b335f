"""clock64 timeline of CTA 0 of conv_tma_kernel at the config-2 layer sizes (development aid)."""
import ctypes as C, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neural_sound_generation_b200 import ops, _cabi
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_ops import geoms, BF, DEV
NAMES = ['entry', 'setup done', 'prod tile0', 'prod tile1', 'prod tile2', 'prod last', 'W resident', 'first box',
         'mma tile0', 'mma tile1', 'mma tile2', 'mma last', 'acc0 ready', 'acc1 ready', 'acc2 ready', 'accL ready',
         'epi0 done', 'epi1 done', 'epi2 done', 'epiL done', 'exit',
         't1g0 box', 't1g0 issued', 't1g0 commit', 't1g1 box', 't1g1 issued', 't1g1 commit', 't1g2 box', 't1g2 issued', 't1g2 commit']
lib = _cabi.load()
for name, H, W, k, s, p, tr in [('3x3', 20, 64, 3, 1, 1, False), ('1x1', 20, 64, 1, 1, 0, False), ('4x4s2', 40, 128, 4, 2, 1, False)]:
    fwd, dgr, wgr, OH, OW = geoms(32, H, W, k, s, p, tr)
    x = torch.randn(32, H, W, 64, device=DEV).to(BF)
    w = torch.randn(64, 64, k, k, device=DEV) * 0.1
    y = torch.empty(32, OH, OW, 64, dtype=BF, device=DEV)
    st = torch.zeros(128, dtype=torch.float64, device=DEV)
    pf = ops.conv_pack_tma(fwd, w)
    for variant in ('stats', 'plain'):
        kw = dict(stats_mode=ops.STATS_MOMENTS, stats=st) if variant == 'stats' else {}
        for _ in range(3):
            ops.conv_gemm_tma(fwd, x, pf, y, **kw)
        torch.cuda.synchronize()
        lib.nsg_debug_timeline(1, None)
        ops.conv_gemm_tma(fwd, x, pf, y, **kw)
        buf = (C.c_ulonglong * 32)()
        lib.nsg_debug_timeline(0, buf)
        t0 = buf[0]
        print(name, variant, ' '.join('%s=%d' % (NAMES[i], buf[i] - t0) for i in range(30) if buf[i]))
