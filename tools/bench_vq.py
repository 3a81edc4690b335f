"""VQ bottleneck kernels alone (BASELINE.json configs[0], [3]): search TFLOP/s = 2*N*K*D / t against the measured
bf16 peak, scatter-add / gather / EMA GB/s against the measured HBM peak (algorithmic bytes of SURVEY.md section 8d).
CUDA events, 256 MiB L2 flush between launches ("cold") or 20 launches in one CUDA graph ("warm").

    python tools/bench_vq.py [--full]        # --full adds the K / N sweep of config 4
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import neural_sound_generation_b200 as nsg  # noqa: E402
from neural_sound_generation_b200 import ops  # noqa: E402
from bench_ops import timed  # noqa: E402

DEV = 'cuda'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return d['hbm_gbs'], d['bf16_tflops']
    return 6650.0, 1590.0


def main():
    full = '--full' in sys.argv
    hbm, tf = peaks()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    lib = nsg._cabi.load()
    torch.manual_seed(0)
    cases = [(2048, 512, 64), (40960, 512, 64), (1 << 20, 512, 64)]
    if full:
        cases += [(1 << 20, k, 64) for k in (1024, 2048, 4096, 8192)]
        cases += [(1 << 20, k, d) for d in (128, 256) for k in (512, 4096)]      # configs[3]: D = 64-256
    for n, k, d in cases:
        for dt in (torch.bfloat16, torch.float32):
            z = torch.randn(n, d, device=DEV).to(dt)
            cb = torch.randn(k, d, device=DEV)
            idx = torch.empty(n, dtype=torch.int64, device=DEV)
            codes = torch.empty_like(z)
            sq = torch.zeros(1, dtype=torch.float64, device=DEV)
            for mode in (1, 0):
                if not ops.vq_mode_serves(k, d, mode):
                    continue
                if mode == 0 and n * k > (1 << 30):
                    continue
                for fl, tag in ((flush, 'cold'), (None, 'warm')):
                    t = timed(lambda: ops.vq_forward(z, cb, idx, codes, sq, mode=mode), flush=fl, iters=10)
                    fl_ = 2.0 * n * k * d
                    print('vq_search  N=%-8d K=%-5d D=%-3d %-5s mode=%d %s %9.1f us  %7.1f TFLOP/s  %5.1f%% of %s bf16 peak'
                          % (n, k, d, str(dt).split('.')[-1], mode, tag, t, fl_ / t / 1e6, 100 * fl_ / t / 1e6 / tf, tf))
            if dt == torch.bfloat16 or n <= 40960:
                es = 2 if dt == torch.bfloat16 else 4
                table = torch.zeros(k, d, device=DEV)
                counts = torch.zeros(k, device=DEV)
                ops.vq_forward(z, cb, idx, codes, sq, mode=0 if n * k <= (1 << 30) else 1)
                skew = (torch.rand(n, device=DEV) ** 8 * min(k, 6)).long()       # an untrained encoder: a few codewords win
                for ix, dist_name in ((idx, 'uniform'), (skew, 'skewed')):
                    for sm, nm in ((0, 'aggregated'), (1, 'per-element')):
                        prev = lib.nsg_scatter_mode(sm)
                        for fl, tag in ((flush, 'cold'), (None, 'warm')):
                            t = timed(lambda: ops.scatter_add(codes, ix, table, alpha=0.5, sub=z, counts=counts), flush=fl, iters=10)
                            nbytes = 2 * n * d * es + n * 8 + 2 * (k * d + k) * 4
                            print('scatter_add N=%-8d K=%-5d D=%-3d %-5s %-11s %-7s %s %9.1f us  %7.1f GB/s  %5.1f%% of %s HBM'
                                  % (n, k, d, str(dt).split('.')[-1], nm, dist_name, tag, t, nbytes / t / 1e3,
                                     100 * nbytes / t / 1e3 / hbm, hbm))
                        lib.nsg_scatter_mode(prev)
                for fl, tag in ((flush, 'cold'), (None, 'warm')):
                    t = timed(lambda: ops.gather_rows(cb, idx, codes), flush=fl, iters=10)
                    nbytes = n * d * es + n * 8 + k * d * 4
                    print('gather     N=%-8d K=%-5d D=%-3d %-5s             %s %9.1f us  %7.1f GB/s  %5.1f%% of %s HBM'
                          % (n, k, d, str(dt).split('.')[-1], tag, t, nbytes / t / 1e3, 100 * nbytes / t / 1e3 / hbm, hbm))
    k, d = 8192, 64
    en, em, cbk = torch.ones(k, device=DEV), torch.randn(k, d, device=DEV), torch.randn(k, d, device=DEV)
    cn, sm_ = torch.ones(k, device=DEV), torch.randn(k, d, device=DEV)
    for fl, tag in ((flush, 'cold'), (None, 'warm')):
        t = timed(lambda: ops.ema_update(cn, sm_, en, em, cbk), flush=fl, iters=10)
        nbytes = 4 * k * d * 4
        print('ema_update K=%d D=%d %s %9.1f us  %7.1f GB/s' % (k, d, tag, t, nbytes / t / 1e3))


if __name__ == '__main__':
    main()
