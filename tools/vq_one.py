"""One vq_forward launch at N = 2^20, K = 512, D = 64 (bf16) for ncu captures."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neural_sound_generation_b200 import ops
n, k, d = int(os.environ.get('VQ_N', 1 << 20)), 512, 64
z = torch.randn(n, d, device='cuda').to(torch.bfloat16)
cb = torch.randn(k, d, device='cuda')
idx = torch.empty(n, dtype=torch.int64, device='cuda')
codes = torch.empty_like(z)
sq = torch.zeros(1, dtype=torch.float64, device='cuda')
for _ in range(3):
    ops.vq_forward(z, cb, idx, codes, sq, mode=1)
torch.cuda.synchronize()
