"""Host-side data-parallel logic (neural_sound_generation_b200/dist.py) on CPU: world size 2 over gloo.

The kernels have no CPU path, so these tests exercise what is backend-agnostic: shard assignment, the gradient
buckets over the model's real flat layout, bucketed all-reduce == whole-buffer all-reduce == mean after 1/N scaling,
and the SyncBN statistic reduction (N shards' [sum | sum sq] == the global batch's moments; oracle BatchNorm on the
global batch == BatchNorm from the reduced statistics)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import neural_sound_generation_b200 as nsg
from neural_sound_generation_b200 import dist as nd
from oracle import vqvae_ref as R


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(fn, world=2):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn), nprocs=world, join=True)


def _entry(rank, world, port, fn):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        fn(rank, world)
    finally:
        dist.destroy_process_group()


def test_shard_range_and_buckets_over_real_layout():
    assert [nd.shard_range(r, 4, 32) for r in range(4)] == [(0, 8), (8, 16), (16, 24), (24, 32)]
    with pytest.raises(ValueError):
        nd.shard_range(0, 3, 32)
    m = nsg.VQVAE(1, 64, 512)
    f = m.flatten_parameters()
    sl = nd.bucket_slices(f['names'], f['offs'], f['total'])
    lo, hi = sl['decoder']
    assert sl['rest'] == (0, lo) and hi == f['total'] and 0 < lo < hi
    for n, o in zip(f['names'], f['offs']):
        assert (o >= lo) == n.startswith('decoder.'), n
    # every parameter's gradient view lies inside exactly one bucket
    g = m.grad_dict()
    base = f['grads'].data_ptr()
    for n, t in g.items():
        o = (t.data_ptr() - base) // 4
        assert (lo <= o and o + t.numel() <= hi) or o + t.numel() <= lo, n
    with pytest.raises(ValueError):
        nd.bucket_slices(['decoder.0.w', 'encoder.0.w'], [0, 64], 128)


def _w_buckets(rank, world):
    torch.manual_seed(100 + rank)
    m = nsg.VQVAE(1, 64, 512)
    f = m.flatten_parameters()
    sl = nd.bucket_slices(f['names'], f['offs'], f['total'])
    g_local = torch.randn(f['total'])
    whole = g_local.clone()
    dist.all_reduce(whole)
    bucketed = nd.allreduce_buckets_(g_local.clone(), sl)
    assert torch.equal(whole, bucketed)
    # mean of the rank gradients after Adam's grad_scale = 1/world
    gathered = [torch.empty_like(g_local) for _ in range(world)]
    dist.all_gather(gathered, g_local)
    assert torch.allclose(bucketed / world, torch.stack(gathered).mean(0), rtol=1e-6, atol=1e-7)


def test_bucketed_allreduce_equals_whole_buffer_mean():
    _run(_w_buckets)


class _StubEngine(object):
    def __init__(self):
        self.calls = []

    def finish_backward(self, ctx, G, part=None):
        self.calls.append(part)


def _w_dp_object(rank, world):
    dp = nd.DataParallel(sync_bn=True, overlap=False)
    assert (dp.world, dp.rank) == (world, rank)
    # SyncBN statistics: shards of one global batch
    g = torch.Generator().manual_seed(7)
    x = torch.randn(8, 64, 5, 6, generator=g, dtype=torch.float64)          # global batch, identical on both ranks
    lo, hi = nd.shard_range(rank, world, 8)
    xs = x[lo:hi]
    stats = torch.cat([xs.sum((0, 2, 3)), (xs * xs).sum((0, 2, 3))])
    dp.reduce_stats(stats)
    count = xs.numel() // 64 * world
    mean = stats[:64] / count
    var = stats[64:] / count - mean * mean
    assert torch.allclose(mean, x.mean((0, 2, 3)), rtol=1e-12, atol=1e-12)
    assert torch.allclose(var, x.var((0, 2, 3), unbiased=False), rtol=1e-10, atol=1e-12)
    # oracle BatchNorm on the global batch == normalisation of the shard with the reduced statistics
    p = {'bn.weight': torch.linspace(0.5, 1.5, 64, dtype=torch.float64), 'bn.bias': torch.linspace(-1, 1, 64, dtype=torch.float64)}
    b = {'bn.running_mean': torch.zeros(64, dtype=torch.float64), 'bn.running_var': torch.ones(64, dtype=torch.float64),
         'bn.num_batches_tracked': torch.zeros((), dtype=torch.int64)}
    nb = {}
    y_ref = R._bn(x, p, b, 'bn', True, nb)
    inv = (var + R.BN_EPS).rsqrt()
    y = (xs - mean[None, :, None, None]) * (inv * p['bn.weight'])[None, :, None, None] + p['bn.bias'][None, :, None, None]
    assert torch.allclose(y, y_ref[lo:hi], rtol=1e-10, atol=1e-10)
    # running statistics from the reduced sums (unbiased variance, momentum 0.1) equal the oracle's on the global batch
    rv = 0.9 * 1.0 + 0.1 * var * count / (count - 1)
    assert torch.allclose(rv, nb['bn.running_var'], rtol=1e-10)
    # backward sums: only rank 0 contributes the (already global) dgamma / dbeta to the gradient all-reduce
    gy = torch.randn(8, 64, 5, 6, generator=g, dtype=torch.float64)
    xhat = (xs - mean[None, :, None, None]) * inv[None, :, None, None]
    bst = torch.cat([gy[lo:hi].sum((0, 2, 3)), (gy[lo:hi] * xhat).sum((0, 2, 3))])
    dp.reduce_stats(bst)
    dgamma = bst[64:].clone() if rank == 0 else torch.zeros(64, dtype=torch.float64)
    dist.all_reduce(dgamma)
    xg = x.clone().requires_grad_(True)
    w = p['bn.weight'].clone().requires_grad_(True)
    yg = torch.nn.functional.batch_norm(xg, None, None, w, p['bn.bias'], True, 0.1, R.BN_EPS)
    yg.backward(gy)
    assert torch.allclose(dgamma, w.grad, rtol=1e-9, atol=1e-9)
    dx = (w.detach() * inv)[None, :, None, None] * (gy[lo:hi] - (bst[:64] / count)[None, :, None, None]
                                                   - xhat * (bst[64:] / count)[None, :, None, None])
    assert torch.allclose(dx, xg.grad[lo:hi], rtol=1e-8, atol=1e-10)
    # finish() without overlap: one fold of every layer, one all-reduce of the whole flat buffer
    eng = _StubEngine()
    flat = dict(names=['encoder.0.weight', 'decoder.0.weight'], offs=[0, 64], total=128, last_numel=64,
                grads=torch.full((128,), float(rank + 1)))
    dp.collectives = 0
    dp.finish(eng, None, None, flat)
    assert eng.calls == [None] and dp.collectives == 1
    assert torch.equal(flat['grads'], torch.full((128,), float(sum(range(1, world + 1)))))
    assert flat['buckets'] == {'decoder': (64, 128), 'rest': (0, 64)}


def test_dataparallel_object_syncbn_and_finish():
    _run(_w_dp_object)


def test_dataparallel_requires_process_group():
    with pytest.raises(RuntimeError):
        nd.DataParallel()
    # a model with a dp object but not the TMA engine refuses to run instead of silently ignoring it
    m = nsg.VQVAE(1, 64, 512).set_precision('fp32')
    m.__dict__['dp'] = object()
    with pytest.raises(Exception):
        m.engine()
    assert np.isfinite(1.0)


# ---------------------------------------------------------------------------------------------------------------------
# W-tiled inference of one long utterance (tiled.py): partition logic on CPU with a stand-in model whose encode /
# decode have the real model's strides and receptive fields (two 4/2/1 convolutions + two 3x3 at latent resolution;
# two 3x3 + two 4/2/1 transposed convolutions), single process and world-size-2 gloo
# ---------------------------------------------------------------------------------------------------------------------
class _StandIn(object):
    def __init__(self):
        g = torch.Generator().manual_seed(0)
        self.w = [torch.randn(4, 1, 4, 4, generator=g, dtype=torch.float64), torch.randn(4, 4, 4, 4, generator=g, dtype=torch.float64),
                  torch.randn(4, 4, 3, 3, generator=g, dtype=torch.float64), torch.randn(1, 4, 3, 3, generator=g, dtype=torch.float64)]
        self.d = [torch.randn(4, 1, 3, 3, generator=g, dtype=torch.float64), torch.randn(4, 4, 3, 3, generator=g, dtype=torch.float64),
                  torch.randn(4, 4, 4, 4, generator=g, dtype=torch.float64), torch.randn(4, 1, 4, 4, generator=g, dtype=torch.float64)]

    def encode(self, x):
        import torch.nn.functional as F
        h = F.conv2d(x.double(), self.w[0], None, 2, 1)
        h = F.conv2d(torch.tanh(h), self.w[1], None, 2, 1)
        h = F.conv2d(torch.tanh(h), self.w[2], None, 1, 1)
        h = F.conv2d(torch.tanh(h), self.w[3], None, 1, 1)
        return (h[:, 0] * 37.0).round().long()

    def decode(self, lat):
        import torch.nn.functional as F
        h = F.conv2d(lat.double().unsqueeze(1) / 37.0, self.d[0], None, 1, 1)
        h = F.conv2d(torch.tanh(h), self.d[1], None, 1, 1)
        h = F.conv_transpose2d(torch.tanh(h), self.d[2], None, 2, 1)
        return torch.tanh(F.conv_transpose2d(torch.tanh(h), self.d[3], None, 2, 1)).float()


def test_tile_bounds_and_single_process_tiling():
    from neural_sound_generation_b200 import tiled as T
    assert T.tile_bounds(10, 3) == [(0, 4), (4, 7), (7, 10)] and T.tile_bounds(2, 5) == [(0, 1), (1, 2)]
    assert T.halo_range(4, 7, 10, 3) == (1, 10, 3, 3) and T.halo_range(0, 4, 10, 3) == (0, 7, 0, 4)
    assert T.my_tiles(5, 1, 2) == [1, 3]
    m = _StandIn()
    x = torch.rand(1, 1, 16, 163, generator=torch.Generator().manual_seed(1))
    want_lat = m.encode(x)
    for n in (1, 2, 3, 7):
        lat, bounds = T.encode_tiled(m, x, n)
        assert torch.equal(lat, want_lat), n
        xt, _ = T.decode_tiled(m, want_lat, n)
        assert torch.allclose(xt, m.decode(want_lat), atol=1e-6), n


def _tiled_worker(rank, world, port, out):
    import os
    import torch.distributed as dist
    from neural_sound_generation_b200 import tiled as T
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        m = _StandIn()
        x = torch.rand(1, 1, 16, 120, generator=torch.Generator().manual_seed(2))
        lat, xt = T.encode_decode_sharded(m, x, n_tiles=3)
        ok = torch.equal(lat, m.encode(x)) and torch.allclose(xt, m.decode(m.encode(x)), atol=1e-6)
        out.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_tiled_inference_sharded_over_two_ranks_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tiled_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]
