#!/usr/bin/env python
"""Benchmark of the VQ-VAE train step (BASELINE.json config 2) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode fp32|bf16]

Prints ONE JSON line (rank 0).  `value` = mel-frames/s of the whole job with inputs resident in HBM;
`e2e` = the same step driven from pinned HOST buffers (H2D of the mel batch + D2H of the losses inside the
timed region); `roofline` = the dominant kernel timed alone with CUDA events; `cpu_baseline` = the oracle
port of the reference's CPU path on this box's host cores (bounded sample).
`--impl reference` times the reference's CPU algorithm (oracle port; /root/reference is not on the GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, N_MELS, FRAMES = 32, 80, 256          # BASELINE.json configs[1]
DIM, K_CODES = 64, 512
TRAIN_GFLOP = 76.84                              # SURVEY.md §8d, per 32x256 step


def ncu_traffic(kernel_key):
    """dram read+write bytes per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/roofline_traffic.json, written from the .ncu-rep by tools/ncu_traffic.py); None when absent."""
    p = os.path.join(ROOT, 'profiles', 'roofline_traffic.json')
    try:
        return json.load(open(p)).get(kernel_key, {}).get('dram_bytes_per_launch')
    except Exception:
        return None


def stock_torch_gpu_run(dev, steps=10, warmup=3, batch=B_PER_GPU):
    """'Stock PyTorch on the same B200' bars.  /root/reference does not exist on the GPU box, so
    the reference model is restated here with the same torch.nn layers and the same loop body (models.py:145-216,
    vector_quantization.py:6-44, train.py:116-136) and run on this GPU four ways: eager fp32 (torch's default cuDNN /
    TF32 settings), eager under torch.autocast(bfloat16), and each of the two captured as ONE CUDA graph (capturable
    Adam) -- the honest launch-bound comparison for a step that is ~100 small kernels.  Baselines only: nothing of
    ours is on this path, and it is never the product."""
    import torch.nn as nn
    import torch.nn.functional as F

    class Res(nn.Module):
        def __init__(self, d):
            super().__init__()
            self.block = nn.Sequential(nn.ReLU(True), nn.Conv2d(d, d, 3, 1, 1), nn.BatchNorm2d(d), nn.ReLU(True),
                                       nn.Conv2d(d, d, 1), nn.BatchNorm2d(d))

        def forward(self, x):
            return x + self.block(x)

    d = DIM

    def build():
        torch.manual_seed(1)
        enc = nn.Sequential(nn.Conv2d(1, d, 4, 2, 1), nn.BatchNorm2d(d), nn.ReLU(True), nn.Conv2d(d, d, 4, 2, 1), Res(d),
                            Res(d))
        dec = nn.Sequential(Res(d), Res(d), nn.ReLU(True), nn.ConvTranspose2d(d, d, 4, 2, 1), nn.BatchNorm2d(d),
                            nn.ReLU(True), nn.ConvTranspose2d(d, 1, 4, 2, 1), nn.Tanh())
        emb = nn.Embedding(K_CODES, d)
        return enc, dec, emb, nn.ModuleList([enc, dec, emb]).to(dev).train()

    c = torch.rand(batch, 1, N_MELS, FRAMES, device=dev)

    def fwd_loss(enc, dec, emb, amp):
        with torch.autocast('cuda', dtype=torch.bfloat16, enabled=amp):
            z_e = enc(c)
            zf = z_e.permute(0, 2, 3, 1).contiguous()
            flat = zf.view(-1, d)
            with torch.no_grad():
                cb = emb.weight
                dist_ = torch.addmm(torch.sum(cb ** 2, dim=1) + torch.sum(flat ** 2, dim=1, keepdim=True), flat, cb.t(),
                                    alpha=-2.0, beta=1.0)
                idx = torch.min(dist_, dim=1)[1]
            z_q_st = zf + (torch.index_select(emb.weight.detach(), 0, idx).view_as(zf) - zf).detach()
            z_q_bar = torch.index_select(emb.weight, 0, idx).view_as(zf).permute(0, 3, 1, 2).contiguous()
            x_t = dec(z_q_st.permute(0, 3, 1, 2).contiguous())
            return (F.mse_loss(x_t.float(), c) + F.mse_loss(z_q_bar.float(), z_e.detach().float())
                    + F.mse_loss(z_e.float(), z_q_bar.detach().float()))

    def time_eager(amp):
        enc, dec, emb, mods = build()
        opt = torch.optim.Adam(mods.parameters(), lr=1e-3)
        ts = []
        for i in range(warmup + steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            opt.zero_grad()
            loss = fwd_loss(enc, dec, emb, amp)
            loss.backward()
            opt.step()
            loss.item()
            torch.cuda.synchronize()
            if i >= warmup:
                ts.append(time.perf_counter() - t0)
        return 1e3 * sorted(ts)[len(ts) // 2]

    def time_graphed(amp):
        enc, dec, emb, mods = build()
        opt = torch.optim.Adam(mods.parameters(), lr=1e-3, capturable=True)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                opt.zero_grad(set_to_none=True)
                fwd_loss(enc, dec, emb, amp).backward()
                opt.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph):
            loss = fwd_loss(enc, dec, emb, amp)
            loss.backward()
            opt.step()
        for _ in range(warmup):
            graph.replay()
        torch.cuda.synchronize()
        evs = []
        for _ in range(steps):
            s_, e_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s_.record()
            graph.replay()
            e_.record()
            evs.append((s_, e_))
        torch.cuda.synchronize()
        assert torch.isfinite(loss).item()
        ts = sorted(a.elapsed_time(b) for a, b in evs)
        return ts[len(ts) // 2]

    out = dict(unit='mel-frames/s', what='torch.nn restatement of the reference step on this GPU (cuDNN / cuBLAS), batch '
               '%d x %d x %d; ms_per_step per variant, value = the FASTEST variant' % (batch, N_MELS, FRAMES))
    variants = {}
    for name, fn, amp in (('eager_fp32_tf32', time_eager, False), ('eager_autocast_bf16', time_eager, True),
                          ('cuda_graph_fp32_tf32', time_graphed, False), ('cuda_graph_autocast_bf16', time_graphed, True)):
        try:
            variants[name] = fn(amp)
        except Exception as e:      # a baseline must never take the bench line down
            variants[name] = None
            out.setdefault('errors', {})[name] = str(e)[:160]
    ok = {k: v for k, v in variants.items() if v}
    out['ms_per_step_by_variant'] = variants
    if ok:
        best = min(ok, key=ok.get)
        out.update(value=batch * FRAMES / (ok[best] / 1e3), ms_per_step=ok[best], fastest=best)
    return out


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d['hbm_gbs'], tf_burst=d['bf16_tflops'], tf_sust=d.get('bf16_tflops_sustained', d['bf16_tflops']),
                    src='measured')
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src='fallback')


class ClockSampler(object):
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.samples, self.proc = [], None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for s in self.samples:
            f = [x.strip() for x in s.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))


def cpu_reference_run(steps, warmup, batch):
    """The reference's CPU path (oracle port, all host threads) on a bounded sample of the workload."""
    from oracle import vqvae_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    p, b = R.make_state(1, DIM, K_CODES, 1, codebook='spread')
    mels = [R.make_mel(batch, N_MELS, FRAMES, 100 + i) for i in range(2)]
    state = R.adam_init(p)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        r = R.step_grads(p, b, mels[i % 2], 1.0)
        p = R.adam_step(p, r['grads'], state, 1e-3)
        b = r['buffers']
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return dict(value=batch * FRAMES / (ms / 1e3), ms_per_step=ms, cores=torch.get_num_threads(),
                sample='%d steps of fwd+bwd+Adam on a %dx%dx%d mel batch (fp32, oracle port of the reference CPU path)'
                       % (steps, batch, N_MELS, FRAMES))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours')
    ap.add_argument('--mode', default=os.environ.get('NSG_BENCH_MODE', 'tma'), choices=['tma', 'bf16', 'tc32', 'fp32'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--sync-bn', action='store_true', default=os.environ.get('NSG_BENCH_SYNC_BN', '0') == '1',
                    help='N > 1: all-reduce BatchNorm statistics (N ranks == the reference on the global batch)')
    ap.add_argument('--no-overlap', action='store_true', help='N > 1: one gradient all-reduce after backward')
    ap.add_argument('--kernel-only', action='store_true',
                    help='launch only the dominant kernel (3x3 64->64 forward) a few times: the target of the ncu '
                         '--set full capture that tools/ncu_traffic.py turns into profiles/roofline_traffic.json')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    config = dict(workload='vqvae_train_step VQVAE(1,64,512) batch %dx%dx%d per GPU, Adam lr 1e-3, beta 1.0'
                           % (B_PER_GPU, N_MELS, FRAMES), parallelism='dp%d' % world, global_batch=B_PER_GPU * world)

    if args.impl == 'reference':
        if rank != 0:
            return
        # the stated configuration (batch 32 x 80 x 256), BASELINE.md section 3: >= 5 steps after >= 2 warm-ups when the
        # caller allows; bounded above so the arm always ends within a few minutes (~1-2 s per step on 16 cores)
        steps, warm = max(1, min(args.steps, 10)), max(1, min(args.warmup, 3))
        r = cpu_reference_run(steps, warm, B_PER_GPU)
        out = dict(metric='vqvae_train_mel_frames_per_sec', value=r['value'], unit='mel-frames/s', n_gpus=args.gpus,
                   steps=steps, warmup=warm, ms_per_step=r['ms_per_step'], higher_is_better=True, scaling='weak',
                   vs_baseline=None, dtype='f32', data='synthetic', config=config, impl='reference',
                   cpu_baseline=dict(value=r['value'], unit='mel-frames/s', cores=r['cores'], kind='port',
                                     sample=r['sample']),
                   e2e=dict(value=r['value'], unit='mel-frames/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        print(json.dumps(out))
        return

    import torch.distributed as dist
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import ops
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    torch.manual_seed(1)                                           # main.py:43 default
    model = nsg.VQVAE(1, DIM, K_CODES)
    with torch.no_grad():
        model.codebook.embedding.weight.normal_()                  # trained-like codebook
    model = model.to(dev)
    if args.mode == 'tma':         # bf16 operand-ready activations, TMA-fed persistent tcgen05 convolutions: the production
        # path IS the model's automatic choice -- nothing is set here, exactly as main.py:120-124 constructs it
        assert (model.conv_impl, model.act_dtype, model.vq_mode) == (2, torch.bfloat16, 1)
    elif args.mode == 'bf16':      # bf16 activations, register-staged tcgen05 convolutions, tcgen05 split-bf16 VQ search
        model.act_dtype, model.conv_impl, model.vq_mode = torch.bfloat16, 1, 1
    elif args.mode == 'tc32':      # fp32 activations, tcgen05 convolutions with split-bf16 operands (3 passes)
        model.act_dtype, model.conv_impl, model.vq_mode = torch.float32, 1, 1
    else:                          # fp32 CUDA-core parity engine
        model.set_precision('fp32')
    model.train()
    if args.kernel_only:
        eng = model.engine()
        x = torch.randn(B_PER_GPU, 20, 64, DIM, device=dev).to(model.act_dtype).relu()
        Pd = {n: q.detach() for n, q in model.named_parameters()}
        kctx = eng.begin(x, True)
        eng._pack(kctx, Pd)
        Bd = dict(model.named_buffers())
        kst = torch.zeros(2 * DIM, dtype=torch.float64, device=dev)
        for _ in range(6):          # the launch the step makes for this layer: convolution + BatchNorm moment sums
            eng._conv_fwd(kctx, Pd, 'encoder.4.block.1', x, stats=kst)
        torch.cuda.synchronize()
        print(json.dumps(dict(kernel_only=True, launches=6)))
        return
    opt = nsg.optim.Adam(model.parameters(), lr=1e-3)                  # main.py:124
    flat = model.flatten_parameters()
    g = torch.Generator(device='cpu').manual_seed(1 + rank)
    n_in = 8
    host = [torch.rand(B_PER_GPU, N_MELS, FRAMES, generator=g).pin_memory() for _ in range(n_in)]
    devin = [h.to(dev) for h in host]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)       # > 126 MB L2

    # the public train-step object: the train_vqvae loop body as one CUDA graph (train.py uses the same class)
    # N > 1: dist.DataParallelStep = the same graph with the bucketed NCCL gradient all-reduce (decoder bucket beside
    # the encoder backward), optional BatchNorm-statistic all-reduces, loss averaging; gradients averaged via grad_scale
    if world > 1:
        graphed = nsg.dist.DataParallelStep(model, opt, (B_PER_GPU, 1, N_MELS, FRAMES), beta=1.0,
                                            sync_bn=args.sync_bn, overlap=not args.no_overlap)
    else:
        graphed = nsg.GraphedTrainStep(model, opt, (B_PER_GPU, 1, N_MELS, FRAMES), beta=1.0)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        evs = []
        for i in range(steps):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn(i)
            e.record()
            evs.append((s, e))
        sync()
        ms = sum(s.elapsed_time(e) for s, e in evs)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item() / steps

    def timed_back_to_back(fn, steps):
        """Average duration of `steps` launches issued back to back between ONE pair of events (no flush kernel in
        between: the caller rotates inputs larger than L2, or wants the warm figure).  The host gets a head start -- a
        ~1 ms spin kernel without memory traffic -- so that the launches are queued when their turn comes and the events
        bracket GPU time, not the host's launch latency."""
        torch.cuda._sleep(2000000)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for i in range(steps):
            fn(i)
        e.record()
        sync()
        return s.elapsed_time(e) / steps

    for i in range(max(3, args.warmup)):
        graphed(devin[i % n_in])
    sync()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ops.launch_count
    model.loss_and_grads(devin[0], beta=1.0)          # count the launches of one (eager) step, outside the timing
    opt.step()
    launches = ops.launch_count - l0 + 1              # + zero_grad memset
    sync()
    ms = timed(lambda i: graphed(devin[i % n_in]), args.steps)
    clocks = sampler.stop() if sampler else None
    collectives = graphed.collectives_per_step if world > 1 else 0

    # end to end: pinned host mel batch -> device, graph replay, losses -> host, every step
    # (double-buffered: the copy of batch i+1 is issued before the host waits for step i's losses, so it runs under
    # the step; every timed step still moves one whole batch host -> device and its three losses device -> host)
    def e2e_step(i):
        graphed(host[i % n_in])
        graphed.prefetch(host[(i + 1) % n_in])
        return graphed.read_losses()        # (device -> pinned host copy queued behind the replay + one stream sync)
    e2e_step(n_in - 1)
    sync()
    ms_e2e = timed(e2e_step, args.steps)
    frames = B_PER_GPU * FRAMES * world
    value = frames / (ms / 1e3)

    # N > 1: the OTHER BatchNorm mode measured in the same run, so that one line (and the driver's SCALE file) carries
    # both curves: `value` is the mode selected by --sync-bn (default: local statistics), `other_bn_mode` the other one
    other = None
    if world > 1 and args.mode == 'tma':
        graphed.close()
        model.__dict__['_grads_dirty'] = True
        g2 = nsg.dist.DataParallelStep(model, opt, (B_PER_GPU, 1, N_MELS, FRAMES), beta=1.0,
                                       sync_bn=not args.sync_bn, overlap=not args.no_overlap)
        for i in range(max(3, args.warmup)):
            g2(devin[i % n_in])
        sync()
        ms2 = timed(lambda i: g2(devin[i % n_in]), args.steps)
        other = dict(sync_bn=not args.sync_bn, ms_per_step=ms2, value=frames / (ms2 / 1e3), unit='mel-frames/s',
                     collectives_per_step=g2.collectives_per_step,
                     bn=('global-batch statistics (reference-equivalent)' if not args.sync_bn
                         else 'local batch statistics per rank (DDP semantics)'))
        graphed = g2

    # dominant kernel alone: 3x3 ResBlock convolution forward (implicit GEMM 40960 x 64 x 576, 3.02 GFLOP)
    pk = peaks()
    eng = model.engine()
    x = torch.randn(B_PER_GPU, 20, 64, DIM, device=dev).to(model.act_dtype)
    w = dict(model.named_parameters())['encoder.4.block.1.weight'].detach()
    bias = dict(model.named_parameters())['encoder.4.block.1.bias'].detach()
    if args.mode == 'tma':
        Pd = {n: q.detach() for n, q in model.named_parameters()}
        kctx = eng.begin(x, True)
        eng._pack(kctx, Pd)
        xr = x.relu()
        Bd = {n: q.clone() for n, q in model.named_buffers()}      # (the running statistics of the model stay untouched)

        kstats = torch.zeros(2 * DIM, dtype=torch.float64, device=dev)

        def kfn(i):                 # the launch the step makes for this layer: convolution + BatchNorm moment sums
            return eng._conv_fwd(kctx, Pd, 'encoder.4.block.1', xr, stats=kstats)

        def kfused(i):              # the BatchNorm-fused form of the same layer (NSG_FUSE_BN=1), timed for the record
            if kctx.nbar >= 30:     # (one zeroed grid-barrier counter per launch)
                kctx.bars.zero_()
                kctx.nbar = 0
            fb, eng.fuse_bn = eng.fuse_bn, True
            try:
                return eng._conv_bn_fwd(kctx, Pd, Bd, 'encoder.4.block.1', 'encoder.4.block.2', xr, relu_out=True)
            finally:
                eng.fuse_bn = fb
        kplain = kfused
    else:
        kfn = lambda i: eng._conv_fwd(x, w, bias, 3, 1, 1, False, ops.prologue(ops.PRO_RELU))
    for _ in range(3):
        kfn(0)
    kms_flushed = timed(kfn, 20)
    kms_warm = timed_back_to_back(kfn, 20)
    kms = kms_flushed
    kl2 = '256 MiB write-flush before every launch'
    if args.mode == 'tma':
        # The roofline figure: inputs larger than L2 instead of the write-flush.  The flush leaves 126 MB of DIRTY lines
        # whose write-back is charged to the ~10 us kernel that follows, and bracketing ONE launch with events adds the
        # event / launch gaps to it (a step-sized region amortises both, a single kernel does not).  Here 64 distinct
        # inputs = 336 MB cycle through the L2 -- launch i reads a tensor last touched 64 launches (>= 336 MB of traffic)
        # ago, i.e. from DRAM; weights and the allocator-reused output stay hot, as they do inside the step -- and the
        # figure is the average of 20 launches issued back to back with programmatic dependent launch, as the step
        # issues them.  The flushed single-launch figure and the warm figure are reported beside it.
        nrot = 64
        xs = [xr.clone() for _ in range(nrot)]

        def kfn_rot(i):
            return eng._conv_fwd(kctx, Pd, 'encoder.4.block.1', xs[i % nrot], stats=kstats)
        for i in range(nrot):
            kfn_rot(i)
        kms = timed_back_to_back(kfn_rot, 20)
        kl2 = ('average of 20 back-to-back launches on 64 rotating inputs (336 MB > 126 MB L2), no flush: every launch '
               'reads its input from DRAM')
        del xs
    kms_plain = None
    if args.mode == 'tma':
        for _ in range(3):
            kplain(0)
        kms_plain = timed(kplain, 20)
    kflop = 2.0 * x.numel() // DIM * DIM * DIM * 9
    achieved = kflop / (kms / 1e3) / 1e12
    # second half of BASELINE.json's metric: the VQ search alone at this workload's bottleneck size
    nlat = B_PER_GPU * 20 * (FRAMES // 4)
    zq_in = torch.randn(nlat, DIM, device=dev).to(model.act_dtype)
    cbk = model.codebook.embedding.weight.detach()
    vidx = torch.empty(nlat, dtype=torch.int64, device=dev)
    vcodes, vsq = torch.empty_like(zq_in), torch.zeros(1, dtype=torch.float64, device=dev)
    vmode = model.vq_mode if ops.vq_mode_serves(K_CODES, DIM, model.vq_mode) else 0
    vfn = lambda i: ops.vq_forward(zq_in, cbk, vidx, vcodes, vsq, mode=vmode)
    for _ in range(3):
        vfn(0)
    vms = timed(vfn, 20)
    vq_tf = 2.0 * nlat * K_CODES * DIM / (vms / 1e3) / 1e12
    vq_search = dict(value=vq_tf, unit='TFLOP/s', frac_of_peak=vq_tf / pk['tf_burst'], us=vms * 1e3, n=nlat, k=K_CODES,
                     d=DIM, what='vq_forward (search + gather + sq-error; split-bf16 passes not counted), L2 flushed')
    roof = dict(bound='tensor', achieved=achieved, peak=pk['tf_burst'], unit='TFLOP/s', frac=achieved / pk['tf_burst'],
                traffic=ncu_traffic('conv_tma_kernel_3x3_fwd'), kernel='conv3x3 64->64 fwd (%s)' % ({0: 'CUDA-core fp32', 1: 'tcgen05 bf16, register-staged', 2: 'tcgen05 bf16, TMA-fed persistent, BatchNorm moment sums in the epilogue'}[model.conv_impl]),
                us=kms * 1e3, l2=kl2, us_l2_write_flushed=kms_flushed * 1e3, us_warm=kms_warm * 1e3,
                us_fused_conv_bn_relu=(kms_plain * 1e3 if kms_plain else None),
                peak_source=pk['src'], step_frac_of_tensor_peak=TRAIN_GFLOP * world / (ms / 1e3) / 1e3 / (pk['tf_sust'] * world))
    if rank != 0:
        if world > 1:
            nsg.dist.shutdown(graphed)
        return
    cpu, stock = None, None
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_run(5, 2, B_PER_GPU)
        cpu = dict(value=r['value'], unit='mel-frames/s', cores=r['cores'], kind='port', sample=r['sample'])
        try:
            stock = stock_torch_gpu_run(dev)
        except Exception as e:          # a baseline must never take the bench line down
            stock = dict(unavailable=str(e)[:200])
    out = dict(metric='vqvae_train_mel_frames_per_sec', value=value, unit='mel-frames/s', n_gpus=world,
               steps=args.steps, warmup=max(3, args.warmup), ms_per_step=ms, higher_is_better=True, scaling='weak',
               vs_baseline=None, dtype='bf16' if args.mode in ('bf16', 'tma') else 'f32', data='synthetic',
               config=dict(config, l2='256 MiB flush between timed steps; 8 rotating input batches',
                           mode=args.mode,
                           bn=('batch statistics' if world == 1 else
                               ('global-batch statistics (all-reduced, reference-equivalent)' if args.sync_bn
                                else 'local batch statistics per rank (DDP semantics)')),
                           collectives_per_step=collectives),
               e2e=dict(value=frames / (ms_e2e / 1e3), unit='mel-frames/s', ms_per_step=ms_e2e,
                        h2d_bytes_per_step=B_PER_GPU * N_MELS * FRAMES * 4, d2h_bytes_per_step=24),
               gpu_launches=launches, clocks=clocks, roofline=roof, cpu_baseline=cpu, vq_search=vq_search,
               stock_torch_gpu=stock)
    if other is not None:
        out['sync_bn' if other['sync_bn'] else 'local_bn'] = other
    print(json.dumps(out), flush=True)
    if world > 1:
        nsg.dist.shutdown(graphed)


if __name__ == '__main__':
    main()
