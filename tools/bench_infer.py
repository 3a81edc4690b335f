"""BASELINE.json configs[4]: inference encode -> quantise -> decode of one ~10 s utterance (batch 1, 861 mel frames x 80
bins), eval mode.  Reports p50 / p99 latency over 200 runs (CUDA events around each run, host->device copy of the
mel and device->host copy of the reconstruction included in the e2e figures) and mel-frames/s, eager vs CUDA graph,
next to the oracle port of the reference on the host cores (median of 20).

    python tools/bench_infer.py [--frames 861] [--runs 200]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import neural_sound_generation_b200 as nsg  # noqa: E402


def pct(ts, q):
    ts = sorted(ts)
    return ts[min(len(ts) - 1, int(q * len(ts)))]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=861)
    ap.add_argument('--runs', type=int, default=200)
    ap.add_argument('--no-cpu', action='store_true')
    a = ap.parse_args()
    dev = torch.device('cuda', 0)
    torch.manual_seed(1)
    model = nsg.VQVAE(1, 64, 512)
    with torch.no_grad():
        model.codebook.embedding.weight.normal_()
    model = model.to(dev).eval()
    model.act_dtype, model.conv_impl, model.vq_mode = torch.bfloat16, 2, 1
    x_host = torch.rand(1, 1, 80, a.frames).pin_memory()
    out_host = torch.empty(1, 1, 80, 4 * (a.frames // 4)).pin_memory()
    x_dev = x_host.to(dev)
    res = {}

    def run(fn, name):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(a.runs):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e))
        res[name] = dict(p50_ms=pct(ts, 0.5), p99_ms=pct(ts, 0.99), frames_per_s=a.frames / (pct(ts, 0.5) / 1e3))

    with torch.no_grad():
        def eager():
            lat = model.encode(x_dev)
            return model.decode(lat)
        run(eager, 'eager_device_resident')
        infer = nsg.GraphedInference(model, (1, 1, 80, a.frames))
        run(lambda: infer(x_dev), 'graph_device_resident')

        def e2e():
            _, xt = infer(x_host)
            out_host.copy_(xt, non_blocking=True)
        run(e2e, 'graph_e2e_pinned_host')
    if not a.no_cpu:
        from oracle import vqvae_ref as R
        torch.set_num_threads(os.cpu_count() or 1)
        p, b = R.make_state(1, 64, 512, 1, codebook='spread')
        ts = []
        with torch.no_grad():
            for i in range(23):
                t0 = time.perf_counter()
                R.decode(p, b, R.encode(p, b, x_host))
                if i >= 3:
                    ts.append(1e3 * (time.perf_counter() - t0))
        res['cpu_oracle_port'] = dict(p50_ms=pct(ts, 0.5), cores=torch.get_num_threads(),
                                      frames_per_s=a.frames / (pct(ts, 0.5) / 1e3))
    print(json.dumps(dict(workload='inference encode->quantise->decode, batch 1 x 80 x %d frames, eval mode' % a.frames,
                          runs=a.runs, results=res)))


if __name__ == '__main__':
    main()
