#!/usr/bin/env python
"""profiles/roofline_traffic.json from an `ncu --set full` capture (what bench.py's `roofline.traffic` reads).

    ncu --set full --clock-control none -k regex:conv_tma_kernel -c 6 -f -o gpurun_out/<tag>_roof \
        python bench.py --kernel-only                  # launches ONLY the dominant kernel (3x3 64->64 forward)
    ncu -i gpurun_out/<tag>_roof.ncu-rep --page raw --csv > gpurun_out/<tag>_roof_raw.csv
    python tools/ncu_traffic.py gpurun_out/<tag>_roof_raw.csv [--out profiles/roofline_traffic.json]

Values are per launch: median over the kernel's launches in the capture of dram__bytes_read.sum and
dram__bytes_write.sum (ncu replays every launch with cold caches).  `--smem-kb X` keeps only launches with that
dynamic shared memory (to pick one layer shape out of a whole-step capture).
"""
import argparse
import csv
import json
import os
import statistics
import sys

ALGORITHMIC = 2 * 32 * 20 * 64 * 64 * 2 + 9 * 8192       # input + output bf16 (32 x 20 x 64 x 64) + packed weights


def col(header, name):
    hits = [i for i, h in enumerate(header) if h == name or h.endswith('.' + name)]
    if not hits:
        raise SystemExit('column %s not in the csv' % name)
    return hits[-1]


def num(v):
    try:
        return float(v.replace(',', ''))
    except ValueError:
        return 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('raw_csv')
    ap.add_argument('--out', default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'profiles',
                                                  'roofline_traffic.json'))
    ap.add_argument('--kernel', default='conv_tma_kernel')
    ap.add_argument('--taps', type=int, default=9, help='k*k of the layer (names the json key)')
    ap.add_argument('--smem-kb', type=float, default=None)
    args = ap.parse_args()
    rows = list(csv.reader(open(args.raw_csv)))
    header, units, body = rows[0], rows[1], rows[2:]
    c_name = col(header, 'Kernel Name')
    c_rd, c_wr = col(header, 'dram__bytes_read.sum'), col(header, 'dram__bytes_write.sum')
    c_t = col(header, 'gpu__time_duration.sum')
    c_sm = col(header, 'launch__shared_mem_per_block_dynamic')
    scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    launches = [r for r in body if args.kernel in r[c_name]]
    if not launches:
        raise SystemExit('no %s launches in %s' % (args.kernel, args.raw_csv))
    sel = launches
    if args.smem_kb is not None:
        sel = [r for r in launches if abs(num(r[c_sm]) - args.smem_kb) < 0.5]
        if not sel:
            raise SystemExit('no launch with %.1f KB of dynamic shared memory' % args.smem_kb)
    pick = statistics.median([num(r[c_sm]) for r in sel])
    rd = [num(r[c_rd]) * scale.get(units[c_rd], 1.0) for r in sel]
    wr = [num(r[c_wr]) * scale.get(units[c_wr], 1.0) for r in sel]
    dur = [num(r[c_t]) for r in sel]
    out = {}
    if os.path.exists(args.out):
        try:
            out = json.load(open(args.out))
        except Exception:
            out = {}
    key = '%s_%s_fwd' % (args.kernel, {1: '1x1', 9: '3x3', 16: '4x4'}.get(args.taps, str(args.taps)))
    out[key] = dict(dram_bytes_per_launch=int(statistics.median(rd) + statistics.median(wr)),
                    dram_read=int(statistics.median(rd)), dram_write=int(statistics.median(wr)),
                    launches=len(sel), duration_under_ncu=statistics.median(dur), duration_unit=units[c_t],
                    dynamic_smem_kb=pick, algorithmic_bytes=ALGORITHMIC if args.taps == 9 else None,
                    source='%s (ncu --set full --clock-control none; cold caches per ncu replay); written by '
                           'tools/ncu_traffic.py' % os.path.basename(args.raw_csv))
    json.dump(out, open(args.out, 'w'), indent=1)
    print(json.dumps(out[key]))


if __name__ == '__main__':
    sys.exit(main())
