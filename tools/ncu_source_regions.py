#!/usr/bin/env python
"""Hot regions of ONE kernel from `ncu -i rep --page source --csv` (SASS view): per block of B instructions the share of
stall samples, of executed warp instructions, the average active threads per warp and the dominant opcodes."""
import csv
import sys


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return 0.0


def main(path, B=300):
    hdr, data = None, []
    for r in csv.reader(open(path)):
        if r and r[0] == 'Address':
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            data.append(dict(zip(hdr, r)))
    tot_s = sum(num(d['# Samples']) for d in data) or 1.0
    tot_i = sum(num(d['Instructions Executed']) for d in data) or 1.0
    print('%d SASS instructions, %d samples, %.0f executed warp instructions' % (len(data), tot_s, tot_i))
    for i in range(0, len(data), B):
        blk = data[i:i + B]
        s = sum(num(d['# Samples']) for d in blk)
        n = sum(num(d['Instructions Executed']) for d in blk)
        if n == 0 and s == 0:
            continue
        ops = {}
        for d in blk:
            t = d['Source'].split()
            op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
            ops[op] = ops.get(op, 0) + num(d['Instructions Executed'])
        top = sorted(ops.items(), key=lambda kv: -kv[1])[:6]
        thr = sum(num(d['Thread Instructions Executed']) for d in blk) / max(n, 1)
        print('%5d  samples %5.1f%%  instr %5.1f%%  thr/warp %4.1f  %s' %
              (i, 100 * s / tot_s, 100 * n / tot_i, thr, [(k, int(v / 1000)) for k, v in top]))


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 300)
