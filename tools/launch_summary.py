"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total us, share."""
import collections
import csv
import sys


def main(path, top=30):
    rows = list(csv.reader(open(path)))
    hdr, agg = None, collections.OrderedDict()
    for r in rows:
        if 'Kernel Name' in r:
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        name = d['Kernel Name'].split('(')[0][:78]
        v = float(d['Metric Value'].replace(',', ''))
        unit = d['Metric Unit']
        v = v / 1e3 if unit in ('ns', 'nsecond') else (v * 1e3 if unit in ('ms', 'msecond') else v)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print('%-80s n=%4d  %10.1f us  avg %8.1f  %5.1f%%' % (k, a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
    print('total us %.1f over %d launches' % (tot, sum(a[0] for a in agg.values())))


if __name__ == '__main__':
    main(sys.argv[1])
