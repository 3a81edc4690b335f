"""Per-launch summary of an `ncu --set full` report exported with `ncu -i X.ncu-rep --page raw --csv`."""
import csv
import sys

COLS = [('gpu__time_duration.sum', 'us'), ('launch__grid_size', 'grid'), ('launch__registers_per_thread', 'regs'),
        ('launch__shared_mem_per_block_dynamic', 'smemKB'), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor%act'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'tensor%el'),
        ('dram__bytes_read.sum', 'dramR'), ('dram__bytes_write.sum', 'dramW'),
        ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'), ('lts__t_sector_hit_rate.pct', 'L2hit%'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'bankconf')]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = [(hdr.index(c), n) for c, n in COLS if c in hdr]
    kn = hdr.index('Kernel Name')
    print('%-44s ' % 'kernel' + ' '.join('%10s' % (n + '[' + units[i][:6] + ']')[:10] for i, n in ix))
    for r in rows[2:]:
        print('%-44s ' % r[kn][:44] + ' '.join('%10s' % r[i][:10] for i, _ in ix))


if __name__ == '__main__':
    main(sys.argv[1])
