#!/usr/bin/env python
"""Condense `ncu --page raw --csv` output to the metrics that matter for HBM-bound kernels."""
import csv
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print('====', r[idx['Kernel Name']].strip())
        for k in KEYS:
            if k in idx:
                print('  %-62s %s %s' % (k, r[idx[k]], units[idx[k]]))
        stalls = []
        for i, h in enumerate(hdr):
            if 'average_warps_issue_stalled' in h and 'not_issued' not in h:
                try:
                    v = float(r[i])
                except ValueError:
                    continue
                if v > 0.3:
                    stalls.append((v, h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
        print('  stalls (warps per issue):', ', '.join('%s %.2f' % (n, v) for v, n in sorted(stalls, reverse=True)))


if __name__ == '__main__':
    main(sys.argv[1])
