#!/usr/bin/env python
"""Summarise ncu output for profiles/: `ncu_summary.py launches <csv>` groups a
gpu__time_duration launch list by kernel; `ncu_summary.py rep <file.ncu-rep>`
prints the roofline-relevant raw metrics of every captured launch."""
import collections
import csv
import subprocess
import sys

KEYS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'l1tex__throughput.avg.pct_of_peak_sustained_active',
    'lts__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__warps_active.avg.pct_of_peak_sustained_active',
    'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
    'launch__shared_mem_per_block_static', 'launch__shared_mem_per_block_dynamic',
    'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
    'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
    'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'l1tex__data_pipe_lsu_wavefronts.sum',
    'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
    'smsp__cycles_active.avg', 'sm__cycles_elapsed.max',
]


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
    tot = collections.OrderedDict()
    for r in rows[1:]:
        k = r[ki].split('(')[0]
        c = tot.setdefault(k, [0, 0.0])
        c[0] += 1
        c[1] += float(r[vi].replace(',', ''))
    total = sum(v for _, v in tot.values())
    print('%-70s %6s %12s %7s' % ('kernel', 'count', 'total us', 'share'))
    for k, (c, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print('%-70s %6d %12.1f %6.1f%%' % (k[:70], c, v / 1e3, 100 * v / total))
    print('%-70s %6d %12.1f' % ('TOTAL', sum(c for c, _ in tot.values()), total / 1e3))


def rep(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('== %s' % r[hdr.index('Kernel Name')].split('(')[0])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print('  %-70s %18s %s' % (k, r[i], units[i]))
        stalls = []
        for i, h in enumerate(hdr):
            if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio'):
                try:
                    stalls.append((float(r[i]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]))
                except ValueError:
                    pass
        for v, n in sorted(stalls, reverse=True)[:8]:
            print('  stall %-64s %18.3f warps/issue' % (n, v))


if __name__ == '__main__':
    {'launches': launches, 'rep': rep}[sys.argv[1]](sys.argv[2])
