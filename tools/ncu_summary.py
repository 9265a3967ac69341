#!/usr/bin/env python3
"""Summarise ncu output (read here, no GPU needed).
  launches:  tools/ncu_summary.py launches gpurun_out/launches.csv [skip_rows]
  raw:       tools/ncu_summary.py raw gpurun_out/prof.ncu-rep
"""
import collections
import csv
import subprocess
import sys

RAW = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
       'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
       'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
       'launch__occupancy_limit_shared_mem', 'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed.sum',
       'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
       'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
       'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum',
       'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
       'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio',
       'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio']


def launches(path):
    lines = [l for l in open(path) if not l.startswith('==')]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = row['Kernel Name'][:90]
        v = float(row['Metric Value'].replace(',', ''))
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
        print(f"{t / 1e6:10.3f} ms {n:5d}x {t / n / 1e3:10.1f} us/launch {100 * t / tot:5.1f}%  {k}")


def raw(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index('Kernel Name')
    print('kernel', [r[ki][:40] for r in data])
    for g in ('Grid Size', 'Block Size'):
        print(g, [r[hdr.index(g)] for r in data])
    for w in RAW:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w} [{units[i]}]", [r[i] for r in data])


if __name__ == '__main__':
    {'launches': launches, 'raw': raw}[sys.argv[1]](sys.argv[2])
