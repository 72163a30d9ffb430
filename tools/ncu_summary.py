#!/usr/bin/env python
"""Print the handful of ncu raw-page metrics we steer by.  usage: ncu_summary.py raw.csv [src.csv]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__occupancy_limit_registers', 'launch__grid_size',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed_pipe_fma.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
for i, h in enumerate(hdr):
    if h in want or ('issue_stalled' in h and 'per_issue_active' in h):
        try:
            if 'stalled' in h and float(vals[i]) < 0.05:
                continue
        except ValueError:
            pass
        print(f"{h:90s} {units[i]:16s} {vals[i]}")
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    tot = sum(int(r[ix['Instructions Executed']] or 0) for r in rows[2:] if len(r) > 10)
    cum = 0
    for r in rows[2:]:
        if len(r) < 10:
            continue
        ie = int(r[ix['Instructions Executed']] or 0)
        cum += ie
        print(f"{r[ix['Address']][-5:]} {ie / 1e6:9.1f}M {100 * ie / tot:5.2f}% cum{100 * cum / tot:6.2f}% "
              f"thr{float(r[ix['Avg. Threads Executed']] or 0):5.1f} smp{int(r[ix['# Samples']] or 0):6d}  {r[ix['Source']][:80]}")
