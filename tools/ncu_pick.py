#!/usr/bin/env python3
"""Print selected metrics of an `ncu --page raw --csv` export, one column per captured launch."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
pats = sys.argv[2:] or ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed.sum', 'smsp__inst_executed.avg.per_cycle_active',
    'smsp__issue_active.avg.pct', 'sm__inst_executed_pipe_fma', 'sm__pipe_fma', 'sm__inst_executed_pipe_alu', 'sm__inst_executed_pipe_lsu',
    'sm__inst_executed_pipe_xu', 'sm__inst_executed_pipe_uniform','sm__inst_executed_pipe_cbu', 'sm__inst_executed_pipe_adu','dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__warps_eligible', 'issue_stalled', 'smsp__cycles_active.avg',
    'sm__throughput.avg.pct']
for i, h in enumerate(hdr):
    if any(p in h for p in pats):
        print("%-92s %-14s %s" % (h[:92], units[i][:14], " | ".join(r[i][:26] for r in rows[2:])))
