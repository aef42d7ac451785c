#!/usr/bin/env python
"""Dev tool: the handful of ncu metrics that decide a kernel, from a .ncu-rep (ncu -i ... --page raw --csv)."""
import csv, subprocess, sys
WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__thread_inst_executed_pred_on_per_inst_executed.ratio']
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines()))
h = r[0]
for row in r[2:]:
    for i, n in enumerate(h):
        if n in WANT or 'issue_stalled' in n and n.endswith('per_issue_active.ratio') and float(row[i] or 0) > 0.3:
            print(f"{n:90s} {row[i]:>16s} {r[1][i]}")
