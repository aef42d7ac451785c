#!/usr/bin/env python
"""Dev tool: one CSV row per .ncu-rep (kernel name, time, DRAM bytes, throughputs, issue/stall summary).
usage: python tools/ncu_summary.py out.csv rep1.ncu-rep rep2.ncu-rep ..."""
import csv, io, subprocess, sys
WANT = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__average_warp_latency_issue_stalled_barrier.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "l1tex__data_bank_conflicts_pipe_lsu.sum"]
out, reps = sys.argv[1], sys.argv[2:]
rows = []
hdr_out = None
for rep in reps:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(txt)))
    h, units, v = r[0], r[1], r[2]
    cols = [c for c in WANT if c in h]
    SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}
    BASE = {"byte": "byte", "Kbyte": "byte", "Mbyte": "byte", "Gbyte": "byte", "ns": "us", "us": "us", "ms": "us", "s": "us"}

    def norm(c):  # ncu picks a unit per report: bring bytes to bytes and times to microseconds
        u, x = units[h.index(c)], v[h.index(c)]
        return str(float(x.replace(",", "")) * SCALE[u]) if u in SCALE else x
    if hdr_out is None:
        hdr_out = cols
        rows.append(["report", "Kernel Name"] + [f"{c} [{BASE.get(units[h.index(c)], units[h.index(c)])}]" for c in cols])
    rows.append([rep.split("/")[-1], v[h.index("Kernel Name")]] + [norm(c) if c in h else "" for c in hdr_out])
csv.writer(open(out, "w")).writerows(rows)
for r in rows: print(",".join(r)[:400])
