#!/usr/bin/env python
"""Dev tool: per-SASS-instruction stall samples from an .ncu-rep (source page), grouped by executed count.
usage: python tools/ncu_src.py rep [start_idx end_idx]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[hi], rows[hi + 1:]
isrc, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[isamp]) for r in data)
print("total samples", tot, "instructions", len(data))
if len(sys.argv) > 3:
    a, b = int(sys.argv[2]), int(sys.argv[3])
    for n, r in enumerate(data[a:b], a):
        s = int(r[isamp])
        st = max(((int(r[i]), hdr[i][6:]) for i in stall), default=(0, ""))
        print(n, r[isrc].strip()[:56].ljust(56), str(s).rjust(6), f"{100*s/tot:5.1f}%", r[iex].rjust(9), st if s else "")
else:
    cur = None
    for n, r in enumerate(data):
        ex, s = int(r[iex]), int(r[isamp])
        if cur and cur[0] == ex: cur[1] += 1; cur[2] += s; cur[4] = n
        else:
            if cur: print(f"instr {cur[3]}-{cur[4]} n={cur[1]} executed={cur[0]} samples={cur[2]} ({100*cur[2]/tot:.1f}%)")
            cur = [ex, 1, s, n, n]
    print(f"instr {cur[3]}-{cur[4]} n={cur[1]} executed={cur[0]} samples={cur[2]} ({100*cur[2]/tot:.1f}%)")
