#!/usr/bin/env python
"""Builds profiles/REPORT_<tag>.md from a bench line (profiles/<tag>_bench.json) and an op sweep
(profiles/<tag>_op_sweep.json).  usage: python tools/make_report.py r01b"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
b = json.load(open(os.path.join(ROOT, "profiles", f"{tag}_bench.json")))
sw = json.load(open(os.path.join(ROOT, "profiles", f"{tag}_op_sweep.json")))
L = [f"# Measured on B200 — {tag}", "",
     "Sources: `profiles/%s_bench.json` (python bench.py), `profiles/%s_op_sweep.json` (python tests/op_sweep.py), "
     "`profiles/%s_launches_bench.csv` + `profiles/%s_ncu_full_summary.csv` (ncu)." % (tag, tag, tag, tag), "",
     "## Bench line (B=64 x N=16384 per GPU, FPS(1024) + gather + ball_query(0.1, 32) + grouping C=3 + grouping C=64)", "",
     f"* device-resident: **{b['value']/1e6:.0f} M points/s** ({b['ms_per_step']:.3f} ms/step)",
     f"* end to end (host buffers, {b['e2e']['h2d_bytes_per_step']/1e6:.0f} MB H2D + {b['e2e']['d2h_bytes_per_step']/1e6:.1f} MB D2H per step): "
     f"**{b['e2e']['value']/1e6:.0f} M points/s** ({b['e2e']['ms_per_step']:.2f} ms/step; the PCIe copy is the bound)",
     f"* CPU port of the reference semantics, {b['cpu_baseline']['cores']} threads: {b['cpu_baseline']['value']/1e6:.2f} M points/s",
     f"* roofline kernel: {b['roofline']['achieved']:.0f} GB/s algorithmic = {100*b['roofline']['frac']:.1f} % of the measured {b['roofline']['peak']:.0f} GB/s HBM peak"
     f" (DRAM traffic {b['roofline']['traffic']/1e6 if b['roofline'].get('traffic') else float('nan'):.0f} MB per launch)", "",
     "| kernel | ms | share | algorithmic GB/s | % HBM |", "|---|---|---|---|---|"]
for k, v in b["kernels"].items():
    L.append(f"| {k} | {v['ms']:.4f} | {100*v['share']:.1f} % | {v['hbm_gbs']:.0f} | {100*v['hbm_frac']:.1f} |")
L += ["", "## Every op against the reference's own kernel (oracle/_ref, same GPU, same inputs)", "",
      "| op | shape | ours ms | reference ms | speed-up | algorithmic GB/s | % HBM | results equal |", "|---|---|---|---|---|---|---|---|"]
for r in sw["rows"]:
    ref = f"{r['ref_ms']:.4f}" if r["ref_ms"] else "-"
    sp = f"{r['speedup']:.2f}x" if r["speedup"] else "-"
    L.append(f"| {r['op']} | {r['shape']} | {r['ours_ms']:.4f} | {ref} | {sp} | {r['gbs']:.0f} | {100*r['hbm_frac']:.1f} | {r['match_ref']} |")
tp = os.path.join(ROOT, "profiles", f"{tag}_transition_sweep.json")
if os.path.exists(tp):
    ts = json.load(open(tp))
    L += ["", "## Transition-matrix path against the reference's torch expressions (oracle/transition_ref.py) on the same GPU", "",
          "`python tests/transition_sweep.py`; forward and forward+backward per op.  `max rel err` of the weight gradient includes "
          "clamp-edge flips: a pre-activation within one ulp of the 1e-5 clamp edge passes gradient on one side and none on the "
          "other (torch fp32 against torch fp64 differs by the same amount).", "",
          "| op | shape | ours ms | torch ms | speed-up | algorithmic GB/s | % HBM | max rel err |", "|---|---|---|---|---|---|---|---|"]
    for r in ts["rows"]:
        L.append(f"| {r['op']} | {r['shape']} | {r['ours_ms']:.4f} | {r['ref_ms']:.4f} | {r['speedup']:.2f}x | {r['gbs']:.0f} | {100*r['hbm_frac']:.1f} | {r['max_rel_err']:.1e} |")
open(os.path.join(ROOT, "profiles", f"REPORT_{tag}.md"), "w").write("\n".join(L) + "\n")
print("\n".join(L[:24]))
