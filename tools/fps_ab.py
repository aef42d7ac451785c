#!/usr/bin/env python
"""Dev tool (GPU box): A/B timing of the FPS kernels, checked against each other.
  mgr = default: bucket-pruned manager/worker kernel (fps.cu)     brute = un-pruned one-CTA kernel
  ref = the reference kernel (oracle/_ref), when present"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F

def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

try:
    from tests import refgpu
    have_ref = refgpu.available("pn2batch")
except Exception:
    have_ref = False
rows = []
shapes = [(64, 16384, 1024), (8, 16384, 1024), (16, 16384, 4096), (6, 16000, 8192), (16, 4096, 1024), (64, 4096, 1024), (64, 1024, 256), (256, 16384, 1024)]
if len(sys.argv) > 1 and sys.argv[1] == "quick":
    shapes = shapes[:4]
if len(sys.argv) > 1 and sys.argv[1] == "small":
    shapes = [(64, 4096, 1024), (64, 3072, 768), (64, 2048, 512), (64, 1536, 384), (64, 1024, 256), (16, 1024, 256), (64, 512, 128), (64, 256, 64), (64, 128, 32), (64, 64, 16)]
for (B, N, M) in shapes:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    variants = {"mgr": 0, "spec": _lib.GEOT_FPS_SPEC, "brute": _lib.GEOT_FPS_NO_PRUNE}
    base = F.furthest_point_sampling(xyz, M, flags=_lib.GEOT_FPS_NO_PRUNE)
    row = {"B": B, "N": N, "M": M}
    for name, fl in variants.items():
        out = F.furthest_point_sampling(xyz, M, flags=fl)
        row[name + "_equal"] = bool((out == base).all())
        row[name + "_ms"] = t(lambda: F.furthest_point_sampling(xyz, M, flags=fl))
    row["mgr_setup_us"] = t(lambda: F.furthest_point_sampling(xyz, 2)) * 1e3
    row["mgr_us_per_round"] = (row["mgr_ms"] * 1e3 - row["mgr_setup_us"]) / (M - 2)
    if have_ref and B * M <= 70000:
        row["ref_equal"] = bool((refgpu.fps(xyz, M) == base).all())
        row["ref_ms"] = t(lambda: refgpu.fps(xyz, M), it=3)
    rows.append(row)
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()}), flush=True)
