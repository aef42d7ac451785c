#!/usr/bin/env python
"""Dev tool (GPU box): whole-kernel ablation timing of the FPS round (prof build; ablated runs give wrong indices)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib
_lib.SO_PATH = os.path.join(os.path.dirname(_lib.SO_PATH), "prof", "libgeot_b200_prof.so")
from geot_b200 import synth
from geot_b200 import functional as F

def t(fn, it=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for (B, N, M) in [(8, 16384, 1024), (8, 4096, 1024)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    for name, base in (("mgr", 0), ("spec", _lib.GEOT_FPS_SPEC)):
        setup = t(lambda: F.furthest_point_sampling(xyz, 2, flags=base)) * 1e3
        row = []
        for abl in (0, 1):
            ms = t(lambda: F.furthest_point_sampling(xyz, M, flags=base | (abl << 8)))
            row.append((ms * 1e3 - setup) / (M - 2) * 1.965e3)
        print(f"N={N} {name:5s} cycles/round: full {row[0]:7.0f}   no updates {row[1]:7.0f}   (setup {setup:.0f} us)", flush=True)
