#!/usr/bin/env python
"""Dev tool (GPU box): A/B timing of the FPS kernels on the bench workload, checked against each other."""
import os, sys
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

for (B, N, M) in [(64, 16384, 1024), (16, 16384, 4096), (6, 16000, 8192), (64, 4096, 1024), (64, 1024, 256), (256, 16384, 1024)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    a = F.furthest_point_sampling(xyz, M)
    b = F.furthest_point_sampling(xyz, M, flags=_lib.GEOT_FPS_NO_PRUNE)
    ok = bool((a == b).all())
    ta = t(lambda: F.furthest_point_sampling(xyz, M))
    tb = t(lambda: F.furthest_point_sampling(xyz, M, flags=_lib.GEOT_FPS_NO_PRUNE))
    t2 = t(lambda: F.furthest_point_sampling(xyz, 2))
    print(f"FPS B={B} N={N} M={M}: mgr {ta:.3f} ms  unpruned {tb:.3f} ms  speedup {tb/ta:.2f}x  equal={ok}  ({(ta-t2)*1e3/(M-2):.3f} us/round, setup(M=2) {t2*1e3:.1f} us)", flush=True)
