#!/usr/bin/env python
"""Dev tool (GPU box): A/B timing of grouping forward (row-staged vs direct gather)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib
from geot_b200 import functional as F

def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for (B, C, N, M, ns) in [(64, 64, 16384, 1024, 32), (64, 3, 16384, 1024, 32), (16, 64, 16384, 4096, 64), (16, 128, 4096, 1024, 32), (64, 64, 4096, 1024, 16)]:
    feat = torch.randn(B, C, N, device="cuda")
    idx = torch.randint(0, N, (B, M, ns), device="cuda", dtype=torch.int32)
    out = torch.empty(B, C, M, ns, device="cuda")
    byt = B * (4 * M * ns + 4 * C * N + 4 * C * M * ns)
    res = []
    for direct in (0, 1):
        _lib.lib().geot_group_points_config(direct)
        ms = t(lambda: F.group_points(feat, idx, out=out))
        res.append(ms)
    _lib.lib().geot_group_points_config(0)
    print(f"group B={B} C={C} N={N} M={M} ns={ns}: row-staged {res[0]:.3f} ms ({byt/res[0]/1e6:.0f} GB/s)  direct {res[1]:.3f} ms ({byt/res[1]/1e6:.0f} GB/s)", flush=True)
