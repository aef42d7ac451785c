#!/usr/bin/env python
"""Dev tool (GPU box): gather_operation C=3 (bare C-ABI calls, CUDA-graph replay so the host is out of the picture)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib
from geot_b200 import functional as F

def graph_time(fn, reps=200):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps): fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); g.replay(); e1.record(s); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1000

for (B, C, N, M) in [(64, 3, 16384, 1024), (16, 3, 16384, 4096), (16, 3, 4096, 1024), (64, 64, 16384, 1024), (16, 32, 16384, 4096)]:
    pts = torch.randn(B, C, N, device="cuda")
    idx = torch.stack([torch.randperm(N, device="cuda")[:M] for _ in range(B)]).int().contiguous()
    out = torch.empty(B, C, M, device="cuda")
    res = {}
    for name, mode in (("default", 0),):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = graph_time(lambda: F.gather_points(pts, idx, out=out))
        finally:
            _lib.lib().geot_group_points_config(0)
    print(f"gather B={B} C={C} N={N} M={M}: " + "  ".join(f"{k} {v:.2f} us" for k, v in res.items()), flush=True)
