#!/usr/bin/env python
"""Dev tool (GPU box): cost of the grid kernel's per-CTA binning = a forced-grid three_nn with 32 queries per cloud,
replayed from a CUDA graph (no host gaps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F

def graph_time(fn, reps=50):
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

for (B, n, m) in [(64, 32, 1024), (64, 1024, 1024), (64, 2048, 1024), (64, 32, 256), (64, 32, 4096), (16, 32, 8192)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, max(n, m, 16384))).cuda()
    unknown = xyz[:, :n].contiguous()
    known = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, m).long()[..., None].expand(-1, -1, 3)).contiguous()
    torch.cuda.synchronize(); print("shape", B, n, m, flush=True)
    d2 = torch.empty(B, n, 3, device="cuda"); ix = torch.empty(B, n, 3, dtype=torch.int32, device="cuda")
    _lib.lib().geot_group_points_config(2)
    try:
        t = graph_time(lambda: _lib.call("geot_three_nn", B, n, m, _lib.ptr(unknown), _lib.ptr(known), _lib.ptr(d2), _lib.ptr(ix), _lib.stream_ptr()))
    finally:
        _lib.lib().geot_group_points_config(0)
    print(f"forced grid three_nn B={B} n={n} m={m}: {t:.1f} us per launch", flush=True)
