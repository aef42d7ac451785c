#!/usr/bin/env python
"""Dev tool (GPU box): three_interpolate forward, persistent interleaved kernel (default) against the grid-per-pair
interleaved kernel (mode 4) and the plain row-staged kernel (mode 3).  One JSON line per shape."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib
from geot_b200 import functional as F

PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]) if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6550.0


def timeit(fn, it=20, reps=5):
    """Device time per call: `it` calls captured into one CUDA graph (no host gaps between 10-us kernels)."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(it): fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(reps): g.replay()
        e1.record(side)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (it * reps)


for (B, C, m, n) in [(64, 64, 1024, 16384), (32, 64, 1024, 16384), (16, 64, 1024, 16384), (8, 64, 1024, 16384), (2, 64, 1024, 16384),
                     (16, 128, 1024, 4096), (6, 384, 512, 2048), (6, 384, 512, 8192), (16, 256, 64, 256), (16, 256, 256, 1024),
                     (16, 128, 256, 1024), (64, 32, 1024, 16384), (16, 512, 256, 1024), (6, 128, 1000, 16000)]:
    g = torch.Generator(device="cuda").manual_seed(1)
    feat = torch.randn(B, C, m, device="cuda", generator=g)
    idx = torch.randint(0, m, (B, n, 3), device="cuda", generator=g, dtype=torch.int32)
    w = torch.rand(B, n, 3, device="cuda", generator=g); w = (w / w.sum(-1, keepdim=True)).contiguous()
    out = torch.empty(B, C, n, device="cuda")
    row = {"B": B, "C": C, "m": m, "n": n}
    nbytes = 4 * B * C * (n + m) + 24 * B * n
    ref = None
    for name, mode in (("persistent", 0), ("grid_rows4", 4), ("rows", 3)):
        _lib.lib().geot_group_points_config(mode)
        try:
            ms = timeit(lambda: F.three_interpolate(feat, idx, w, out=out))
            cur = out.clone()
        finally:
            _lib.lib().geot_group_points_config(0)
        ref = cur if ref is None else ref
        row[name + "_ms"] = round(ms, 5)
        row[name + "_hbm_frac"] = round(nbytes / ms / 1e6 / PEAK, 3)
        row[name + "_equal"] = bool(torch.equal(cur, ref))
    print(json.dumps(row), flush=True)
