#!/usr/bin/env python
"""GPU box: three_interpolate forward, default dispatch (channel-interleaved rows) vs direct (mode 1) vs row-staged without the
interleaved kernel (mode 3)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, functional as F


def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / it


for (B, C, m, n) in [(6, 384, 8192, 16000), (6, 384, 512, 8192), (6, 384, 512, 4096), (16, 128, 4096, 6000), (64, 64, 1024, 16384), (16, 32, 4096, 16384), (16, 128, 1024, 4096)]:
    pts = torch.randn(B, C, m, device="cuda")
    idx = torch.randint(0, m, (B, n, 3), device="cuda", dtype=torch.int32)
    w = torch.rand(B, n, 3, device="cuda"); w = w / w.sum(-1, keepdim=True)
    out = torch.empty(B, C, n, device="cuda")
    res = []
    for mode in (0, 1, 3):
        _lib.lib().geot_group_points_config(mode)
        res.append(t(lambda: F.three_interpolate(pts, idx, w, out=out)))
    _lib.lib().geot_group_points_config(0)
    gb = B * (24 * n + 4 * C * m + 4 * C * n) / 1e6
    print(f"B={B} C={C} m={m} n={n}: default {res[0]:.4f}  direct {res[1]:.4f}  rows(old) {res[2]:.4f} ms   ({gb / res[0]:.0f} / {gb / res[1]:.0f} / {gb / res[2]:.0f} GB/s)", flush=True)
