#!/usr/bin/env python
"""GPU box: three_interpolate backward, default dispatch vs direct (mode 1), on real three_nn indices."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, synth, functional as F


def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / it


for (B, C, m, n) in [(6, 384, 8192, 16000), (6, 384, 512, 8192), (6, 384, 512, 4096), (16, 128, 4096, 16384), (64, 64, 1024, 16384)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, 16384)).cuda()
    unknown = xyz[:, :n].contiguous(); known = xyz[:, torch.randperm(16384, device="cuda")[:m]].contiguous()
    d2, idx = F.three_nn(unknown, known)
    w = 1.0 / (d2 + 1e-8); w = (w / w.sum(-1, keepdim=True)).contiguous()
    g = torch.randn(B, C, n, device="cuda")
    plan = F.ScatterPlan(idx, B, m)
    out = torch.empty(B, C, m, device="cuda")
    res = []
    for mode in (0, 1):
        _lib.lib().geot_group_points_config(mode)
        res.append(t(lambda: F.three_interpolate_grad(g, w, plan, m, out=out)))
    _lib.lib().geot_group_points_config(0)
    gb = B * (24 * n + 4 * C * m + 4 * C * n) / 1e6
    print(f"B={B} C={C} m={m} n={n}: default {res[0]:.4f}  direct {res[1]:.4f} ms  plan build {t(lambda: F.ScatterPlan(idx, B, m)):.4f} ms  ({gb / res[0]:.0f} / {gb / res[1]:.0f} GB/s)", flush=True)
