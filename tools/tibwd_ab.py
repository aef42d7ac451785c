#!/usr/bin/env python
"""Dev tool (GPU box): three_interpolate backward (plan cached) on real three_nn lists, a few shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F


def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


for (B, C, m, n) in [(64, 64, 1024, 16384), (16, 128, 1024, 4096), (6, 384, 512, 2048), (16, 256, 256, 1024), (8, 64, 1024, 16384), (2, 64, 1000, 16000)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, n)).cuda()
    cen = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, m).long()[..., None].expand(-1, -1, 3)).contiguous()
    d2, i3 = F.three_nn(xyz, cen)
    w = 1.0 / (d2 + 1e-8); w = (w / w.sum(-1, keepdim=True)).contiguous()
    go = torch.randn(B, C, n, device="cuda")
    plan = F.ScatterPlan(i3, B, m)
    out = torch.empty(B, C, m, device="cuda")
    res = {}
    for name, mode in (("default", 0), ("stream16", 8), ("stream8", 5), ("direct", 1)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.three_interpolate_grad(go, w, plan, m, out=out))
            res[name + "_out"] = out.clone()
        finally:
            _lib.lib().geot_group_points_config(0)
    print(f"B={B} C={C} m={m} n={n}: default {res['default']:.4f}  stream16 {res['stream16']:.4f}  row-resident {res['stream8']:.4f}  direct {res['direct']:.4f} ms;"
          f" stream == direct bits: {bool(torch.equal(res['stream16_out'], res['direct_out']))}"
          f"  max |stream - default| {(res['stream16_out'] - res['default_out']).abs().max().item():.2e}", flush=True)
