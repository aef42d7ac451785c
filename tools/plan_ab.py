#!/usr/bin/env python
"""Dev tool (GPU box): scatter-plan build, fused one-launch builder (default where it applies) against the multi-kernel
path (mode 2), on kNN / ball-query index tensors of a few shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F
from geot_b200 import transition as T


def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


for (B, N, k) in [(2, 16000, 32), (1, 16000, 32), (4, 8000, 16), (2, 16000, 16), (8, 16000, 32), (16, 4096, 32)]:
    pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    idx = T._neighbours(k, pos)
    res = {}
    for name, mode in (("default", 0), ("multi_kernel", 2)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.ScatterPlan(idx, B, N))
        finally:
            _lib.lib().geot_group_points_config(0)
    print(f"B={B} N={N} k={k} E={N*k}: default {res['default']:.4f} ms   multi-kernel {res['multi_kernel']:.4f} ms", flush=True)

# ball-query index tensors as the grouping backward builds them (grouping_plan: split into pieces of <= 65535 sources):
# E = M * ns sources into N targets, most targets empty, the lowest indices of every ball listed by many queries
for (B, N, M, ns, r) in [(64, 16384, 1024, 32, 0.1), (16, 16384, 4096, 32, 0.1), (16, 4096, 1024, 64, 0.4), (16, 1024, 256, 64, 0.8),
                          (16, 16384, 4096, 64, 0.2), (16, 4096, 1024, 32, 0.2), (16, 16384, 4096, 16, 0.05)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    cen = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, M).long()[..., None].expand(-1, -1, 3)).contiguous()
    idx = F.ball_query(cen, xyz, r, ns)
    res = {}
    for name, mode in (("default", 0), ("multi_kernel", 2)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.grouping_plan(idx, B, N))
        finally:
            _lib.lib().geot_group_points_config(0)
    print(f"ball query B={B} N={N} M={M} ns={ns} E={M*ns}: default {res['default']:.4f} ms   multi-kernel {res['multi_kernel']:.4f} ms", flush=True)

# three_nn index tensors as three_interpolate's backward builds them: E = 3 n sources into m coarse targets, ~3n/m sources
# per target - the ordered builder (stable counting sort, default) against the fused builder (mode 15) and the multi-kernel path
for (B, n, m) in [(64, 16384, 1024), (16, 16384, 1024), (16, 4096, 1024), (16, 4096, 256), (6, 8192, 2048), (16, 1024, 256), (2, 16000, 1000), (16, 16384, 4096)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, n)).cuda()
    cen = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, m).long()[..., None].expand(-1, -1, 3)).contiguous()
    _, i3 = F.three_nn(xyz, cen)
    res = {}
    for name, mode in (("default", 0), ("fused", 15), ("multi_kernel", 2)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.ScatterPlan(i3, B, m))
            pl = F.ScatterPlan(i3, B, m)
            res[name + "_plan"] = (pl.row_ptr().clone(), pl.perm().clone())
        finally:
            _lib.lib().geot_group_points_config(0)
    eq = all(torch.equal(a, b) for a, b in zip(res["default_plan"], res["multi_kernel_plan"])) and all(torch.equal(a, b) for a, b in zip(res["fused_plan"], res["multi_kernel_plan"]))
    print(f"three_nn B={B} n={n} m={m} E={3*n}: default {res['default']:.4f} ms   fused {res['fused']:.4f} ms   multi-kernel {res['multi_kernel']:.4f} ms   equal={eq}", flush=True)
