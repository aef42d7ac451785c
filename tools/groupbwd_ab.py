#!/usr/bin/env python
"""Dev tool (GPU box): grouping_operation backward (plan cached) on real ball-query lists: the half-row pipeline of
seg_reduce_planrow_kernel (default) against the same kernel waiting for whole rows (mode 9) and the direct kernel."""
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


for (B, C, N, M, r, ns) in [(64, 64, 16384, 1024, 0.1, 32), (16, 32, 16384, 4096, 0.1, 32), (16, 128, 4096, 1024, 0.4, 64),
                            (64, 128, 1024, 256, 0.8, 32), (8, 64, 16384, 1024, 0.1, 32), (16, 64, 16384, 2048, 0.05, 16)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    cen = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, M).long()[..., None].expand(-1, -1, 3)).contiguous()
    idx = F.ball_query(cen, xyz, r, ns)
    go = torch.randn(B, C, M, ns, device="cuda")
    plan = F.grouping_plan(idx, B, N)
    out = torch.empty(B, C, N, device="cuda")
    res = {}
    for name, mode in (("halves", 11), ("whole", 9), ("direct", 1), ("pf", 13), ("pfhalves", 10), ("pfgrid", 12), ("default", 0)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.group_points_grad(go, plan, N, out=out))
            res[name + "_out"] = out.clone()
        finally:
            _lib.lib().geot_group_points_config(0)
    print(f"B={B} C={C} N={N} M={M} ns={ns} r={r}: persistent+halves {res['halves']:.4f}  r02 kernel (grid, whole rows) {res['whole']:.4f}  direct {res['direct']:.4f} ms;"
          f" | DEFAULT {res['default']:.4f}, persistent+prefetch {res['pf']:.4f}, persistent+prefetch+halves {res['pfhalves']:.4f}, grid+prefetch {res['pfgrid']:.4f}; {bool(torch.equal(res['pf_out'], res['direct_out']))},{bool(torch.equal(res['pfhalves_out'], res['direct_out']))}"
          f" bits equal: {bool(torch.equal(res['halves_out'], res['direct_out']))},{bool(torch.equal(res['whole_out'], res['direct_out']))}", flush=True)
