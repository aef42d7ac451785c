#!/usr/bin/env python
"""Dev tool (GPU box): manifold loss forward / backward, tile kernels (default) against the warp-per-point kernels (mode 6)
at the cfg4 size (B=2 x 16000, C=33, k=32) and two smaller ones.  CUDA-event times over 20 calls."""
import json, os, sys
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


for (B, N, C, k) in [(2, 16000, 33, 32), (2, 16000, 17, 32), (4, 8000, 33, 16), (2, 16000, 5, 32)]:
    g = torch.Generator(device="cuda").manual_seed(1)
    pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
    insT = torch.softmax(torch.randn((B * N, C, C), device="cuda", generator=g), 2).contiguous()
    idx = T._neighbours(k, pos)
    order = F.spatial_order(pos)
    plan = F.ScatterPlan(idx, B, N)
    gl = torch.ones((1,), device="cuda")
    row = {"B": B, "N": N, "C": C, "k": k, "order_ms": round(timeit(lambda: F.spatial_order(pos)), 4)}
    for name, m in (("tile", 7), ("warp", 6)):
        _lib.lib().geot_group_points_config(m)
        try:
            loss, coef = F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)
            row[name + "_fwd_ms"] = round(timeit(lambda: F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)), 4)
            row[name + "_bwd_ms"] = round(timeit(lambda: F.manifold_loss_backward(insT, idx, coef, plan, gl, order=order)), 4)
        finally:
            _lib.lib().geot_group_points_config(0)
    print(json.dumps(row), flush=True)
