#!/usr/bin/env python
"""Dev tool (GPU box): pieces of threeD_space_loss fwd+bwd at the cfg4 size, with and without the Morton processing order."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import synth, transition as T
from geot_b200 import functional as F

def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for C in (17, 33):
    B, N, k = 2, 16000, 32
    g = torch.Generator(device="cuda").manual_seed(0)
    pos = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=2)).cuda()
    labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
    insT = torch.softmax(torch.randn((B * N, C, C), device="cuda", generator=g), 2).contiguous()
    idx = T._neighbours(k, pos)
    order = F.spatial_order(pos)
    plan = F.ScatterPlan(idx, B, N)
    gl = torch.ones((1,), device="cuda")
    loss, coef = F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)
    print(f"C={C}: kNN(k+1) {t(lambda: T._neighbours(k, pos)):.3f} ms | spatial_order {t(lambda: F.spatial_order(pos)):.3f} | plan {t(lambda: F.ScatterPlan(idx, B, N)):.3f} | "
          f"fwd dataset order {t(lambda: F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D)):.3f}  Morton order {t(lambda: F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)):.3f} | "
          f"bwd dataset order {t(lambda: F.manifold_loss_backward(insT, idx, coef, plan, gl)):.3f}  Morton order {t(lambda: F.manifold_loss_backward(insT, idx, coef, plan, gl, order=order)):.3f}", flush=True)
