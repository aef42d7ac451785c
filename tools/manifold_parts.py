#!/usr/bin/env python
"""Dev tool (GPU box): where the time of threeD_space_loss forward + backward goes at the cfg4 size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import synth
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


B, N, C, k = 2, 16000, 33, 32
g = torch.Generator(device="cuda").manual_seed(1)
pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
insT = torch.softmax(torch.randn((B * N, C, C), device="cuda", generator=g), 2).contiguous()
idx = T._neighbours(k, pos)
order = F.spatial_order(pos)
plan = F.ScatterPlan(idx, B, N)
gl = torch.ones((1,), device="cuda")
loss, coef = F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)
print(f"kNN k+1=33          {timeit(lambda: T._neighbours(k, pos)):.4f} ms")
print(f"spatial_order       {timeit(lambda: F.spatial_order(pos)):.4f} ms")
print(f"forward             {timeit(lambda: F.manifold_loss_forward(pos, labels, insT, idx, 1.0, F.MANIFOLD_3D, order=order)):.4f} ms")
print(f"scatter plan build  {timeit(lambda: F.ScatterPlan(idx, B, N)):.4f} ms")
print(f"backward            {timeit(lambda: F.manifold_loss_backward(insT, idx, coef, plan, gl, order=order)):.4f} ms")
mod = T.threeD_space_loss(k=k, sigma=1.0, num_classes=C)
t1 = insT.clone().requires_grad_(True)
def full():
    t1.grad = None
    mod(pos, labels, t1, idx=idx).backward()
print(f"module fwd+bwd (idx given) {timeit(full):.4f} ms")
def full_knn():
    t1.grad = None
    mod(pos, labels, t1).backward()
print(f"module fwd+bwd (with kNN)  {timeit(full_knn):.4f} ms")
