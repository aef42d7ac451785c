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
