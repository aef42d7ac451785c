#!/usr/bin/env python
"""Dev tool (GPU box): three_nn grid kernel with the ring-2 queue (default) against every query finished in place (mode 14)."""
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


for (B, n, m) in [(64, 16384, 1024), (16, 4096, 1024), (6, 8192, 2048), (16, 16384, 256), (8, 16384, 1024), (6, 16000, 512), (16, 16384, 4096), (6, 16000, 8192), (1, 100000, 16000)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, max(n, m))).cuda()
    unknown = xyz[:, :n].contiguous()
    known = torch.gather(xyz, 1, F.furthest_point_sampling(xyz, m).long()[..., None].expand(-1, -1, 3)).contiguous()
    res = {}
    for name, mode in (("queue", 0), ("inplace", 14)):
        _lib.lib().geot_group_points_config(mode)
        try:
            res[name] = timeit(lambda: F.three_nn(unknown, known))
            res[name + "_out"] = [x.clone() for x in F.three_nn(unknown, known)]
        finally:
            _lib.lib().geot_group_points_config(0)
    eq = all(torch.equal(a, b) for a, b in zip(res["queue_out"], res["inplace_out"]))
    print(f"three_nn B={B} n={n} m={m}: ring-2 queue {res['queue']:.4f} ms  in place {res['inplace']:.4f} ms  equal={eq}", flush=True)
