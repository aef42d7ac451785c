#!/usr/bin/env python
"""Dev tool (GPU box): A/B timing of ball query (grid path vs index-order scan) on the bench workload."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F

def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

for (B, N, M, r, ns) in [(64, 16384, 1024, 0.1, 32), (16, 16384, 4096, 0.1, 16), (16, 16384, 4096, 0.2, 64), (16, 4096, 1024, 0.4, 32), (64, 1024, 256, 0.8, 32),
                          (16, 16384, 4096, 0.05, 16), (16, 16384, 4096, 0.1, 32), (16, 16384, 4096, 0.2, 64), (16, 4096, 1024, 0.2, 32), (8, 65536, 4096, 0.05, 32)]:
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    idx = F.furthest_point_sampling(xyz, M)
    q = torch.gather(xyz, 1, idx.long()[..., None].expand(-1, -1, 3)).contiguous()
    ws = torch.empty(int(_lib.lib().geot_ball_query_workspace_bytes(B, N)), dtype=torch.uint8, device="cuda")
    out = torch.empty((B, M, ns), dtype=torch.int32, device="cuda")
    a = F.ball_query(q, xyz, r, ns).clone()
    b = F.ball_query(q, xyz, r, ns, use_grid=False).clone()
    c = F.ball_query(q, xyz, r, ns, flags=_lib.GEOT_BQ_FORCE_GRID).clone()
    ta = t(lambda: F.ball_query(q, xyz, r, ns, out=out, workspace=ws))
    tb = t(lambda: F.ball_query(q, xyz, r, ns, out=out, use_grid=False))
    tc = t(lambda: F.ball_query(q, xyz, r, ns, out=out, workspace=ws, flags=_lib.GEOT_BQ_FORCE_GRID))
    lv = []
    for L in (1, 2, 4, 8):
        fl = _lib.GEOT_BQ_FORCE_GRID | _lib.GEOT_BQ_LEVELS(L)
        d = F.ball_query(q, xyz, r, ns, flags=fl).clone()
        lv.append(f"L{L} {t(lambda: F.ball_query(q, xyz, r, ns, out=out, workspace=ws, flags=fl)):.3f}{'' if bool((a == d).all()) else ' MISMATCH'}")
    print(f"ball B={B} N={N} M={M} r={r} ns={ns}: auto {ta:.3f} ms  scan {tb:.3f} ms  forced-grid {tc:.3f} ms  equal={bool((a==b).all())},{bool((a==c).all())}  | " + "  ".join(lv), flush=True)
