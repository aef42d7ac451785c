#!/usr/bin/env python
"""Dev tool (GPU box): phase times of the fused scatter-plan builder (block 0), prof build."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib
_lib.SO_PATH = os.path.join(os.path.dirname(_lib.SO_PATH), "prof", "libgeot_b200_prof.so")
from geot_b200 import synth
from geot_b200 import functional as F
L = _lib.lib()
for (B, N, M, ns, r, what) in [(64, 16384, 1024, 32, 0.1, "ball"), (16, 4096, 1024, 64, 0.4, "ball"), (64, 1024, 16384, 3, 0, "nn3"), (16, 4096, 16384, 3, 0, "nn3")]:
    xyz = torch.from_numpy(synth.tooth_batch(B, 16384)).cuda()
    if what == "ball":
        x = xyz[:, :N].contiguous()
        cen = F.gather_rows(x, F.furthest_point_sampling(x, M))
        idx = F.ball_query(cen, x, r, ns)
        tgt = N
    else:
        cen = F.gather_rows(xyz, F.furthest_point_sampling(xyz, N))
        _, idx = F.three_nn(xyz, cen)
        tgt = N
    for _ in range(3):
        F.ScatterPlan(idx, B, tgt)
    torch.cuda.synchronize()
    buf = (C.c_longlong * 16)()
    L.geot_debug_plan_prof.argtypes = [C.c_void_p]
    assert L.geot_debug_plan_prof(C.addressof(buf)) == 0
    s = list(buf)
    names = ["zero", "count", "scan", "fill", "sort (thread+warp tiers)", "CTA bitmap tier"]
    print(f"{what} B={B} targets={tgt} E={idx.numel() // B}: " + "  ".join(f"{n} {(s[i + 1] - s[i]) / 1965:.1f}us" for i, n in enumerate(names) if s[i + 1] >= s[i]) + f"  very long segments {s[7]}")
