#!/usr/bin/env python
"""Dev tool (GPU box): where do the cycles of one FPS round go?  Loads the -DGEOT_FPS_PROF build of the library
(make -C geot_b200/csrc prof), runs one cloud batch and prints the per-phase SM-clock deltas of block 0."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib
_lib.SO_PATH = os.path.join(os.path.dirname(_lib.SO_PATH), "prof", "libgeot_b200_prof.so")
from geot_b200 import synth
from geot_b200 import functional as F

which = sys.argv[1] if len(sys.argv) > 1 else "mgr"
B, N, M = 8, 16384, 1024
xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
buf = torch.zeros(((M + 1) * 32 * 8,), dtype=torch.int64, device="cuda")
L = _lib.lib()
L.geot_debug_fps_prof.argtypes = [C.c_void_p]
flags = 0
F.furthest_point_sampling(xyz, M, flags=flags)
assert L.geot_debug_fps_prof(buf.data_ptr()) == 0
F.furthest_point_sampling(xyz, M, flags=flags)
torch.cuda.synchronize()
L.geot_debug_fps_prof(None)
s = buf.cpu().numpy().reshape(M + 1, 32, 8).astype(np.int64)
R = M - 1
if which == "mgr":
    mg = s[:, 0, :]             # manager stamps, iterations 1..R
    wk = s[:, 1:, :]            # workers, rounds 0..R-1
    rows = []
    for j in range(20, R - 1):  # steady state
        pub = mg[j, 3]                              # manager published sample j (arrive on barrier 1)
        w0 = wk[j, :, 0]                            # workers woke
        act = wk[j, :, 2] > 0
        w2 = np.where(wk[j, :, 2] >= w0, wk[j, :, 2], w0)   # update done (only stamped when active this round)
        stale = wk[j, :, 2] < w0
        w3 = wk[j, :, 3]
        nxt = mg[j + 1, 0]                          # manager woke for iteration j+1
        rows.append([mg[j, 1] - mg[j, 0], mg[j, 2] - mg[j, 1], mg[j, 3] - mg[j, 2], w0.max() - pub, np.median(w0) - pub,
                     (w3 - w0).max(), nxt - w3.max(), nxt - mg[j, 0], (~stale).sum()])
    a = np.array(rows, dtype=np.float64)
    names = ["mgr: wake->pick done", "mgr: pick->coords", "mgr: coords->published", "publish->last worker awake", "publish->median worker awake",
             "slowest worker awake->arrive", "last worker arrive->mgr awake", "whole round", "workers that updated"]
    for i, nme in enumerate(names):
        print(f"{nme:34s} mean {a[:, i].mean():8.1f}  median {np.median(a[:, i]):8.1f}  p90 {np.percentile(a[:, i], 90):8.1f}")
    # one-bucket update cost: rounds in which the slowest worker did exactly... report the distribution of (w2-w1) per active worker
    d = []
    for j in range(20, R - 1):
        ok = wk[j, :, 2] >= wk[j, :, 0]
        d.extend(list((wk[j, ok, 2] - wk[j, ok, 1])))
        dd = (wk[j, ok, 1] - wk[j, ok, 0])
    d = np.array(d)
    print(f"active worker: flags seen -> update(s) done: mean {d.mean():.1f} median {np.median(d):.1f} p10 {np.percentile(d,10):.1f} p90 {np.percentile(d,90):.1f}  (n={len(d)})")

    for a_, b_, nm in [(1, 4, "flags seen -> distances done (last active bucket)"), (4, 5, "distances -> REDUX max + rank select"), (5, 6, "rank select -> REDUX min + key stored"), (6, 2, "key stored -> loop exit")]:
        d = []
        for j in range(20, R - 1):
            ok = (wk[j, :, 2] >= wk[j, :, 1]) & (wk[j, :, 1] > 0)
            d.extend(list(wk[j, ok, b_] - wk[j, ok, a_]))
        d = np.array(d)
        print(f"{nm:52s} median {np.median(d):7.1f}  p10 {np.percentile(d,10):7.1f}  p90 {np.percentile(d,90):7.1f}")
