#!/usr/bin/env python
"""GPU box: kNN grid path vs brute-force scan (GEOT_KNN_NO_GRID) vs the reference kernel, offset layout."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F
import refgpu

RP = refgpu.lib("pointops") if refgpu.available("pointops") else None
p = lambda t: C.c_void_p(t.data_ptr())


def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


for (b, n_per, m_per, k) in [(8, 16384, 16384, 16), (8, 4096, 4096, 16), (8, 16384, 4096, 16), (8, 16384, 16384, 32),
                             (8, 16384, 16384, 3), (8, 16384, 16384, 64), (2, 2048, 2048, 16), (64, 16384, 1024, 16)]:
    n, m = b * n_per, b * m_per
    xyz = torch.from_numpy(synth.tooth_batch(b, n_per)).cuda().reshape(n, 3).contiguous()
    q = xyz.view(b, n_per, 3)[:, :m_per].reshape(m, 3).contiguous()
    off = torch.arange(1, b + 1, device="cuda", dtype=torch.int32) * n_per
    noff = torch.arange(1, b + 1, device="cuda", dtype=torch.int32) * m_per
    idx = torch.empty((m, k), dtype=torch.int32, device="cuda"); d = torch.empty((m, k), device="cuda")
    idx2 = torch.empty_like(idx); d2 = torch.empty_like(d)
    tg = timeit(lambda: F.knnquery(k, xyz, q, off, noff, m_per=m_per, idx=idx, dist2=d))
    ts = timeit(lambda: F.knnquery(k, xyz, q, off, noff, m_per=m_per, idx=idx2, dist2=d2, flags=_lib.GEOT_KNN_NO_GRID))
    tr = timeit(lambda: RP.ref_knnquery(m, k, p(xyz), p(q), p(off), p(noff), p(idx2), p(d2))) if RP else float("nan")
    ws = torch.empty(int(_lib.lib().geot_knnquery_workspace_bytes(b, n, m)), dtype=torch.uint8, device="cuda")
    F.knnquery(k, xyz, q, off, noff, m_per=m_per, idx=idx, dist2=d)
    print(f"b={b} n_per={n_per} m_per={m_per} k={k}: grid {tg:.4f} ms  scan {ts:.4f} ms  ref {tr:.4f} ms  match={bool((idx == idx2).all())}", flush=True)
