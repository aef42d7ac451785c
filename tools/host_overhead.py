#!/usr/bin/env python
"""Dev tool (GPU box): where do the host microseconds of a 5-us op go?  (gather_operation C=3 is launch-bound.)"""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F
from tests import refgpu

B, N, M = 64, 16384, 1024
xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
xyz_t = xyz.transpose(1, 2).contiguous()
idx = F.furthest_point_sampling(xyz, M)
out = torch.empty((B, 3, M), device="cuda")
R = refgpu.lib("pn2batch")
p = lambda t: C.c_void_p(t.data_ptr())
L = _lib.lib()

def bench(name, fn, n=20000):
    for _ in range(200): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    host = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    tot = (time.perf_counter() - t0) / n * 1e6
    print(f"{name:56s} host {host:6.2f} us/call   incl. drain {tot:6.2f} us/call", flush=True)

bench("python no-op lambda", lambda: None)
bench("check_cuda(points, out, i32=(idx,))", lambda: _lib.check_cuda(xyz_t, out, i32=(idx,)))
bench("on_device enter/exit", lambda: _lib.on_device(xyz_t.device).__enter__())
bench("stream_ptr()", _lib.stream_ptr)
bench("4 x ptr()", lambda: (_lib.ptr(xyz_t), _lib.ptr(idx), _lib.ptr(out), _lib.ptr(None)))
bench("increment_version", lambda: torch.autograd.graph.increment_version(out))
bench("ctypes call geot_abi_version()", L.geot_abi_version)
bench("F.gather_points(xyz_t, idx, out=out)  [ours, full path]", lambda: F.gather_points(xyz_t, idx, out=out))
bench("F.gather_rows(xyz, idx, out=cen)", (lambda cen=torch.empty((B, M, 3), device='cuda'): F.gather_rows(xyz, idx, out=cen)))
bench("raw ctypes geot_gather_points (pointers precomputed)", (lambda a=xyz_t.data_ptr(), b=idx.data_ptr(), c=out.data_ptr(), s=_lib.stream_ptr(): L.geot_gather_points(B, 3, N, M, a, b, c, s)))
bench("ref launcher via ctypes (what op_sweep times)", lambda: R.ref_gather_points(B, 3, N, M, p(xyz_t), p(idx), p(out)))
