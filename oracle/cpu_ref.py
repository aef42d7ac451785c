"""ctypes/numpy front-end of oracle/cpu_ref.c (TEST INFRASTRUCTURE ONLY — see oracle/__init__.py).

Each function mirrors one reference entry point and cites it; arguments are numpy arrays, results
are new numpy arrays.  Parity status: pinned against the reference's own kernels (oracle/_ref, built
unchanged from /root/reference) on a B200 — see tests/test_ref_parity_gpu.py and tests/golden/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgeot_oracle.so")
_lib = None

_f = np.float32
_i = np.int32


def build(force: bool = False) -> str:
    """Compile cpu_ref.c (gcc, OpenMP).  Called by __graft_entry__.build() and lazily by load()."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "cpu_ref.c")):
        subprocess.run(["make", "-C", _HERE, "_build/libgeot_oracle.so"], check=True, capture_output=True)
    return _SO


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_opt_n_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def num_threads() -> int:
    return int(load().oracle_num_threads())


def set_num_threads(n: int) -> None:
    """torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline (bench.py) asks for all host threads."""
    load().oracle_set_num_threads(C.c_int(int(n)))


def opt_n_threads(n: int, cap: int = 1024) -> int:
    """cuda_utils.h:10-14 block-size rule."""
    return int(load().oracle_opt_n_threads(C.c_int(n), C.c_int(cap)))


# --- batched API (openpoints/cpp/pointnet2_batch, pointnet2/_ext) -----------------------------------
def furthest_point_sample(xyz, npoint, cap=1024, mag_skip=False):
    """subsample.py:76-98 → sampling_gpu.cu:101-216.  cap=512, mag_skip=True is pointnet2/_ext (variant #4)."""
    xyz = _c(xyz, _f)
    B, N, _ = xyz.shape
    temp = np.full((B, N), 1e10, _f)
    idx = np.zeros((B, npoint), _i)
    load().oracle_fps_batched(C.c_int(B), C.c_int(N), C.c_int(npoint), _p(xyz), _p(temp), _p(idx), C.c_int(cap), C.c_int(int(mag_skip)))
    return idx


def ball_query(radius, nsample, xyz, new_xyz):
    """group.py:177-203 → ball_query_gpu.cu:15-51.  Argument order as the reference's Python API."""
    xyz, new_xyz = _c(xyz, _f), _c(new_xyz, _f)
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    idx = np.zeros((B, M, nsample), _i)
    load().oracle_ball_query(C.c_int(B), C.c_int(N), C.c_int(M), C.c_float(radius), C.c_int(nsample), _p(new_xyz), _p(xyz), _p(idx))
    return idx


def grouping_operation(features, idx):
    """group.py:76-117 fwd → group_points_gpu.cu:53-72."""
    features, idx = _c(features, _f), _c(idx, _i)
    B, Cc, N = features.shape
    _, M, ns = idx.shape
    out = np.empty((B, Cc, M, ns), _f)
    load().oracle_group_points(C.c_int(B), C.c_int(Cc), C.c_int(N), C.c_int(M), C.c_int(ns), _p(features), _p(idx), _p(out))
    return out


def grouping_operation_grad(grad_out, idx, N):
    """group.py:100-114 → group_points_gpu.cu:14-31 (double accumulation)."""
    grad_out, idx = _c(grad_out, _f), _c(idx, _i)
    B, Cc, M, ns = grad_out.shape
    g = np.empty((B, Cc, N), _f)
    load().oracle_group_points_grad(C.c_int(B), C.c_int(Cc), C.c_int(N), C.c_int(M), C.c_int(ns), _p(grad_out), _p(idx), _p(g))
    return g


def gather_operation(features, idx):
    """subsample.py:108-144 fwd → sampling_gpu.cu:15-31."""
    features, idx = _c(features, _f), _c(idx, _i)
    B, Cc, N = features.shape
    M = idx.shape[1]
    out = np.empty((B, Cc, M), _f)
    load().oracle_gather_points(C.c_int(B), C.c_int(Cc), C.c_int(N), C.c_int(M), _p(features), _p(idx), _p(out))
    return out


def gather_operation_grad(grad_out, idx, N):
    grad_out, idx = _c(grad_out, _f), _c(idx, _i)
    B, Cc, M = grad_out.shape
    g = np.empty((B, Cc, N), _f)
    load().oracle_gather_points_grad(C.c_int(B), C.c_int(Cc), C.c_int(N), C.c_int(M), _p(grad_out), _p(idx), _p(g))
    return g


def three_nn(unknown, known):
    """upsampling.py:11-40 → interpolate_gpu.cu:16-59.  Returns (sqrt(dist2), idx) like the Python API, plus dist2."""
    unknown, known = _c(unknown, _f), _c(known, _f)
    B, n, _ = unknown.shape
    m = known.shape[1]
    d2 = np.empty((B, n, 3), _f)
    idx = np.empty((B, n, 3), _i)
    load().oracle_three_nn(C.c_int(B), C.c_int(n), C.c_int(m), _p(unknown), _p(known), _p(d2), _p(idx))
    return np.sqrt(d2), idx, d2


def three_interpolate(features, idx, weight):
    """upsampling.py:43-67 → interpolate_gpu.cu:84-104."""
    features, idx, weight = _c(features, _f), _c(idx, _i), _c(weight, _f)
    B, Cc, m = features.shape
    n = idx.shape[1]
    out = np.empty((B, Cc, n), _f)
    load().oracle_three_interpolate(C.c_int(B), C.c_int(Cc), C.c_int(m), C.c_int(n), _p(features), _p(idx), _p(weight), _p(out))
    return out


def three_interpolate_grad(grad_out, idx, weight, m):
    """upsampling.py:69-86 → interpolate_gpu.cu:127-149."""
    grad_out, idx, weight = _c(grad_out, _f), _c(idx, _i), _c(weight, _f)
    B, Cc, n = grad_out.shape
    g = np.empty((B, Cc, m), _f)
    load().oracle_three_interpolate_grad(C.c_int(B), C.c_int(Cc), C.c_int(n), C.c_int(m), _p(grad_out), _p(idx), _p(weight), _p(g))
    return g


# --- offset API (pointops) ---------------------------------------------------------------------------
def furthestsampling(xyz, offset, new_offset, weights=None):
    """pointops/functions/pointops.py:61-98 → sampling_cuda_kernel.cu:15-129 / :175-307."""
    xyz, offset, new_offset = _c(xyz, _f), _c(offset, _i), _c(new_offset, _i)
    n, b = xyz.shape[0], offset.shape[0]
    n_max = int(np.max(np.diff(np.concatenate([[0], offset]))))
    tmp = np.full((n,), 1e10, _f)
    idx = np.zeros((int(new_offset[-1]),), _i)
    w = _c(weights, _f) if weights is not None else None
    load().oracle_furthestsampling(C.c_int(b), C.c_int(n_max), _p(xyz), _p(offset), _p(new_offset), _p(w) if w is not None else None, _p(tmp), _p(idx))
    return idx


def knnquery(nsample, xyz, new_xyz, offset, new_offset):
    """pointops/functions/pointops.py:101-116 → knnquery_cuda_kernel.cu:65-108.  Returns (idx, sqrt(dist2), dist2)."""
    xyz = _c(xyz, _f)
    new_xyz = xyz if new_xyz is None else _c(new_xyz, _f)
    offset, new_offset = _c(offset, _i), _c(new_offset, _i)
    m = new_xyz.shape[0]
    idx = np.zeros((m, nsample), _i)
    d2 = np.zeros((m, nsample), _f)
    load().oracle_knnquery(C.c_int(m), C.c_int(nsample), _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx), _p(d2))
    return idx, np.sqrt(d2), d2


def ballquery(radius, nsample, xyz, new_xyz, offset, new_offset):
    """openpoints/cpp/pointops/functions/pointops.py:53-70 → ballquery_cuda_kernel.cu:26-78."""
    xyz = _c(xyz, _f)
    new_xyz = xyz if new_xyz is None else _c(new_xyz, _f)
    offset, new_offset = _c(offset, _i), _c(new_offset, _i)
    m = new_xyz.shape[0]
    idx = np.zeros((m, nsample), _i)
    load().oracle_ballquery_offset(C.c_int(m), C.c_float(radius), C.c_int(nsample), _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx))
    return idx


def grouping(input, idx):
    """openpoints/cpp/pointops/functions/pointops.py:73-103 fwd."""
    input, idx = _c(input, _f), _c(idx, _i)
    m, ns = idx.shape
    c = input.shape[1]
    out = np.empty((m, ns, c), _f)
    load().oracle_grouping_forward(C.c_int(m), C.c_int(ns), C.c_int(c), _p(input), _p(idx), _p(out))
    return out


def grouping_grad(grad_output, idx, n):
    grad_output, idx = _c(grad_output, _f), _c(idx, _i)
    m, ns, c = grad_output.shape
    g = np.empty((n, c), _f)
    load().oracle_grouping_backward(C.c_int(m), C.c_int(ns), C.c_int(c), C.c_int(n), _p(grad_output), _p(idx), _p(g))
    return g


def interpolation_forward(input, idx, weight):
    """openpoints/cpp/pointops/functions/pointops.py:262-299 fwd core."""
    input, idx, weight = _c(input, _f), _c(idx, _i), _c(weight, _f)
    n, k = idx.shape
    c = input.shape[1]
    out = np.zeros((n, c), _f)
    load().oracle_interpolation_forward(C.c_int(n), C.c_int(c), C.c_int(k), _p(input), _p(idx), _p(weight), _p(out))
    return out


def interpolation_backward(grad_output, idx, weight, m):
    grad_output, idx, weight = _c(grad_output, _f), _c(idx, _i), _c(weight, _f)
    n, c = grad_output.shape
    k = idx.shape[1]
    g = np.empty((m, c), _f)
    load().oracle_interpolation_backward(C.c_int(n), C.c_int(c), C.c_int(k), C.c_int(m), _p(grad_output), _p(idx), _p(weight), _p(g))
    return g


def subtraction_forward(input1, input2, idx):
    input1, input2, idx = _c(input1, _f), _c(input2, _f), _c(idx, _i)
    n, ns = idx.shape
    c = input1.shape[1]
    out = np.empty((n, ns, c), _f)
    load().oracle_subtraction_forward(C.c_int(n), C.c_int(ns), C.c_int(c), _p(input1), _p(input2), _p(idx), _p(out))
    return out


def subtraction_backward(idx, grad_output):
    idx, grad_output = _c(idx, _i), _c(grad_output, _f)
    n, ns, c = grad_output.shape
    g1, g2 = np.empty((n, c), _f), np.empty((n, c), _f)
    load().oracle_subtraction_backward(C.c_int(n), C.c_int(ns), C.c_int(c), _p(idx), _p(grad_output), _p(g1), _p(g2))
    return g1, g2


def aggregation_forward(input, position, weight, idx):
    input, position, weight, idx = _c(input, _f), _c(position, _f), _c(weight, _f), _c(idx, _i)
    n, ns, c = position.shape
    w_c = weight.shape[2]
    out = np.zeros((n, c), _f)
    load().oracle_aggregation_forward(C.c_int(n), C.c_int(ns), C.c_int(c), C.c_int(w_c), _p(input), _p(position), _p(weight), _p(idx), _p(out))
    return out


def aggregation_backward(input, position, weight, idx, grad_output):
    input, position, weight, idx, grad_output = _c(input, _f), _c(position, _f), _c(weight, _f), _c(idx, _i), _c(grad_output, _f)
    n, ns, c = position.shape
    w_c = weight.shape[2]
    gi, gp, gw = np.empty((n, c), _f), np.empty((n, ns, c), _f), np.empty((n, ns, w_c), _f)
    load().oracle_aggregation_backward(C.c_int(n), C.c_int(ns), C.c_int(c), C.c_int(w_c), _p(input), _p(position), _p(weight), _p(idx), _p(grad_output), _p(gi), _p(gp), _p(gw))
    return gi, gp, gw
