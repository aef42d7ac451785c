"""ctypes front-end of oracle/_ref/*.so — the reference's OWN CUDA kernels, compiled unchanged from
/root/reference by oracle/ref_build/Makefile.  Test infrastructure: used by the -m gpu tests as the
strongest parity pin (and by bench.py to time the kernels we replace).  Operates on torch CUDA tensors;
the reference launches on the legacy default stream, so every call is followed by a device sync."""
import ctypes as C
import os

import torch

_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
_libs = {}


def available(name="pn2batch"):
    return os.path.exists(os.path.join(_DIR, f"libref_{name}.so"))


def lib(name):
    if name not in _libs:
        _libs[name] = C.CDLL(os.path.join(_DIR, f"libref_{name}.so"))
    return _libs[name]


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _sync(name):
    torch.cuda.current_stream().synchronize()
    rc = lib(name).ref_sync()
    assert rc == 0, f"reference kernel failed: cudaError {rc}"


def _pre():
    torch.cuda.synchronize()


def fps(xyz, M, which="pn2batch"):
    B, N, _ = xyz.shape
    temp = torch.full((B, N), 1e10, device=xyz.device)
    idx = torch.zeros((B, M), dtype=torch.int32, device=xyz.device)
    _pre(); lib(which).ref_furthest_point_sampling(B, N, M, _p(xyz), _p(temp), _p(idx)); _sync(which)
    return idx


def ball_query(radius, nsample, xyz, new_xyz, which="pn2batch"):
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    idx = torch.zeros((B, M, nsample), dtype=torch.int32, device=xyz.device)
    _pre(); lib(which).ref_ball_query(B, N, M, C.c_float(radius), nsample, _p(new_xyz), _p(xyz), _p(idx)); _sync(which)
    return idx


def group_points(points, idx, which="pn2batch"):
    B, Cc, N = points.shape
    _, M, ns = idx.shape
    out = torch.empty((B, Cc, M, ns), device=points.device)
    _pre(); lib(which).ref_group_points(B, Cc, N, M, ns, _p(points), _p(idx), _p(out)); _sync(which)
    return out


def group_points_grad(grad_out, idx, N, which="pn2batch"):
    B, Cc, M, ns = grad_out.shape
    g = torch.zeros((B, Cc, N), device=grad_out.device)
    _pre(); lib(which).ref_group_points_grad(B, Cc, N, M, ns, _p(grad_out), _p(idx), _p(g)); _sync(which)
    return g


def gather_points(points, idx, which="pn2batch"):
    B, Cc, N = points.shape
    M = idx.shape[1]
    out = torch.empty((B, Cc, M), device=points.device)
    _pre(); lib(which).ref_gather_points(B, Cc, N, M, _p(points), _p(idx), _p(out)); _sync(which)
    return out


def gather_points_grad(grad_out, idx, N, which="pn2batch"):
    B, Cc, M = grad_out.shape
    g = torch.zeros((B, Cc, N), device=grad_out.device)
    _pre(); lib(which).ref_gather_points_grad(B, Cc, N, M, _p(grad_out), _p(idx), _p(g)); _sync(which)
    return g


def three_nn(unknown, known, which="pn2batch"):
    B, n, _ = unknown.shape
    m = known.shape[1]
    d2 = torch.empty((B, n, 3), device=unknown.device)
    idx = torch.empty((B, n, 3), dtype=torch.int32, device=unknown.device)
    _pre(); lib(which).ref_three_nn(B, n, m, _p(unknown), _p(known), _p(d2), _p(idx)); _sync(which)
    return d2, idx


def three_interpolate(points, idx, weight, which="pn2batch"):
    B, Cc, m = points.shape
    n = idx.shape[1]
    out = torch.empty((B, Cc, n), device=points.device)
    _pre(); lib(which).ref_three_interpolate(B, Cc, m, n, _p(points), _p(idx), _p(weight), _p(out)); _sync(which)
    return out


def three_interpolate_grad(grad_out, idx, weight, m, which="pn2batch"):
    B, Cc, n = grad_out.shape
    g = torch.zeros((B, Cc, m), device=grad_out.device)
    _pre(); lib(which).ref_three_interpolate_grad(B, Cc, n, m, _p(grad_out), _p(idx), _p(weight), _p(g)); _sync(which)
    return g


def knnquery(nsample, xyz, new_xyz, offset, new_offset, which="pointops"):
    m = new_xyz.shape[0]
    idx = torch.zeros((m, nsample), dtype=torch.int32, device=xyz.device)
    d2 = torch.zeros((m, nsample), device=xyz.device)
    _pre(); lib(which).ref_knnquery(m, nsample, _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx), _p(d2)); _sync(which)
    return idx, d2


def furthestsampling(xyz, offset, new_offset, weights=None, which="pointops"):
    n, b = xyz.shape[0], offset.shape[0]
    o = offset.cpu()
    n_max = int(torch.diff(o, prepend=o.new_zeros(1)).max())
    idx = torch.zeros((int(new_offset[-1]),), dtype=torch.int32, device=xyz.device)
    tmp = torch.full((n,), 1e10, device=xyz.device)
    _pre()
    if weights is None:
        lib(which).ref_furthestsampling(b, n_max, _p(xyz), _p(offset), _p(new_offset), _p(tmp), _p(idx))
    else:
        lib(which).ref_furthestsampling_weights(b, n_max, _p(xyz), _p(offset), _p(new_offset), _p(weights), _p(tmp), _p(idx))
    _sync(which)
    return idx


def ballquery_offset(radius, nsample, xyz, new_xyz, offset, new_offset):
    m = new_xyz.shape[0]
    idx = torch.zeros((m, nsample), dtype=torch.int32, device=xyz.device)
    _pre(); lib("pointops_full").ref_ballquery(m, C.c_float(radius), nsample, _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx)); _sync("pointops_full")
    return idx


def grouping_forward(input, idx):
    m, ns = idx.shape
    c = input.shape[1]
    out = torch.empty((m, ns, c), device=input.device)
    _pre(); lib("pointops_full").ref_grouping_forward(m, ns, c, _p(input), _p(idx), _p(out)); _sync("pointops_full")
    return out


def grouping_backward(grad_output, idx, n):
    m, ns, c = grad_output.shape
    g = torch.zeros((n, c), device=grad_output.device)
    _pre(); lib("pointops_full").ref_grouping_backward(m, ns, c, _p(grad_output), _p(idx), _p(g)); _sync("pointops_full")
    return g


def interpolation_forward(input, idx, weight):
    n, k = idx.shape
    c = input.shape[1]
    out = torch.zeros((n, c), device=input.device)
    _pre(); lib("pointops_full").ref_interpolation_forward(n, c, k, _p(input), _p(idx), _p(weight), _p(out)); _sync("pointops_full")
    return out


def interpolation_backward(grad_output, idx, weight, m):
    n, c = grad_output.shape
    k = idx.shape[1]
    g = torch.zeros((m, c), device=grad_output.device)
    _pre(); lib("pointops_full").ref_interpolation_backward(n, c, k, _p(grad_output), _p(idx), _p(weight), _p(g)); _sync("pointops_full")
    return g


def subtraction_forward(in1, in2, idx):
    n, ns = idx.shape
    c = in1.shape[1]
    out = torch.zeros((n, ns, c), device=in1.device)
    _pre(); lib("pointops_full").ref_subtraction_forward(n, ns, c, _p(in1), _p(in2), _p(idx), _p(out)); _sync("pointops_full")
    return out


def subtraction_backward(idx, grad_output):
    n, ns, c = grad_output.shape
    g1 = torch.zeros((n, c), device=grad_output.device)
    g2 = torch.zeros((n, c), device=grad_output.device)
    _pre(); lib("pointops_full").ref_subtraction_backward(n, ns, c, _p(idx), _p(grad_output), _p(g1), _p(g2)); _sync("pointops_full")
    return g1, g2


def aggregation_forward(input, position, weight, idx):
    n, ns, c = position.shape
    w_c = weight.shape[2]
    out = torch.zeros((n, c), device=input.device)
    _pre(); lib("pointops_full").ref_aggregation_forward(n, ns, c, w_c, _p(input), _p(position), _p(weight), _p(idx), _p(out)); _sync("pointops_full")
    return out


def aggregation_backward(input, position, weight, idx, grad_output):
    n, ns, c = position.shape
    w_c = weight.shape[2]
    gi = torch.zeros((n, c), device=input.device)
    gp = torch.zeros((n, ns, c), device=input.device)
    gw = torch.zeros((n, ns, w_c), device=input.device)
    _pre(); lib("pointops_full").ref_aggregation_backward(n, ns, c, w_c, _p(input), _p(position), _p(weight), _p(idx), _p(grad_output), _p(gi), _p(gp), _p(gw)); _sync("pointops_full")
    return gi, gp, gw
