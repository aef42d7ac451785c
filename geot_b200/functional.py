"""Tensor-level entry points: allocate outputs with torch, enqueue the sm_100a kernels through the C ABI.

One function per C-ABI entry point of include/geot_b200.h.  These are what the API mirrors
(layers.py, pointops.py, pointnet2_utils.py) and the pybind-shaped facades (backends.py) call.
No CPU path, no PyTorch re-implementation: a missing library or a CPU tensor raises.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import call, check_cuda, on_device, ptr, stream_ptr

_i32 = torch.int32
_f32 = torch.float32


def _wrote(*tensors):
    """The kernels write through raw pointers, which does not touch torch's version counter.  Caches keyed on a
    tensor's version (the scatter plan of an idx tensor, the FPS prefix cache) would not notice a caller-provided
    `out=` buffer being refilled, so those paths bump the counter explicitly."""
    for t in tensors:
        if t is not None:
            torch.autograd.graph.increment_version(t)


def _as_f32(t):
    return t if t.dtype == _f32 else t.float()


def _as_i32(t):
    return t if t.dtype == _i32 else t.int()


# --------------------------------------------------------------------------------------------------
# sampling
# --------------------------------------------------------------------------------------------------
def furthest_point_sampling(xyz, npoint, flags=0, temp=None, out=None):
    """xyz (B,N,3) fp32 -> idx (B,npoint) int32.  flags: _lib.GEOT_FPS_* (which reference kernel to match)."""
    dev = check_cuda(xyz, temp, i32=(out,))
    B, N, _ = xyz.shape
    _wrote(out)
    with on_device(dev):
        if out is None:
            out = torch.empty((B, npoint), dtype=_i32, device=dev)
        needs_temp = N > 16384 or (flags & _lib.GEOT_FPS_FORCE_GENERIC)
        if temp is None and needs_temp:
            temp = torch.full((B, N), 1e10, dtype=_f32, device=dev)
        call("geot_furthest_point_sampling", B, N, npoint, ptr(xyz), ptr(temp), ptr(out), flags, stream_ptr())
    return out


def furthestsampling(xyz, offset, new_offset, weights=None, n_max=None, m_total=None, tmp=None, out=None, flags=0):
    """Offset-format FPS: xyz (n,3), cumulative int32 offset/new_offset (b) -> GLOBAL idx (m_total)."""
    dev = check_cuda(xyz, weights, tmp, i32=(offset, new_offset, out))
    n, b = xyz.shape[0], offset.shape[0]
    _wrote(out)
    with on_device(dev):
        if n_max is None or m_total is None:
            # one host sync, like the reference's own Python (pointops.py:69-72 reads offset on the host)
            o = offset.cpu()
            seg = torch.diff(o, prepend=o.new_zeros(1))
            n_max = int(seg.max()) if n_max is None else n_max
            m_total = int(new_offset[-1]) if m_total is None else m_total
        if out is None:
            out = torch.empty((m_total,), dtype=_i32, device=dev)
        if tmp is None and (weights is not None or n_max > 16384 or (flags & _lib.GEOT_FPS_FORCE_GENERIC)):
            tmp = torch.full((n,), 1e10, dtype=_f32, device=dev)
        call("geot_furthestsampling", b, n_max, ptr(xyz), ptr(offset), ptr(new_offset), ptr(weights), ptr(tmp), ptr(out), flags, stream_ptr())
    return out


# --------------------------------------------------------------------------------------------------
# neighbour search
# --------------------------------------------------------------------------------------------------
def ball_query(new_xyz, xyz, radius, nsample, out=None, workspace=None, use_grid=True, flags=0):
    """new_xyz (B,M,3), xyz (B,N,3) -> idx (B,M,nsample) int32.  `workspace`: a uint8 tensor of at least
    geot_ball_query_workspace_bytes(B,N) bytes (allocated here when None); use_grid=False forces the
    index-order scan for every cloud (same result; testing / A-B timing)."""
    dev = check_cuda(new_xyz, xyz, i32=(out,))
    B, N, _ = xyz.shape
    M = new_xyz.shape[1]
    _wrote(out)
    with on_device(dev):
        if out is None:
            out = torch.empty((B, M, nsample), dtype=_i32, device=dev)
        ws_bytes = int(_lib.lib().geot_ball_query_workspace_bytes(B, N)) if use_grid else 0
        if ws_bytes and (workspace is None or workspace.numel() < ws_bytes):
            workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        if not ws_bytes:
            flags &= ~_lib.GEOT_BQ_PREBUILT
        call("geot_ball_query", B, N, M, float(radius), nsample, ptr(new_xyz), ptr(xyz), ptr(out),
             ptr(workspace) if ws_bytes else None, ws_bytes, flags, stream_ptr())
    return out


_side_streams = {}


def fps_ball_query(xyz, npoint, radius, nsample, fps_flags=0, bq_flags=0, idx=None, centres=None, nbr=None, workspace=None):
    """The sampling half of a PointNet++ set-abstraction level (pointnet2_modules.py: furthest_point_sample ->
    gather_operation -> QueryAndGroup's ball_query) with the ball query's binning of the cloud - which needs xyz, radius
    and nsample but not the centres - running on a side stream while FPS (one CTA per cloud, most SMs idle) computes
    them.  Returns (fps idx (B,npoint), centres (B,npoint,3), neighbour idx (B,npoint,nsample)): the values of the three
    separate calls."""
    dev = check_cuda(xyz)
    cur = torch.cuda.current_stream(dev)
    side = _side_streams.get(dev)
    if side is None:
        side = _side_streams[dev] = torch.cuda.Stream(device=dev)
    B, N, _ = xyz.shape
    ws_bytes = int(_lib.lib().geot_ball_query_workspace_bytes(B, N))
    if workspace is None or workspace.numel() < ws_bytes:
        with on_device(dev):
            workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    ready, built = torch.cuda.Event(), torch.cuda.Event()
    ready.record(cur)                      # xyz (and the workspace) as the current stream leaves them
    side.wait_event(ready)
    with torch.cuda.stream(side):
        ball_query_prepare(xyz, radius, nsample, workspace, flags=bq_flags)
        built.record(side)
    idx = furthest_point_sampling(xyz, npoint, flags=fps_flags, out=idx)
    centres = gather_rows(xyz, idx, out=centres)
    cur.wait_event(built)                  # joined again: nothing of this call is left on the side stream afterwards
    nbr = ball_query(centres, xyz, radius, nsample, out=nbr, workspace=workspace, flags=bq_flags | _lib.GEOT_BQ_PREBUILT)
    return idx, centres, nbr


def ball_query_prepare(xyz, radius, nsample, workspace=None, flags=0):
    """Bins the clouds for ball_query(.., radius, nsample) into `workspace` (allocated when None) and returns it.  The
    binning does not depend on the query centres, so it can run on another stream while they are being computed; the
    query itself is then ball_query(..., workspace=ws, flags=flags | GEOT_BQ_PREBUILT) with the same xyz."""
    dev = check_cuda(xyz)
    B, N, _ = xyz.shape
    with on_device(dev):
        ws_bytes = int(_lib.lib().geot_ball_query_workspace_bytes(B, N))
        if workspace is None or workspace.numel() < ws_bytes:
            workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        if ws_bytes:  # 0: a cloud size that never takes the binned path - nothing to prepare
            call("geot_ball_query", B, N, 0, float(radius), nsample, None, ptr(xyz), None, ptr(workspace), ws_bytes,
                 flags | _lib.GEOT_BQ_BUILD_ONLY, stream_ptr())
    return workspace


def ballquery_offset(radius, nsample, xyz, new_xyz, offset, new_offset, out=None):
    dev = check_cuda(xyz, new_xyz, i32=(offset, new_offset, out))
    m, b = new_xyz.shape[0], offset.shape[0]
    _wrote(out)
    with on_device(dev):
        if out is None:
            out = torch.empty((m, nsample), dtype=_i32, device=dev)
        call("geot_ballquery_offset", b, m, float(radius), nsample, ptr(xyz), ptr(new_xyz), ptr(offset), ptr(new_offset), ptr(out), stream_ptr())
    return out


def knnquery(nsample, xyz, new_xyz, offset=None, new_offset=None, n_per=0, m_per=None, idx=None, dist2=None, flags=0):
    """k nearest in the same segment: GLOBAL int32 idx (m,k), SQUARED dist2 (m,k), ascending.

    offset/new_offset given: offset format (ragged).  Otherwise batched: xyz is (b*n_per,3) and
    new_xyz (b*m_per,3)."""
    dev = check_cuda(xyz, new_xyz, dist2, i32=(offset, new_offset, idx))
    m = new_xyz.shape[0]
    _wrote(idx, dist2)
    with on_device(dev):
        if offset is not None:
            b = offset.shape[0]
            if m_per is None:
                m_per = m if b == 1 else int(torch.diff(new_offset, prepend=new_offset.new_zeros(1)).max())
        else:
            b = m // m_per
        if idx is None:
            idx = torch.empty((m, nsample), dtype=_i32, device=dev)
        if dist2 is None:
            dist2 = torch.empty((m, nsample), dtype=_f32, device=dev)
        n = xyz.shape[0]
        ws_bytes = int(_lib.lib().geot_knnquery_workspace_bytes(b, n, m))
        ws = torch.empty((max(ws_bytes, 1),), dtype=torch.uint8, device=dev)
        call("geot_knnquery", b, n, m, nsample, ptr(xyz), ptr(new_xyz), ptr(offset), ptr(new_offset), n_per, max(m_per, 1), ptr(idx), ptr(dist2), ptr(ws), ws_bytes, int(flags), stream_ptr())
    return idx, dist2


# --------------------------------------------------------------------------------------------------
# grouping / gather
# --------------------------------------------------------------------------------------------------
def group_points(points, idx, out=None):
    """points (B,C,N), idx (B,M,ns) -> (B,C,M,ns)."""
    dev = check_cuda(points, out, i32=(idx,))
    B, Cc, N = points.shape
    _, M, ns = idx.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, M, ns), dtype=_f32, device=dev)
        call("geot_group_points", B, Cc, N, M, ns, ptr(points), ptr(idx), ptr(out), stream_ptr())
    return out


def gather_rows(src, idx, out=None):
    """Point-major gather: src (B,N,D), idx (B,M) -> (B,M,D); D=3: new_xyz = xyz[fps_idx] without the two transposes."""
    dev = check_cuda(src, out, i32=(idx,))
    _wrote(out)
    B, N, D = src.shape
    M = idx.shape[1]
    with on_device(dev):
        if out is None:
            out = torch.empty((B, M, D), dtype=_f32, device=dev)
        call("geot_gather_rows", B, N, M, D, ptr(src), ptr(idx), ptr(out), stream_ptr())
    return out


def group_xyz_relative(xyz, new_xyz, idx, relative=True, divisor=0.0, out=None):
    """Fused xyz grouping of QueryAndGroup: xyz (B,N,3), new_xyz (B,M,3), idx (B,M,ns) -> (B,3,M,ns) =
    (xyz[idx] - new_xyz) / divisor, without the (B,3,N) transpose copy or separate subtract / divide passes."""
    dev = check_cuda(xyz, new_xyz, out, i32=(idx,))
    B, N, _ = xyz.shape
    M, ns = idx.shape[1], idx.shape[2]
    with on_device(dev):
        if out is None:
            out = torch.empty((B, 3, M, ns), dtype=_f32, device=dev)
        call("geot_group_xyz_relative", B, N, M, ns, ptr(xyz), ptr(new_xyz), ptr(idx), int(bool(relative)), float(divisor), ptr(out), stream_ptr())
    return out


def gather_points(points, idx, out=None):
    """points (B,C,N), idx (B,M) -> (B,C,M)."""
    dev = check_cuda(points, out, i32=(idx,))
    B, Cc, N = points.shape
    M = idx.shape[1]
    _wrote(out)
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, M), dtype=_f32, device=dev)
        call("geot_gather_points", B, Cc, N, M, ptr(points), ptr(idx), ptr(out), stream_ptr())
    return out


class ScatterPlan:
    """Deterministic scatter-add plan (CSR of source positions by target) for one idx tensor.

    Built lazily on the first backward that needs it and cached on the idx tensor's autograd context,
    so e.g. the xyz- and feature-grouping of one QueryAndGroup share it."""

    def __init__(self, idx, B, N, pieces=1):
        """pieces > 1 (only for group_points_grad, see grouping_plan): one CSR per (cloud, piece of the neighbour
        list) - the same build over the same idx buffer viewed as (B*pieces, E/pieces)."""
        dev = check_cuda(i32=(idx,))
        self.pieces = int(pieces)
        self.clouds = B
        E = idx.numel() // max(B, 1)
        if self.pieces > 1:
            assert E % self.pieces == 0
            B, E = B * self.pieces, E // self.pieces
        self.B, self.N = B, N
        self.E = E
        with on_device(dev):
            nbytes = int(_lib.lib().geot_scatter_plan_bytes(B, N, self.E))
            self.ws = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=dev)
            call("geot_scatter_plan_build", B, N, self.E, ptr(idx), ptr(self.ws), nbytes, stream_ptr())

    @property
    def ptr(self):
        return ptr(self.ws)

    def row_ptr(self):
        return self.ws[: self.B * (self.N + 1) * 4].view(_i32).view(self.B, self.N + 1)

    def perm(self):
        off = (self.B * (self.N + 1) * 4 + 255) // 256 * 256
        return self.ws[off: off + self.B * self.E * 4].view(_i32).view(self.B, self.E)


def grouping_plan(idx, B, N):
    """Scatter plan for group_points_grad: split into pieces when the neighbour list of a cloud (M*ns entries) is
    too long for the shared-memory backward kernel in one go (geot_scatter_plan_pieces)."""
    E = idx.numel() // max(B, 1)
    return ScatterPlan(idx, B, N, pieces=int(_lib.lib().geot_scatter_plan_pieces(N, E)))


def group_points_grad(grad_out, plan, N, out=None):
    """grad_out (B,C,M,ns) -> grad_points (B,C,N), fixed summation order."""
    dev = check_cuda(grad_out, out)
    B, Cc, M, ns = grad_out.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, N), dtype=_f32, device=dev)
        if getattr(plan, "pieces", 1) > 1:
            call("geot_group_points_grad_pieces", B, Cc, N, M, ns, plan.pieces, ptr(grad_out), plan.ptr, ptr(out), stream_ptr())
        else:
            call("geot_group_points_grad", B, Cc, N, M, ns, ptr(grad_out), plan.ptr, ptr(out), stream_ptr())
    return out


def gather_points_grad(grad_out, plan, N, out=None):
    dev = check_cuda(grad_out, out)
    B, Cc, M = grad_out.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, N), dtype=_f32, device=dev)
        call("geot_gather_points_grad", B, Cc, N, M, ptr(grad_out), plan.ptr, ptr(out), stream_ptr())
    return out


# --------------------------------------------------------------------------------------------------
# three_nn / three_interpolate
# --------------------------------------------------------------------------------------------------
def three_nn(unknown, known, dist2=None, idx=None):
    """unknown (B,n,3), known (B,m,3) -> SQUARED dist2 (B,n,3), idx (B,n,3) int32."""
    dev = check_cuda(unknown, known, dist2, i32=(idx,))
    _wrote(dist2, idx)
    B, n, _ = unknown.shape
    m = known.shape[1]
    with on_device(dev):
        if dist2 is None:
            dist2 = torch.empty((B, n, 3), dtype=_f32, device=dev)
        if idx is None:
            idx = torch.empty((B, n, 3), dtype=_i32, device=dev)
        call("geot_three_nn", B, n, m, ptr(unknown), ptr(known), ptr(dist2), ptr(idx), stream_ptr())
    return dist2, idx


def three_interpolate(points, idx, weight, out=None):
    """points (B,C,m), idx/weight (B,n,3) -> (B,C,n)."""
    dev = check_cuda(points, weight, out, i32=(idx,))
    B, Cc, m = points.shape
    n = idx.shape[1]
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, n), dtype=_f32, device=dev)
        call("geot_three_interpolate", B, Cc, m, n, ptr(points), ptr(idx), ptr(weight), ptr(out), stream_ptr())
    return out


def three_interpolate_grad(grad_out, weight, plan, m, out=None):
    dev = check_cuda(grad_out, weight, out)
    B, Cc, n = grad_out.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((B, Cc, m), dtype=_f32, device=dev)
        call("geot_three_interpolate_grad", B, Cc, n, m, ptr(grad_out), ptr(weight), plan.ptr, ptr(out), stream_ptr())
    return out


def interp_weights(dist2, want_dist=False):
    """dist2 (..., 3) SQUARED three_nn distances -> inverse-distance weights (same shape) [, sqrt distances]."""
    dev = check_cuda(dist2)
    with on_device(dev):
        weight = torch.empty_like(dist2)
        dist = torch.empty_like(dist2) if want_dist else None
        call("geot_interp_weights", dist2.numel() // 3, ptr(dist2), ptr(dist), ptr(weight), stream_ptr())
    return (weight, dist) if want_dist else weight


def interp_argmax(points, idx, weight, want_score=False):
    """points (B,C,m), idx/weight (B,n,3) -> pred (B,n) int64 = argmax_c of the 3-neighbour blend [, score (B,n)]."""
    dev = check_cuda(points, weight, i32=(idx,))
    B, Cc, m = points.shape
    n = idx.shape[1]
    with on_device(dev):
        pred = torch.empty((B, n), dtype=torch.int64, device=dev)
        score = torch.empty((B, n), dtype=_f32, device=dev) if want_score else None
        call("geot_interp_argmax", B, Cc, m, n, ptr(points), ptr(idx), ptr(weight), ptr(pred), ptr(score), stream_ptr())
    return (pred, score) if want_score else pred


def seg_counts(pred, label, num_classes):
    """pred, label (n) int64 -> counts (3C+2) int64: inter[C] | pred_cnt[C] | label_cnt[C] | correct | out-of-range."""
    dev = check_cuda(i64=(pred, label))
    with on_device(dev):
        counts = torch.empty((3 * num_classes + 2,), dtype=torch.int64, device=dev)
        call("geot_seg_counts", pred.numel(), num_classes, ptr(pred), ptr(label), ptr(counts), stream_ptr())
    return counts


# --------------------------------------------------------------------------------------------------
# offset-format (channel-last) ops
# --------------------------------------------------------------------------------------------------
def grouping_forward(input, idx, out=None):
    dev = check_cuda(input, out, i32=(idx,))
    m, ns = idx.shape
    c = input.shape[1]
    with on_device(dev):
        if out is None:
            out = torch.empty((m, ns, c), dtype=_f32, device=dev)
        call("geot_grouping_forward", m, ns, c, ptr(input), ptr(idx), ptr(out), stream_ptr())
    return out


def grouping_backward(grad_output, plan, n, out=None):
    dev = check_cuda(grad_output, out)
    m, ns, c = grad_output.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((n, c), dtype=_f32, device=dev)
        call("geot_grouping_backward", m, ns, c, n, ptr(grad_output), plan.ptr, ptr(out), stream_ptr())
    return out


def interpolation_forward(input, idx, weight, out=None):
    dev = check_cuda(input, weight, out, i32=(idx,))
    n, k = idx.shape
    c = input.shape[1]
    with on_device(dev):
        if out is None:
            out = torch.empty((n, c), dtype=_f32, device=dev)
        call("geot_interpolation_forward", n, c, k, ptr(input), ptr(idx), ptr(weight), ptr(out), stream_ptr())
    return out


def interpolation_backward(grad_output, weight, plan, m, k, out=None):
    dev = check_cuda(grad_output, weight, out)
    n, c = grad_output.shape
    with on_device(dev):
        if out is None:
            out = torch.empty((m, c), dtype=_f32, device=dev)
        call("geot_interpolation_backward", n, c, k, m, ptr(grad_output), ptr(weight), plan.ptr, ptr(out), stream_ptr())
    return out


def subtraction_forward(input1, input2, idx, out=None):
    dev = check_cuda(input1, input2, out, i32=(idx,))
    n, ns = idx.shape
    c = input1.shape[1]
    with on_device(dev):
        if out is None:
            out = torch.empty((n, ns, c), dtype=_f32, device=dev)
        call("geot_subtraction_forward", n, ns, c, ptr(input1), ptr(input2), ptr(idx), ptr(out), stream_ptr())
    return out


def subtraction_backward(grad_output, plan, g1=None, g2=None):
    dev = check_cuda(grad_output, g1, g2)
    n, ns, c = grad_output.shape
    with on_device(dev):
        if g1 is None:
            g1 = torch.empty((n, c), dtype=_f32, device=dev)
        if g2 is None:
            g2 = torch.empty((n, c), dtype=_f32, device=dev)
        call("geot_subtraction_backward", n, ns, c, ptr(grad_output), plan.ptr, ptr(g1), ptr(g2), stream_ptr())
    return g1, g2


def aggregation_forward(input, position, weight, idx, out=None):
    dev = check_cuda(input, position, weight, out, i32=(idx,))
    n, ns, c = position.shape
    w_c = weight.shape[2]
    with on_device(dev):
        if out is None:
            out = torch.empty((n, c), dtype=_f32, device=dev)
        call("geot_aggregation_forward", n, ns, c, w_c, ptr(input), ptr(position), ptr(weight), ptr(idx), ptr(out), stream_ptr())
    return out


def aggregation_backward(input, position, weight, idx, grad_output, plan, gi=None, gp=None, gw=None):
    dev = check_cuda(input, position, weight, grad_output, gi, gp, gw, i32=(idx,))
    n, ns, c = position.shape
    w_c = weight.shape[2]
    with on_device(dev):
        if gi is None:
            gi = torch.empty((n, c), dtype=_f32, device=dev)
        if gp is None:
            gp = torch.empty((n, ns, c), dtype=_f32, device=dev)
        if gw is None:
            gw = torch.empty((n, ns, w_c), dtype=_f32, device=dev)
        call("geot_aggregation_backward", n, ns, c, w_c, ptr(input), ptr(position), ptr(weight), ptr(idx), ptr(grad_output), plan.ptr, ptr(gi), ptr(gp), ptr(gw), stream_ptr())
    return gi, gp, gw


# --------------------------------------------------------------------------------------------------
# instance-dependent transition-matrix path
# --------------------------------------------------------------------------------------------------
def _ws(nbytes, dev):
    return torch.empty((max(int(nbytes), 1),), dtype=torch.uint8, device=dev)


def sigt_forward(x, W, cm=None, out=None):
    """x (B,C,N), W (C,C,2C) with cm (C,C) [sig_t_mean] or W (C*C,C) with cm None [sig_t] -> insT (B*N,C,C)."""
    dev = check_cuda(x, W, cm, out)
    B, Cc, N = x.shape
    stride = 2 * Cc if cm is not None else Cc
    with on_device(dev):
        if out is None:
            out = torch.empty((B * N, Cc, Cc), dtype=_f32, device=dev)
        nb = int(_lib.lib().geot_sigt_workspace_bytes(Cc))
        ws = _ws(nb, dev)
        call("geot_sigt_forward", B, Cc, N, ptr(x), ptr(W), stride, ptr(cm), ptr(out), ptr(ws), nb, stream_ptr())
    return out


def sigt_backward(x, W, cm, grad_insT):
    """-> (grad_W like W, g_raw (B*N, C*C) gradient w.r.t. the pre-clamp products)."""
    dev = check_cuda(x, W, cm, grad_insT)
    B, Cc, N = x.shape
    stride = 2 * Cc if cm is not None else Cc
    with on_device(dev):
        grad_W = torch.empty_like(W)
        g_raw = torch.empty((B * N, Cc * Cc), dtype=_f32, device=dev)
        nb = int(_lib.lib().geot_sigt_backward_workspace_bytes(B, Cc, N))
        ws = _ws(nb, dev)
        call("geot_sigt_backward", B, Cc, N, ptr(x), ptr(W), stride, ptr(cm), ptr(grad_insT), ptr(g_raw), ptr(grad_W), ptr(ws), nb, stream_ptr())
    return grad_W, g_raw


def transition_apply_forward(logits, ema_t, insT, lam, out=None):
    """logits (B,C,N), ema_t (C,C), insT (B*N,C,C) -> corrected logits (B,C,N)."""
    dev = check_cuda(logits, ema_t, insT, out)
    B, Cc, N = logits.shape
    with on_device(dev):
        if out is None:
            out = torch.empty_like(logits)
        call("geot_transition_apply_forward", B, Cc, N, ptr(logits), ptr(ema_t), ptr(insT), float(lam), ptr(out), stream_ptr())
    return out


def transition_apply_backward(logits, ema_t, insT, lam, grad_out, need_logits=True, need_insT=True):
    dev = check_cuda(logits, ema_t, insT, grad_out)
    B, Cc, N = logits.shape
    with on_device(dev):
        gl = torch.empty_like(logits) if need_logits else None
        gi = torch.empty_like(insT) if need_insT else None
        call("geot_transition_apply_backward", B, Cc, N, ptr(logits), ptr(ema_t), ptr(insT), float(lam), ptr(grad_out), ptr(gl), ptr(gi), stream_ptr())
    return gl, gi


MANIFOLD_3D, MANIFOLD_FEATURE = 0, 1


def spatial_order(points):
    """points (B,N,D) -> order (B,N) int32: the points of every cloud listed cell by cell of a 16^3 Morton grid over
    their first three coordinates (a processing order for the manifold kernels; not a deterministic permutation
    inside a cell, and no result depends on it)."""
    dev = check_cuda(points)
    B, N, D = points.shape
    with on_device(dev):
        order = torch.empty((B, N), dtype=_i32, device=dev)
        call("geot_spatial_order", B, N, D, ptr(points), ptr(order), stream_ptr())
    return order


def manifold_loss_forward(points, labels, insT, idx, sigma, mode, order=None):
    """points (B,N,D), labels (B,N) int64, insT (B*N,C,C), idx (B,N,k) int32 local -> (loss (1,), coef (B*N,k)).
    order: optional (B,N) int32 processing order (spatial_order)."""
    dev = check_cuda(points, insT, i32=(idx, order), i64=(labels,))
    B, N, D = points.shape
    Cc, k = insT.shape[1], idx.shape[2]
    with on_device(dev):
        coef = torch.empty((B * N, k), dtype=_f32, device=dev)
        loss = torch.empty((1,), dtype=_f32, device=dev)
        nb = int(_lib.lib().geot_manifold_workspace_bytes(B * N))
        ws = _ws(nb, dev)
        call("geot_manifold_loss_forward", B, N, D, Cc, k, ptr(points), ptr(labels), ptr(insT), ptr(idx), ptr(order), float(sigma), mode, ptr(coef), ptr(loss), ptr(ws), nb, stream_ptr())
    return loss, coef


def manifold_loss_backward(insT, idx, coef, plan, grad_loss, order=None):
    dev = check_cuda(insT, coef, grad_loss, i32=(idx, order))
    B, N, k = idx.shape
    Cc = insT.shape[1]
    with on_device(dev):
        g = torch.empty_like(insT)
        call("geot_manifold_loss_backward", B, N, Cc, k, ptr(insT), ptr(idx), ptr(coef), plan.ptr, ptr(order), ptr(grad_loss), ptr(g), stream_ptr())
    return g


def class_pick(pred_u):
    """pred_u (B,C,N) softmax -> class_T (C,C), best_val (C), best_idx (C) int64 (flat b*N+n)."""
    dev = check_cuda(pred_u)
    B, Cc, N = pred_u.shape
    with on_device(dev):
        class_T = torch.empty((Cc, Cc), dtype=_f32, device=dev)
        best_v = torch.empty((Cc,), dtype=_f32, device=dev)
        best_i = torch.empty((Cc,), dtype=torch.int64, device=dev)
        nb = int(_lib.lib().geot_class_pick_workspace_bytes(Cc))
        ws = _ws(nb, dev)
        call("geot_class_pick", B, Cc, N, ptr(pred_u), ptr(class_T), ptr(best_v), ptr(best_i), ptr(ws), nb, stream_ptr())
    return class_T, best_v, best_i


def class_t_step(class_T, ema_t, sigma, label_proj, geo_lambda, ema_decay):
    """-> (ema_t_corr, new ema_t); label_proj int32 (C) on the device."""
    dev = check_cuda(class_T, ema_t, sigma, i32=(label_proj,))
    Cc = class_T.shape[0]
    with on_device(dev):
        corr = torch.empty_like(class_T)
        ema_out = torch.empty_like(class_T)
        call("geot_class_t_step", Cc, ptr(class_T), ptr(ema_t), ptr(sigma), ptr(label_proj), float(geo_lambda), float(ema_decay), ptr(corr), ptr(ema_out), stream_ptr())
    return corr, ema_out


def mean_feature_update(probs, target, cm, c_num):
    """In-place running class-mean update (train.py:884-895 semantics); probs (BN,C), target (BN) int64."""
    dev = check_cuda(probs, cm, c_num, i64=(target,))
    BN, Cc = probs.shape
    with on_device(dev):
        nb = max(int(_lib.lib().geot_class_pick_workspace_bytes(Cc)), 512)
        ws = _ws(nb, dev)
        call("geot_mean_feature_update", BN, Cc, ptr(probs), ptr(target), ptr(cm), ptr(c_num), ptr(ws), nb, stream_ptr())
    return cm, c_num


# --------------------------------------------------------------------------------------------------
# unsupervised poly-1 focal loss (Poly1FocalLoss_U_corr)
# --------------------------------------------------------------------------------------------------
def poly1_focal_forward(logits, labels, conf, mask, thresh, epsilon, alpha, gamma):
    """logits (B,C,N) f32, labels (B,N) int64, conf (B,N) f32 or None, mask (B,N) f32 or None -> (loss (1,), denom (1,))."""
    dev = check_cuda(logits, conf, mask, i64=(labels,))
    B, Cc, N = logits.shape
    with on_device(dev):
        loss = torch.empty((1,), dtype=_f32, device=dev)
        denom = torch.empty((1,), dtype=_f32, device=dev)
        nb = int(_lib.lib().geot_poly1_focal_workspace_bytes(B * N))
        ws = _ws(nb, dev)
        call("geot_poly1_focal_forward", B, Cc, N, ptr(logits), ptr(labels), ptr(conf), ptr(mask), float(thresh), float(epsilon),
             float(alpha), float(gamma), ptr(loss), ptr(denom), ptr(ws), nb, stream_ptr())
    return loss, denom


def poly1_focal_backward(logits, labels, conf, mask, thresh, epsilon, alpha, gamma, denom, grad_loss):
    dev = check_cuda(logits, conf, mask, denom, grad_loss, i64=(labels,))
    B, Cc, N = logits.shape
    with on_device(dev):
        g = torch.empty_like(logits)
        call("geot_poly1_focal_backward", B, Cc, N, ptr(logits), ptr(labels), ptr(conf), ptr(mask), float(thresh), float(epsilon),
             float(alpha), float(gamma), ptr(denom), ptr(grad_loss), ptr(g), stream_ptr())
    return g


def poly1_focal_t_forward(logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma):
    """Poly1FocalLoss_U_T: as poly1_focal_forward with every point scaled by conf / pred_t[b, label, n]."""
    dev = check_cuda(logits, conf, mask, pred_t, i64=(labels,))
    B, Cc, N = logits.shape
    with on_device(dev):
        loss = torch.empty((1,), dtype=_f32, device=dev)
        denom = torch.empty((1,), dtype=_f32, device=dev)
        nb = int(_lib.lib().geot_poly1_focal_workspace_bytes(B * N))
        ws = _ws(nb, dev)
        call("geot_poly1_focal_t_forward", B, Cc, N, ptr(logits), ptr(labels), ptr(conf), ptr(mask), ptr(pred_t), float(thresh),
             float(epsilon), float(alpha), float(gamma), ptr(loss), ptr(denom), ptr(ws), nb, stream_ptr())
    return loss, denom


def poly1_focal_t_backward(logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma, denom, grad_loss):
    dev = check_cuda(logits, conf, mask, pred_t, denom, grad_loss, i64=(labels,))
    B, Cc, N = logits.shape
    with on_device(dev):
        g = torch.empty_like(logits)
        gt = torch.empty_like(pred_t)
        call("geot_poly1_focal_t_backward", B, Cc, N, ptr(logits), ptr(labels), ptr(conf), ptr(mask), ptr(pred_t), float(thresh),
             float(epsilon), float(alpha), float(gamma), ptr(denom), ptr(grad_loss), ptr(g), ptr(gt), stream_ptr())
    return g, gt


def graph_feature(x_k, x_q, idx):
    """x_k (B,C,Nk), x_q (B,C,Nq), idx (B,Nq,k) int32 -> (B,2C,Nq,k) = cat(x_k[.., idx] - x_q, x_q)."""
    dev = check_cuda(x_k, x_q, i32=(idx,))
    B, Cc, Nk = x_k.shape
    Nq, k = idx.shape[1], idx.shape[2]
    with on_device(dev):
        out = torch.empty((B, 2 * Cc, Nq, k), dtype=_f32, device=dev)
        call("geot_graph_feature", B, Cc, Nk, Nq, k, ptr(x_k), ptr(x_q), ptr(idx), ptr(out), stream_ptr())
    return out


def graph_feature_grad_q(grad_out, Cc):
    """grad_out (B,2C,Nq,k) -> grad_x_q (B,C,Nq)."""
    dev = check_cuda(grad_out)
    B, _, Nq, k = grad_out.shape
    with on_device(dev):
        g = torch.empty((B, Cc, Nq), dtype=_f32, device=dev)
        call("geot_graph_feature_grad_q", B, Cc, Nq, k, ptr(grad_out), ptr(g), stream_ptr())
    return g
