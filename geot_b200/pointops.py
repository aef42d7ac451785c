"""Mirror of the reference's offset-format op API.

  pointops/functions/pointops.py:7-116               knn, fps, fps_weight, index_points, furthestsampling,
                                                     furthestsampling_weight, knnquery   (the module GeoT imports)
  openpoints/cpp/pointops/functions/pointops.py:10-299   + ballquery, grouping, interpolation, interpolation2,
                                                     subtraction, aggregation, querygroup, queryandgroup
Same names, argument order and return conventions (global int32 indices, sqrt'ed kNN distances,
int64 local indices from `knn`).  Backward passes are deterministic (scatter plan) instead of atomics.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import functional as F_


def _cum_offsets(b, n, device):
    return torch.arange(1, b + 1, device=device, dtype=torch.int32) * n


def knn(x, src, k, transpose=False):
    """pointops.py:7-21: x (b,n,3) queries, src (b,m,3) -> (idx (b,n,k) int64 LOCAL, dist (b,n,k))."""
    if transpose:
        x = x.transpose(1, 2).contiguous()
        src = src.transpose(1, 2).contiguous()
    b, n, _ = x.shape
    m = src.shape[1]
    idx, d2 = F_.knnquery(k, src.contiguous().view(-1, 3), x.contiguous().view(-1, 3), None, None, n_per=m, m_per=n)
    base = (torch.arange(b, device=idx.device, dtype=torch.int32) * m)[:, None, None]
    return (idx.view(b, n, k) - base).long(), torch.sqrt(d2).view(b, n, k)


class _FpsPrefixCache:
    """FPS is greedy from index 0, so the first k samples of a longer run ARE the k-sample run (round j only depends on
    the j-1 earlier picks).  GeoT samples the same cloud three times per forward with decreasing targets
    (transformer.py:1037-1039, downsample_targets [8192, 4096, 2048]): the second and third calls are prefixes of the
    first.  One entry: the last tensor sampled (held alive, so its storage cannot be recycled under the cache) with its
    version counter; any in-place write or a different tensor misses."""

    def __init__(self):
        self.x = None
        self.version = -1
        self.idx = None  # (b, k_max) global indices

    def get(self, x, k):
        if self.x is x and self.version == x._version and self.idx is not None and self.idx.shape[1] >= k:
            return self.idx[:, :k]
        return None

    def put(self, x, idx):
        self.x, self.version, self.idx = x, x._version, idx


_fps_cache = _FpsPrefixCache()


def fps(x, k):
    """pointops.py:24-32: x (b,n,3) -> sampled xyz (b,k,3)."""
    b, n, _ = x.shape
    idx = _fps_cache.get(x, k)
    flat = x.contiguous().view(-1, 3)
    if idx is None:
        idx = F_.furthestsampling(flat, _cum_offsets(b, n, x.device), _cum_offsets(b, k, x.device), n_max=n, m_total=b * k).long().view(b, k)
        _fps_cache.put(x, idx)
    return flat[idx.reshape(-1)].view(b, k, 3)


def fps_weight(x, k, weight=None):
    """pointops.py:34-44."""
    assert weight is not None, "the weight should be defined if using weighted fps"
    b, n, _ = x.shape
    x = x.contiguous().view(-1, 3)
    idx = F_.furthestsampling(x, _cum_offsets(b, n, x.device), _cum_offsets(b, k, x.device),
                              weights=weight.contiguous().view(-1).float(), n_max=n, m_total=b * k).long()
    return x[idx].view(b, k, 3)


def index_points(points, idx):
    """pointops.py:47-58 (pure torch)."""
    raw_size = idx.size()
    idx = idx.reshape(raw_size[0], -1)
    res = torch.gather(points, 1, idx[..., None].expand(-1, -1, points.size(-1)))
    return res.reshape(*raw_size, -1)


class FurthestSampling(Function):
    @staticmethod
    def forward(ctx, xyz, offset, new_offset):
        """xyz (n,3), offset (b), new_offset (b) -> idx (m) int32 GLOBAL."""
        assert xyz.is_contiguous()
        idx = F_.furthestsampling(xyz, offset.int(), new_offset.int())
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None


furthestsampling = FurthestSampling.apply


class FurthestSamplingWeight(Function):
    @staticmethod
    def forward(ctx, xyz, offset, new_offset, weights):
        assert xyz.is_contiguous()
        idx = F_.furthestsampling(xyz, offset.int(), new_offset.int(), weights=weights.contiguous().float())
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


furthestsampling_weight = FurthestSamplingWeight.apply


class KNNQuery(Function):
    @staticmethod
    def forward(ctx, nsample, xyz, new_xyz, offset, new_offset):
        """xyz (n,3), new_xyz (m,3) -> idx (m,nsample) int32 GLOBAL, dist (m,nsample) = sqrt(d2)."""
        if new_xyz is None:
            new_xyz = xyz
        assert xyz.is_contiguous() and new_xyz.is_contiguous()
        idx, d2 = F_.knnquery(nsample, xyz, new_xyz, offset.int(), new_offset.int())
        dist = torch.sqrt(d2)
        ctx.mark_non_differentiable(idx, dist)
        return idx, dist

    @staticmethod
    def backward(ctx, a=None, b=None):
        return None, None, None, None, None


knnquery = KNNQuery.apply


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius, nsample, xyz, new_xyz, offset, new_offset):
        """openpoints/cpp/pointops/functions/pointops.py:53-70."""
        if new_xyz is None:
            new_xyz = xyz
        assert xyz.is_contiguous() and new_xyz.is_contiguous()
        idx = F_.ballquery_offset(radius, nsample, xyz, new_xyz, offset.int(), new_offset.int())
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None, None, None


ballquery = BallQuery.apply


class Grouping(Function):
    @staticmethod
    def forward(ctx, input, idx):
        """input (n,c), idx (m,nsample) -> (m,nsample,c)."""
        assert input.is_contiguous() and idx.is_contiguous()
        ctx.n = input.shape[0]
        ctx.save_for_backward(idx)
        return F_.grouping_forward(input, idx)

    @staticmethod
    def backward(ctx, grad_output):
        idx, = ctx.saved_tensors
        plan = F_.ScatterPlan(idx, 1, ctx.n)
        return F_.grouping_backward(grad_output.contiguous(), plan, ctx.n), None


grouping = Grouping.apply


def _neighbour_rows(nsample, xyz, new_xyz, offset, new_offset, radius, query_method):
    if query_method in ('knn', 'knnquery'):
        return knnquery(nsample, xyz, new_xyz, offset, new_offset)[0]
    return ballquery(radius, nsample, xyz, new_xyz, offset, new_offset)


def querygroup(nsample, xyz, new_xyz, feat, offset, new_offset, radius=None, query_method='knn',
               normalize_dp=False, idx=None):
    """Neighbour search + relative positions + grouped features, offset format; the role of
    openpoints/cpp/pointops/functions/pointops.py:106-147.  -> (grouped_xyz (m,nsample,3), grouped_feat (m,nsample,c)).

    Built on the grouping kernel (deterministic backward) rather than fancy indexing.  Differences from the
    reference function, both supersets: a caller-supplied `idx` is used (the reference falls off the end and returns
    None in that case), and `feat=None` is accepted (the reference asserts on it before its own None check).
    normalize_dp divides by the group's largest offset norm (+1e-8) for query_method 'knn' and by `radius` otherwise."""
    new_xyz = xyz if new_xyz is None else new_xyz
    for t in (xyz, new_xyz) + (() if feat is None else (feat,)):
        if not t.is_contiguous():
            raise AssertionError("querygroup needs contiguous xyz / new_xyz / feat")
    if nsample is None:                              # group-all: every point of the batch, one group
        return xyz.transpose(1, 2).unsqueeze(2), (None if feat is None else feat.unsqueeze(2))
    if idx is None:
        idx = _neighbour_rows(nsample, xyz, new_xyz, offset, new_offset, radius, query_method)
    idx = idx.view(new_xyz.shape[0], nsample).int().contiguous()
    rel = grouping(xyz, idx) - new_xyz.unsqueeze(1)
    if normalize_dp:
        scale = rel.norm(dim=-1, keepdim=True).amax(dim=1, keepdim=True) + 1.0e-8 if query_method == 'knn' else radius
        rel = rel / scale
    return rel, (None if feat is None else grouping(feat, idx))


def queryandgroup(nsample, xyz, new_xyz, feat, idx, offset, new_offset, use_xyz=True):
    """openpoints/cpp/pointops/functions/pointops.py:151-172 -> (m, nsample, 3+c) or (m, nsample, c).

    The two row gathers go through the grouping kernel (deterministic backward) instead of fancy indexing."""
    assert xyz.is_contiguous() and new_xyz.is_contiguous() and feat.is_contiguous()
    if new_xyz is None:
        new_xyz = xyz
    if idx is None:
        idx, _ = knnquery(nsample, xyz, new_xyz, offset, new_offset)
    idx = idx.contiguous()
    grouped_xyz = grouping(xyz, idx) - new_xyz.unsqueeze(1)
    grouped_feat = grouping(feat, idx)
    if use_xyz:
        return torch.cat((grouped_xyz, grouped_feat), -1)
    return grouped_feat


class Subtraction(Function):
    @staticmethod
    def forward(ctx, input1, input2, idx):
        """input1 (n,c), input2 (n,c), idx (n,nsample) -> (n,nsample,c)."""
        assert input1.is_contiguous() and input2.is_contiguous()
        idx = idx.contiguous()
        ctx.save_for_backward(idx)
        return F_.subtraction_forward(input1, input2, idx)

    @staticmethod
    def backward(ctx, grad_output):
        idx, = ctx.saved_tensors
        plan = F_.ScatterPlan(idx, 1, idx.shape[0])
        g1, g2 = F_.subtraction_backward(grad_output.contiguous(), plan)
        return g1, g2, None


subtraction = Subtraction.apply


class Aggregation(Function):
    @staticmethod
    def forward(ctx, input, position, weight, idx):
        """input (n,c), position (n,nsample,c), weight (n,nsample,c'), idx (n,nsample) -> (n,c)."""
        assert input.is_contiguous() and position.is_contiguous() and weight.is_contiguous()
        idx = idx.contiguous()
        ctx.save_for_backward(input, position, weight, idx)
        return F_.aggregation_forward(input, position, weight, idx)

    @staticmethod
    def backward(ctx, grad_output):
        input, position, weight, idx = ctx.saved_tensors
        plan = F_.ScatterPlan(idx, 1, input.shape[0])
        gi, gp, gw = F_.aggregation_backward(input, position, weight, idx, grad_output.contiguous(), plan)
        return gi, gp, gw, None


aggregation = Aggregation.apply


def interpolation(xyz, new_xyz, feat, offset, new_offset, k=3):
    """openpoints/cpp/pointops/functions/pointops.py:245-259 (torch accumulation over k)."""
    assert xyz.is_contiguous() and new_xyz.is_contiguous() and feat.is_contiguous()
    idx, dist = knnquery(k, xyz, new_xyz, offset, new_offset)
    dist_recip = 1.0 / (dist + 1e-8)
    norm = torch.sum(dist_recip, dim=1, keepdim=True)
    weight = dist_recip / norm
    new_feat = torch.zeros((new_xyz.shape[0], feat.shape[1]), dtype=torch.float32, device=feat.device)
    for i in range(k):
        new_feat += feat[idx[:, i].long(), :] * weight[:, i].unsqueeze(-1)
    return new_feat


class Interpolation(Function):
    @staticmethod
    def forward(ctx, xyz, new_xyz, input, offset, new_offset, k=3):
        """xyz (m,3), new_xyz (n,3), input (m,c) -> (n,c): kNN(k) + inverse-distance weights + weighted gather."""
        assert xyz.is_contiguous() and new_xyz.is_contiguous() and input.is_contiguous()
        idx, d2 = F_.knnquery(k, xyz, new_xyz, offset.int(), new_offset.int())
        if k == 3:
            weight = F_.interp_weights(d2)
        else:
            dist_recip = 1.0 / (torch.sqrt(d2) + 1e-8)
            norm = torch.sum(dist_recip, dim=1, keepdim=True)
            weight = (dist_recip / norm).contiguous()
        ctx.m, ctx.k = input.shape[0], k
        ctx.save_for_backward(idx, weight)
        return F_.interpolation_forward(input, idx, weight)

    @staticmethod
    def backward(ctx, grad_output):
        idx, weight = ctx.saved_tensors
        plan = F_.ScatterPlan(idx, 1, ctx.m)
        grad_input = F_.interpolation_backward(grad_output.contiguous(), weight, plan, ctx.m, ctx.k)
        return None, None, grad_input, None, None, None


interpolation2 = Interpolation.apply
