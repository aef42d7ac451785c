"""Mirror of the reference's batched op API (openpoints/models/layers/{subsample,group,upsampling,knn}.py).

Same names, argument order, dtypes, shapes and autograd contracts as the reference so that code
written against `openpoints.models.layers` runs unchanged on top of the sm_100a kernels:
  furthest_point_sample  subsample.py:76-105      gather_operation   subsample.py:108-144 / group.py:140-174
  ball_query             group.py:177-203         grouping_operation group.py:76-117
  three_nn               upsampling.py:11-40      three_interpolate  upsampling.py:43-89
  three_interpolation    upsampling.py:92-102     knn_point / KNN    knn.py:7-61, group.py:12-28
  QueryAndGroup / KNNGroup / GroupAll / create_grouper   group.py:206-352
Differences, all deliberate: outputs are allocated on the inputs' device (the reference uses the
current device, subsample.py:93-94); backward passes are deterministic; CPU tensors raise.
"""
from __future__ import annotations

import copy
from typing import Tuple

import torch
import torch.nn as nn
from torch.autograd import Function

from . import _lib
from . import functional as F_


class FurthestPointSampling(Function):
    @staticmethod
    def forward(ctx, xyz: torch.Tensor, npoint: int) -> torch.Tensor:
        """xyz (B,N,3) -> idx (B,npoint) int32; same indices as furthest_point_sampling_wrapper."""
        assert xyz.is_contiguous()
        out = F_.furthest_point_sampling(xyz, npoint, flags=0)
        ctx.mark_non_differentiable(out)
        return out

    @staticmethod
    def backward(ctx, a=None):
        return None, None


furthest_point_sample = FurthestPointSampling.apply


class GatherOperation(Function):
    @staticmethod
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B,C,N), idx (B,npoint) int32 -> (B,C,npoint)."""
        assert features.is_contiguous()
        assert idx.is_contiguous()
        _, C, N = features.size()
        ctx.for_backwards = (idx, C, N)
        return F_.gather_points(features, idx)

    @staticmethod
    def backward(ctx, grad_out):
        idx, C, N = ctx.for_backwards
        plan = F_.ScatterPlan(idx, idx.shape[0], N)
        return F_.gather_points_grad(grad_out.contiguous(), plan, N), None


gather_operation = GatherOperation.apply


def fps(data, number):
    """subsample.py:148-156: data (B,N,C) -> sampled rows (B,number,C)."""
    fps_idx = furthest_point_sample(data[:, :, :3].contiguous(), number)
    return torch.gather(data, 1, fps_idx.unsqueeze(-1).long().expand(-1, -1, data.shape[-1]))


class _PlanCache:
    """One scatter plan per idx tensor, shared by every grouping_operation that used it.  The plan is only valid for
    the CONTENT it was built from: it remembers the tensor's version counter (every in-place write - torch ops and the
    library's own `out=` paths, which bump it explicitly - changes it) and storage address, and is rebuilt otherwise."""

    def __init__(self):
        self.plan = None
        self.version = -1
        self.data_ptr = 0

    def get(self, idx, N, version_at_forward):
        if idx._version != version_at_forward:
            raise RuntimeError("grouping_operation: idx was modified in place between forward and backward")
        if self.plan is None or self.plan.N != N or self.version != idx._version or self.data_ptr != idx.data_ptr():
            self.plan = F_.grouping_plan(idx, idx.shape[0], N)
            self.version, self.data_ptr = idx._version, idx.data_ptr()
        return self.plan


class GroupingOperation(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """features (B,C,N), idx (B,npoint,nsample) int32 -> (B,C,npoint,nsample)."""
        assert features.is_contiguous()
        assert idx.is_contiguous()
        N = features.size(2)
        cache = getattr(idx, "_geot_plan_cache", None)
        if cache is None:
            cache = _PlanCache()
            try:
                idx._geot_plan_cache = cache
            except Exception:  # tensor subclasses that refuse attributes: no sharing, still correct
                pass
        ctx.for_backwards = (idx, N, cache, idx._version)
        return F_.group_points(features, idx)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        idx, N, cache, version = ctx.for_backwards
        return F_.group_points_grad(grad_out.contiguous(), cache.get(idx, N, version), N), None


grouping_operation = GroupingOperation.apply


def torch_grouping_operation(features, idx):
    """group.py:120-137 (pure torch, kept because utils/insT_loss.py imports the name)."""
    all_idx = idx.reshape(idx.shape[0], -1)
    all_idx = all_idx.unsqueeze(1).repeat(1, features.shape[1], 1)
    grouped_features = features.gather(2, all_idx)
    return grouped_features.reshape(idx.shape[0], features.shape[1], idx.shape[1], idx.shape[2])


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius: float, nsample: int, xyz: torch.Tensor, new_xyz: torch.Tensor) -> torch.Tensor:
        """xyz (B,N,3) points, new_xyz (B,npoint,3) ball centres -> idx (B,npoint,nsample) int32."""
        assert new_xyz.is_contiguous()
        assert xyz.is_contiguous()
        idx = F_.ball_query(new_xyz, xyz, radius, nsample)
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


ball_query = BallQuery.apply


class ThreeNN(Function):
    @staticmethod
    def forward(ctx, unknown: torch.Tensor, known: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """unknown (B,n,3), known (B,m,3) -> (dist (B,n,3) = sqrt(d2), idx (B,n,3) int32)."""
        assert unknown.is_contiguous()
        assert known.is_contiguous()
        dist2, idx = F_.three_nn(unknown, known)
        dist = torch.sqrt(dist2)
        ctx.mark_non_differentiable(dist, idx)
        return dist, idx

    @staticmethod
    def backward(ctx, a=None, b=None):
        return None, None


three_nn = ThreeNN.apply


class ThreeInterpolate(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, features: torch.Tensor, idx: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """features (B,C,m), idx/weight (B,n,3) -> (B,C,n)."""
        assert features.is_contiguous()
        assert idx.is_contiguous()
        assert weight.is_contiguous()
        m = features.size(2)
        ctx.three_interpolate_for_backward = (idx, weight, m)
        return F_.three_interpolate(features, idx, weight)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        idx, weight, m = ctx.three_interpolate_for_backward
        plan = F_.ScatterPlan(idx, idx.shape[0], m)
        return F_.three_interpolate_grad(grad_out.contiguous(), weight, plan, m), None, None


three_interpolate = ThreeInterpolate.apply


def three_interpolation(unknown_xyz, known_xyz, know_feat):
    """upsampling.py:92-102: three_nn -> inverse-distance weights -> three_interpolate."""
    assert unknown_xyz.is_contiguous()
    assert known_xyz.is_contiguous()
    with torch.no_grad():  # dist / idx are non-differentiable in the reference too (upsampling.py:38)
        dist2, idx = F_.three_nn(unknown_xyz, known_xyz)
        weight = F_.interp_weights(dist2)  # sqrt, 1/(d + 1e-8), sum, divide: one kernel instead of five
    return three_interpolate(know_feat, idx, weight)


# --------------------------------------------------------------------------------------------------
# kNN (knn.py:7-61, group.py:12-28).  The reference builds a dense (B,M,N) torch.cdist matrix and
# runs topk over it (2 GB at B=2, N=16000).  For 3-D points the sm_100a kNN kernel gives the same
# neighbours without the matrix; distances are sqrt(d2) of the kernel's fp32 d2 (cdist's matmul
# expansion differs in the last bits, and topk's tie order is unspecified — SURVEY.md §7.3-8 — so the
# contract is "same neighbour sets / loss values", not bit-equal indices).  Other feature widths keep
# the reference's own torch expression (it is pure torch in the reference too).
# --------------------------------------------------------------------------------------------------
def _knn3d(k, query, support):
    B, M, _ = query.shape
    N = support.shape[1]
    q = query.contiguous().float().view(-1, 3)
    s = support.contiguous().float().view(-1, 3)
    idx, d2 = F_.knnquery(k, s, q, None, None, n_per=N, m_per=M, flags=_lib.GEOT_KNN_TIES_BY_INDEX)
    base = (torch.arange(B, device=idx.device, dtype=torch.int32) * N).view(B, 1, 1)
    return torch.sqrt(d2).view(B, M, k), (idx.view(B, M, k) - base)


@torch.no_grad()
def knn_point(k, query, support=None):
    """knn.py:7-20 -> (dist (B,M,k), idx (B,M,k) int64), ascending."""
    if support is None:
        support = query
    if query.shape[-1] == 3 and query.is_cuda and k <= 127:
        dist, idx = _knn3d(k, query, support)
        return dist, idx.long()
    dist = torch.cdist(query, support)
    k_dist = dist.topk(k=k, dim=-1, largest=False, sorted=True)
    return k_dist.values, k_dist.indices


class KNN(nn.Module):
    """knn.py:23-61 (query, support) -> (dist, idx int32 (B,M,k)).  `farthest=True` keeps the torch path."""

    def __init__(self, neighbors, farthest=False, sorted=True, **kwargs):
        super().__init__()
        self.neighbors = neighbors
        self.farthest = farthest
        self.sorted = sorted

    @torch.no_grad()
    def forward(self, query, support=None):
        if support is None:
            support = query
        if not self.farthest and query.shape[-1] == 3 and query.is_cuda and self.neighbors <= 127:
            return _knn3d(self.neighbors, query, support)
        dist = torch.cdist(query, support)
        k_dist = dist.topk(k=self.neighbors, dim=-1, largest=self.farthest, sorted=self.sorted)
        return k_dist.values, k_dist.indices.int()


class GroupKNN(nn.Module):
    """group.py:12-28 `KNN(neighbors)(support, query)` -> (dist (B,k,M)-shaped values, idx (B,M,k) int32)."""

    def __init__(self, neighbors, transpose_mode=True):
        super().__init__()
        self.neighbors = neighbors

    @torch.no_grad()
    def forward(self, support, query):
        if query.shape[-1] == 3 and query.is_cuda and self.neighbors <= 127:
            dist, idx = _knn3d(self.neighbors, query, support)
            return dist.transpose(1, 2).contiguous(), idx.contiguous()
        dist = torch.cdist(support, query)
        k_dist = dist.topk(k=self.neighbors, dim=1, largest=False)
        return k_dist.values, k_dist.indices.transpose(1, 2).contiguous().int()


def _fusable(query_xyz, support_xyz):
    """The fused xyz grouping has no backward: used when neither position tensor takes part in autograd."""
    return (support_xyz.is_cuda and support_xyz.dtype == torch.float32 and query_xyz.dtype == torch.float32
            and support_xyz.is_contiguous() and query_xyz.is_contiguous()
            and not (torch.is_grad_enabled() and (support_xyz.requires_grad or query_xyz.requires_grad)))


class _Grouper(nn.Module):
    """Shared body of the two neighbourhood groupers: neighbour search (subclass) -> positions of the neighbours,
    relative to their centre when `relative_xyz` -> optional scale -> grouped features.  With positions outside
    autograd (the usual case) the position part is ONE kernel reading the (B,N,3) cloud directly; otherwise it is
    grouping_operation on the transposed cloud, which has the deterministic backward."""

    relative_xyz = True
    return_only_idx = False

    def _neighbours(self, query_xyz, support_xyz):
        raise NotImplementedError

    def _fused_divisor(self):
        return 0.0

    def _rescale(self, grouped_xyz, fused):
        return grouped_xyz

    def forward(self, query_xyz, support_xyz, features=None):
        idx = self._neighbours(query_xyz, support_xyz)
        if self.return_only_idx:
            return idx
        idx = idx.int().contiguous()
        fused = _fusable(query_xyz, support_xyz)
        if fused:
            grouped_xyz = F_.group_xyz_relative(support_xyz, query_xyz, idx, relative=self.relative_xyz,
                                                divisor=self._fused_divisor())
        else:
            grouped_xyz = grouping_operation(support_xyz.transpose(1, 2).contiguous(), idx)   # (B,3,npoint,nsample)
            if self.relative_xyz:
                grouped_xyz = grouped_xyz - query_xyz.transpose(1, 2).unsqueeze(-1)
        grouped_xyz = self._rescale(grouped_xyz, fused)
        return grouped_xyz, (None if features is None else grouping_operation(features, idx))


class QueryAndGroup(_Grouper):
    """Ball-query grouper, the role of openpoints/models/layers/group.py:206-255.  normalize_dp divides the
    relative positions by the radius (inside the fused kernel when that path is taken)."""

    def __init__(self, radius: float, nsample: int, relative_xyz=True, normalize_dp=False,
                 normalize_by_std=False, normalize_by_allstd=False, normalize_by_allstd2=False,
                 return_only_idx=False, **kwargs):
        super().__init__()
        if int(bool(normalize_dp)) + int(bool(normalize_by_std)) + int(bool(normalize_by_allstd)) > 1:
            raise AssertionError("at most one of normalize_dp / normalize_by_std / normalize_by_allstd")
        self.radius, self.nsample = radius, nsample
        self.relative_xyz, self.return_only_idx = relative_xyz, return_only_idx
        self.normalize_dp = normalize_dp
        # accepted and stored like the reference's constructor; its forward never reads them either
        self.normalize_by_std, self.normalize_by_allstd, self.normalize_by_allstd2 = \
            normalize_by_std, normalize_by_allstd, normalize_by_allstd2

    def _neighbours(self, query_xyz, support_xyz):
        return ball_query(self.radius, self.nsample, support_xyz, query_xyz)

    def _fused_divisor(self):
        return self.radius if (self.relative_xyz and self.normalize_dp) else 0.0

    def _rescale(self, grouped_xyz, fused):
        if fused or not (self.relative_xyz and self.normalize_dp):
            return grouped_xyz
        return grouped_xyz / self.radius


class GroupAll(nn.Module):
    """One group holding the whole cloud (group.py:258-272): (B,3,1,N) positions, (B,C,1,N) features."""

    def forward(self, new_xyz, xyz, features=None):
        return xyz.transpose(1, 2).unsqueeze(2), (None if features is None else features.unsqueeze(2))


class KNNGroup(_Grouper):
    """kNN grouper, the role of group.py:275-320.  normalize_dp divides by each cloud's largest neighbour offset."""

    def __init__(self, nsample: int, relative_xyz=True, normalize_dp=False, return_only_idx=False, **kwargs):
        super().__init__()
        self.nsample = nsample
        self.knn = GroupKNN(nsample, transpose_mode=True)
        self.relative_xyz, self.normalize_dp, self.return_only_idx = relative_xyz, normalize_dp, return_only_idx

    def _neighbours(self, query_xyz, support_xyz):
        return self.knn(support_xyz, query_xyz)[1]

    def _rescale(self, grouped_xyz, fused):
        if not self.normalize_dp:
            return grouped_xyz
        reach = grouped_xyz.square().sum(dim=1).amax(dim=(1, 2)).sqrt()
        return grouped_xyz / reach.view(-1, 1, 1, 1)


# --------------------------------------------------------------------------------------------------
# DGCNN edge features (transformer.py:343-364)
# --------------------------------------------------------------------------------------------------
class GraphFeature(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, x_k: torch.Tensor, x_q: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """x_k (B,C,Nk), x_q (B,C,Nq), idx (B,Nq,k) int32 -> (B,2C,Nq,k) = cat(x_k[.., idx] - x_q, x_q)."""
        x_k, x_q = x_k.contiguous(), x_q.contiguous()
        assert idx.is_contiguous() and idx.dtype == torch.int32
        ctx.for_backwards = (idx, x_k.shape[1], x_k.shape[2])
        return F_.graph_feature(x_k, x_q, idx)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        idx, C, Nk = ctx.for_backwards
        grad_out = grad_out.contiguous()
        g_k = g_q = None
        if ctx.needs_input_grad[0]:  # deterministic scatter over the same plan machinery as grouping_operation
            plan = F_.grouping_plan(idx, idx.shape[0], Nk)
            g_k = F_.group_points_grad(grad_out[:, :C].contiguous(), plan, Nk)
        if ctx.needs_input_grad[1]:
            g_q = F_.graph_feature_grad_q(grad_out, C)
        return g_k, g_q, None


def knn_neighbours(k, coor_q, coor_k):
    """coor_q (B,3,Nq), coor_k (B,3,Nk) -> idx (B,Nq,k) int32, cloud-local, ascending distance (the neighbours
    knn_cuda.KNN(k, transpose_mode=False)(coor_k, coor_q) returns, in the layout the kNN kernel writes)."""
    with torch.no_grad():
        ref = coor_k.transpose(1, 2).contiguous().float()
        query = coor_q.transpose(1, 2).contiguous().float()
        B, Nk, _ = ref.shape
        Nq = query.shape[1]
        idx, _ = F_.knnquery(k, ref.view(-1, 3), query.view(-1, 3), None, None, n_per=Nk, m_per=Nq, flags=_lib.GEOT_KNN_TIES_BY_INDEX)
        base = (torch.arange(B, device=idx.device, dtype=torch.int32) * Nk).view(B, 1, 1)
        return (idx.view(B, Nq, k) - base).contiguous()


def get_graph_feature(coor_q, x_q, coor_k, x_k, k=4, idx=None):
    """DGCNN_Propagation.get_graph_feature (transformer.py:343-364) as a function: coor (B,3,N), x (B,C,N) ->
    (B,2C,Nq,k).  kNN, gather, `feature - x_q` and the concat are two kernels; no (B,Nk,C) transpose, no
    (B,C,Nq,k) intermediates.  `idx` (B,Nq,k) int32 may be passed when several calls share the neighbours."""
    if idx is None:
        idx = knn_neighbours(k, coor_q, coor_k)
    return GraphFeature.apply(x_k, x_q, idx)


class Group(nn.Module):
    """transformer.py:274-303 (`Group`: FPS + kNN patches of the Point-Transformer front end; same constructor and
    `forward(xyz) -> (neighborhood (B,G,M,3), center (B,G,3), idx (B*G*M) int64 into the flattened batch)`).
    The reference runs FPS, two transposes around gather_operation, knn_cuda, an index-base add, a fancy-index row gather
    and a subtraction; here: FPS (the 512-thread / |p|^2 skip variant of pointnet2/_ext_src, transformer.py:271), two
    point-major row gathers (`geot_gather_rows`: centres and patches straight from the (B,N,3) cloud, no transposes) and
    the kNN kernel, whose indices already are the flattened-batch indices the reference builds with `idx + idx_base`."""

    def __init__(self, num_group, group_size):
        super().__init__()
        self.num_group = num_group
        self.group_size = group_size

    @torch.no_grad()  # xyz is input data: no gradient flows through the patch coordinates in GeoT either
    def forward(self, xyz):
        from . import _lib
        batch_size, num_points, _ = xyz.shape
        xyz = xyz.contiguous().float()
        fps_idx = F_.furthest_point_sampling(xyz, self.num_group, flags=_lib.GEOT_FPS_CAP512 | _lib.GEOT_FPS_MAGSKIP)
        center = F_.gather_rows(xyz, fps_idx)  # B G 3
        idx, _ = F_.knnquery(self.group_size, xyz.view(-1, 3), center.view(-1, 3), None, None, n_per=num_points, m_per=self.num_group,
                             flags=_lib.GEOT_KNN_TIES_BY_INDEX)
        assert idx.shape == (batch_size * self.num_group, self.group_size)
        flat = xyz.view(1, batch_size * num_points, 3)
        neighborhood = F_.gather_rows(flat, idx.view(1, -1)).view(batch_size, self.num_group, self.group_size, 3)
        neighborhood = neighborhood - center.unsqueeze(2)
        return neighborhood, center, idx.view(-1).long()


def create_grouper(group_args):
    """group.py:338-352."""
    group_args_copy = copy.deepcopy(group_args)
    method = group_args_copy.pop('NAME', 'ballquery')
    radius = group_args_copy.pop('radius', 0.1)
    nsample = group_args_copy.pop('nsample', 20)
    if nsample is not None:
        if method == 'ballquery':
            return QueryAndGroup(radius, nsample, **group_args_copy)
        elif method == 'knn':
            return KNNGroup(nsample, **group_args_copy)
    return GroupAll()
