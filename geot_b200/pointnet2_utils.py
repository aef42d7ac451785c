"""Mirror of pointnet2/pointnet2_utils.py:48-422 (the erikwijmans/VoteNet-style API GeoT's transformer uses).

Same names and call conventions as the reference module (`pointnet2_utils.furthest_point_sample`,
`gather_operation`, `three_nn`, `three_interpolate`, `grouping_operation`, `ball_query`,
`QueryAndGroup`, `GroupAll`).  FPS here reproduces variant #4 of the reference bit for bit: a
512-thread tie order and the |p|^2 <= 1e-3 skip (pointnet2/_ext_src/src/sampling_gpu.cu:103-104,
include/cuda_utils.h:15); everything else shares the kernels of geot_b200.layers.
Also serves `pointnet2_ops.pointnet2_utils` (train.py:39,792-796), whose three_nn/three_interpolate
have the same semantics.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.autograd import Function

from . import _lib
from . import functional as F_
from . import layers as _layers
from .layers import (GatherOperation, GroupingOperation, ThreeInterpolate, ThreeNN)

_FPS_EXT = _lib.GEOT_FPS_CAP512 | _lib.GEOT_FPS_MAGSKIP


class FurthestPointSampling(Function):
    @staticmethod
    def forward(ctx, xyz, npoint):
        """xyz (B,N,3) -> idx (B,npoint) int32 (pointnet2_utils.py:48-74)."""
        out = F_.furthest_point_sampling(xyz.contiguous(), npoint, flags=_FPS_EXT)
        ctx.mark_non_differentiable(out)
        return out

    @staticmethod
    def backward(ctx, a=None):
        return None, None


furthest_point_sample = FurthestPointSampling.apply
gather_operation = GatherOperation.apply          # pointnet2_utils.py:80-114
three_nn = ThreeNN.apply                          # :117-146
three_interpolate = ThreeInterpolate.apply        # :149-203
grouping_operation = GroupingOperation.apply      # :206-254


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius, nsample, xyz, new_xyz):
        """pointnet2_utils.py:257-285: xyz (B,N,3), new_xyz (B,npoint,3) -> idx (B,npoint,nsample) int32."""
        idx = F_.ball_query(new_xyz.contiguous(), xyz.contiguous(), radius, nsample)
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


ball_query = BallQuery.apply


def _resample_unique(idx):
    """sample_uniformly of pointnet2_utils.py:291-373, for all (batch, region) rows at once on the device: each
    row becomes its distinct ids in ascending order followed by uniform draws (with replacement) from those ids.
    Same distribution as the reference's per-region torch.unique / torch.randint loop, not the same random stream
    (the reference draws on the CPU generator, one region at a time).  -> (idx int32, distinct count per row)."""
    nsample = idx.shape[-1]
    ordered = idx.sort(dim=-1).values
    is_first = torch.ones_like(ordered, dtype=torch.bool)
    is_first[..., 1:] = ordered[..., 1:] != ordered[..., :-1]
    rank = is_first.cumsum(-1) - 1                               # repeated ids share a rank, so the scatter is race-free
    distinct = torch.zeros_like(ordered).scatter_(-1, rank, ordered)
    count = rank[..., -1:] + 1
    slot = torch.arange(nsample, device=idx.device).expand_as(ordered)
    draw = (torch.rand(ordered.shape, device=idx.device) * count).long().minimum(count - 1)
    out = distinct.gather(-1, torch.where(slot < count, slot, draw))
    return out.to(idx.dtype).contiguous(), count.squeeze(-1).float().cpu()


class QueryAndGroup(nn.Module):
    """pointnet2_utils.py:291-373 (argument order xyz, new_xyz — the opposite of openpoints' grouper)."""

    def __init__(self, radius, nsample, use_xyz=True, ret_grouped_xyz=False, normalize_xyz=False,
                 sample_uniformly=False, ret_unique_cnt=False):
        super().__init__()
        self.radius, self.nsample, self.use_xyz = radius, nsample, use_xyz
        self.ret_grouped_xyz = ret_grouped_xyz
        self.normalize_xyz = normalize_xyz
        self.sample_uniformly = sample_uniformly
        self.ret_unique_cnt = ret_unique_cnt
        if ret_unique_cnt and not sample_uniformly:
            raise AssertionError("ret_unique_cnt needs sample_uniformly")

    def forward(self, xyz, new_xyz, features=None):
        idx = ball_query(self.radius, self.nsample, xyz, new_xyz)
        if self.sample_uniformly:
            idx, unique_cnt = _resample_unique(idx)
        if _layers._fusable(new_xyz, xyz):  # positions outside autograd: one fused kernel, same bits
            grouped_xyz = F_.group_xyz_relative(xyz, new_xyz, idx, relative=True,
                                                divisor=self.radius if self.normalize_xyz else 0.0)
        else:
            xyz_trans = xyz.transpose(1, 2).contiguous()
            grouped_xyz = grouping_operation(xyz_trans, idx)
            grouped_xyz = grouped_xyz - new_xyz.transpose(1, 2).unsqueeze(-1)
            if self.normalize_xyz:
                grouped_xyz = grouped_xyz / self.radius
        if features is not None:
            grouped_features = grouping_operation(features, idx)
            new_features = torch.cat([grouped_xyz, grouped_features], dim=1) if self.use_xyz else grouped_features
        else:
            if not self.use_xyz:
                raise AssertionError("QueryAndGroup without features needs use_xyz=True")
            new_features = grouped_xyz
        ret = [new_features]
        if self.ret_grouped_xyz:
            ret.append(grouped_xyz)
        if self.ret_unique_cnt:
            ret.append(unique_cnt)
        return ret[0] if len(ret) == 1 else tuple(ret)


class GroupAll(nn.Module):
    """pointnet2_utils.py:376-422.  The reference reads self.ret_grouped_xyz without ever setting it
    (SURVEY.md appendix D); here it is set from the constructor argument."""

    def __init__(self, use_xyz=True, ret_grouped_xyz=False):
        super().__init__()
        self.use_xyz = use_xyz
        self.ret_grouped_xyz = ret_grouped_xyz

    def forward(self, xyz, new_xyz, features=None):
        grouped_xyz = xyz.transpose(1, 2).unsqueeze(2)
        if features is not None:
            grouped_features = features.unsqueeze(2)
            new_features = torch.cat([grouped_xyz, grouped_features], dim=1) if self.use_xyz else grouped_features
        else:
            new_features = grouped_xyz
        return (new_features, grouped_xyz) if self.ret_grouped_xyz else new_features
