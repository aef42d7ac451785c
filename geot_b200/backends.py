"""pybind-shaped facades: objects with the exact callables of the reference's three extension modules,
so UNMODIFIED reference Python (openpoints/models/layers/*.py, pointops/functions/pointops.py,
pointnet2/pointnet2_utils.py) can run on the sm_100a kernels.  compat/ installs them under the import
names the reference uses.

  pointnet2_batch_cuda   openpoints/cpp/pointnet2_batch/src/pointnet2_api.cpp:10-24   caller allocates outputs
  pointops_cuda          pointops/src/pointops_api.cpp:8-12 + openpoints/cpp/pointops/src/pointops_api.cpp:13-25
  pointnet2._ext         pointnet2/_ext_src/src/bindings.cpp:9-21                      callee allocates, returns tensors

The backward wrappers receive idx (the reference API has no plan argument): the scatter plan is built
per call here.  The reference pre-zeroes gradient buffers because its kernels accumulate; ours
overwrite every element, which gives the same result on a zeroed buffer.
"""
from __future__ import annotations

import torch

from . import _lib
from . import functional as F_


class _PointNet2Batch:
    """`from openpoints.cpp.pointnet2_batch import pointnet2_cuda`"""

    @staticmethod
    def ball_query_wrapper(B, N, M, radius, nsample, new_xyz, xyz, idx):
        F_.ball_query(new_xyz, xyz, radius, nsample, out=idx)
        return 1

    @staticmethod
    def group_points_wrapper(B, C, N, M, nsample, points, idx, out):
        F_.group_points(points, idx, out=out)
        return 1

    @staticmethod
    def group_points_grad_wrapper(B, C, N, M, nsample, grad_out, idx, grad_points):
        F_.group_points_grad(grad_out, F_.grouping_plan(idx, B, N), N, out=grad_points)
        return 1

    @staticmethod
    def gather_points_wrapper(B, C, N, M, points, idx, out):
        F_.gather_points(points, idx, out=out)
        return 1

    @staticmethod
    def gather_points_grad_wrapper(B, C, N, M, grad_out, idx, grad_points):
        F_.gather_points_grad(grad_out, F_.ScatterPlan(idx, B, N), N, out=grad_points)
        return 1

    @staticmethod
    def furthest_point_sampling_wrapper(B, N, M, xyz, temp, idx):
        F_.furthest_point_sampling(xyz, M, flags=0, temp=temp, out=idx)
        return 1

    @staticmethod
    def three_nn_wrapper(B, n, m, unknown, known, dist2, idx):
        F_.three_nn(unknown, known, dist2=dist2, idx=idx)

    @staticmethod
    def three_interpolate_wrapper(B, C, m, n, points, idx, weight, out):
        F_.three_interpolate(points, idx, weight, out=out)

    @staticmethod
    def three_interpolate_grad_wrapper(B, C, n, m, grad_out, idx, weight, grad_points):
        F_.three_interpolate_grad(grad_out, weight, F_.ScatterPlan(idx, B, m), m, out=grad_points)


class _PointOps:
    """`import pointops_cuda` (both the trimmed top-level build and the full openpoints/cpp build)."""

    @staticmethod
    def knnquery_cuda(m, nsample, xyz, new_xyz, offset, new_offset, idx, dist2):
        F_.knnquery(nsample, xyz, new_xyz, offset, new_offset, idx=idx, dist2=dist2)

    @staticmethod
    def furthestsampling_cuda(b, n_max, xyz, offset, new_offset, tmp, idx):
        F_.furthestsampling(xyz, offset, new_offset, n_max=int(n_max), m_total=idx.numel(), tmp=tmp, out=idx)

    @staticmethod
    def furthestsampling_weights_cuda(b, n_max, xyz, offset, new_offset, weights, tmp, idx):
        F_.furthestsampling(xyz, offset, new_offset, weights=weights, n_max=int(n_max), m_total=idx.numel(), tmp=tmp, out=idx)

    @staticmethod
    def ballquery_cuda(m, radius, nsample, xyz, new_xyz, offset, new_offset, idx):
        F_.ballquery_offset(radius, nsample, xyz, new_xyz, offset, new_offset, out=idx)

    @staticmethod
    def grouping_forward_cuda(m, nsample, c, input, idx, output):
        F_.grouping_forward(input, idx, out=output)

    @staticmethod
    def grouping_backward_cuda(m, nsample, c, grad_output, idx, grad_input):
        n = grad_input.shape[0]
        F_.grouping_backward(grad_output, F_.ScatterPlan(idx, 1, n), n, out=grad_input)

    @staticmethod
    def interpolation_forward_cuda(n, c, k, input, idx, weight, output):
        F_.interpolation_forward(input, idx, weight, out=output)

    @staticmethod
    def interpolation_backward_cuda(n, c, k, grad_output, idx, weight, grad_input):
        m = grad_input.shape[0]
        F_.interpolation_backward(grad_output, weight, F_.ScatterPlan(idx, 1, m), m, k, out=grad_input)

    @staticmethod
    def subtraction_forward_cuda(n, nsample, c, input1, input2, idx, output):
        F_.subtraction_forward(input1, input2, idx, out=output)

    @staticmethod
    def subtraction_backward_cuda(n, nsample, c, idx, grad_output, grad_input1, grad_input2):
        F_.subtraction_backward(grad_output, F_.ScatterPlan(idx, 1, n), g1=grad_input1, g2=grad_input2)

    @staticmethod
    def aggregation_forward_cuda(n, nsample, c, w_c, input, position, weight, idx, output):
        F_.aggregation_forward(input, position, weight, idx, out=output)

    @staticmethod
    def aggregation_backward_cuda(n, nsample, c, w_c, input, position, weight, idx, grad_output,
                                  grad_input, grad_position, grad_weight):
        F_.aggregation_backward(input, position, weight, idx, grad_output, F_.ScatterPlan(idx, 1, n),
                                gi=grad_input, gp=grad_position, gw=grad_weight)


class _PointNet2Ext:
    """`import pointnet2._ext as _ext` — tensor-returning, dtype/contiguity checked like utils.h:8-28."""

    @staticmethod
    def _chk(t, dtype, name):
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be a contiguous tensor")
        if t.dtype != dtype:
            raise RuntimeError(f"{name} must be a {'float' if dtype == torch.float32 else 'int'} tensor")
        if not t.is_cuda:
            raise RuntimeError("CPU not supported")

    @classmethod
    def gather_points(cls, points, idx):
        cls._chk(points, torch.float32, "points"); cls._chk(idx, torch.int32, "idx")
        return F_.gather_points(points, idx)

    @classmethod
    def gather_points_grad(cls, grad_out, idx, n):
        cls._chk(grad_out, torch.float32, "grad_out"); cls._chk(idx, torch.int32, "idx")
        return F_.gather_points_grad(grad_out, F_.ScatterPlan(idx, idx.shape[0], n), n)

    @classmethod
    def furthest_point_sampling(cls, points, nsamples):
        cls._chk(points, torch.float32, "points")
        return F_.furthest_point_sampling(points, nsamples, flags=_lib.GEOT_FPS_CAP512 | _lib.GEOT_FPS_MAGSKIP)

    @classmethod
    def three_nn(cls, unknowns, knows):
        cls._chk(unknowns, torch.float32, "unknowns"); cls._chk(knows, torch.float32, "knows")
        dist2, idx = F_.three_nn(unknowns, knows)
        return [dist2, idx]

    @classmethod
    def three_interpolate(cls, points, idx, weight):
        cls._chk(points, torch.float32, "points"); cls._chk(idx, torch.int32, "idx"); cls._chk(weight, torch.float32, "weight")
        return F_.three_interpolate(points, idx, weight)

    @classmethod
    def three_interpolate_grad(cls, grad_out, idx, weight, m):
        cls._chk(grad_out, torch.float32, "grad_out"); cls._chk(idx, torch.int32, "idx"); cls._chk(weight, torch.float32, "weight")
        return F_.three_interpolate_grad(grad_out, weight, F_.ScatterPlan(idx, idx.shape[0], m), m)

    @classmethod
    def ball_query(cls, new_xyz, xyz, radius, nsample):
        cls._chk(new_xyz, torch.float32, "new_xyz"); cls._chk(xyz, torch.float32, "xyz")
        return F_.ball_query(new_xyz, xyz, radius, nsample)

    @classmethod
    def group_points(cls, points, idx):
        cls._chk(points, torch.float32, "points"); cls._chk(idx, torch.int32, "idx")
        return F_.group_points(points, idx)

    @classmethod
    def group_points_grad(cls, grad_out, idx, n):
        cls._chk(grad_out, torch.float32, "grad_out"); cls._chk(idx, torch.int32, "idx")
        return F_.group_points_grad(grad_out, F_.grouping_plan(idx, idx.shape[0], n), n)


pointnet2_batch_cuda = _PointNet2Batch()
pointops_cuda = _PointOps()
pointnet2_ext = _PointNet2Ext()
