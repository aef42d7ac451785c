"""The reference's OWN kernels (oracle/_ref/*.so, built unchanged from /root/reference by oracle/ref_build/Makefile)
behind the same three pybind-shaped module objects geot_b200.backends provides.  Test infrastructure: lets the SAME
unmodified reference model run once per backend (tests/refenv.py switches between the two), so that a whole
forward/backward is compared, not only single ops.

  pointnet2_batch_cuda   <- libref_pn2batch.so       (openpoints/cpp/pointnet2_batch/src/pointnet2_api.cpp:10-24)
  pointops_cuda          <- libref_pointops_full.so  (openpoints/cpp/pointops/src/pointops_api.cpp:13-25) and
                            libref_pointops.so for furthestsampling_weights (pointops/src/pointops_api.cpp:8-12)
  pointnet2._ext         <- libref_pn2ext.so         (pointnet2/_ext_src/src/bindings.cpp:9-21; allocations follow
                            the .cpp hosts: torch::zeros outputs, sampling.cpp:71-77 temp = 1e10)

The reference launches on the legacy default stream; torch's default stream is that stream, so no extra sync is
needed while the test runs on it (asserted).
"""
import ctypes as C

import torch

from tests import refgpu

_f = C.c_float


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _L(name):
    assert torch.cuda.current_stream() == torch.cuda.default_stream(), "reference kernels run on the default stream"
    return refgpu.lib(name)


class _PointNet2Batch:
    @staticmethod
    def ball_query_wrapper(B, N, M, radius, nsample, new_xyz, xyz, idx):
        _L("pn2batch").ref_ball_query(B, N, M, _f(radius), nsample, _p(new_xyz), _p(xyz), _p(idx))
        return 1

    @staticmethod
    def group_points_wrapper(B, Cc, N, M, nsample, points, idx, out):
        _L("pn2batch").ref_group_points(B, Cc, N, M, nsample, _p(points), _p(idx), _p(out))
        return 1

    @staticmethod
    def group_points_grad_wrapper(B, Cc, N, M, nsample, grad_out, idx, grad_points):
        _L("pn2batch").ref_group_points_grad(B, Cc, N, M, nsample, _p(grad_out), _p(idx), _p(grad_points))
        return 1

    @staticmethod
    def gather_points_wrapper(B, Cc, N, M, points, idx, out):
        _L("pn2batch").ref_gather_points(B, Cc, N, M, _p(points), _p(idx), _p(out))
        return 1

    @staticmethod
    def gather_points_grad_wrapper(B, Cc, N, M, grad_out, idx, grad_points):
        _L("pn2batch").ref_gather_points_grad(B, Cc, N, M, _p(grad_out), _p(idx), _p(grad_points))
        return 1

    @staticmethod
    def furthest_point_sampling_wrapper(B, N, M, xyz, temp, idx):
        _L("pn2batch").ref_furthest_point_sampling(B, N, M, _p(xyz), _p(temp), _p(idx))
        return 1

    @staticmethod
    def three_nn_wrapper(B, n, m, unknown, known, dist2, idx):
        _L("pn2batch").ref_three_nn(B, n, m, _p(unknown), _p(known), _p(dist2), _p(idx))

    @staticmethod
    def three_interpolate_wrapper(B, Cc, m, n, points, idx, weight, out):
        _L("pn2batch").ref_three_interpolate(B, Cc, m, n, _p(points), _p(idx), _p(weight), _p(out))

    @staticmethod
    def three_interpolate_grad_wrapper(B, Cc, n, m, grad_out, idx, weight, grad_points):
        _L("pn2batch").ref_three_interpolate_grad(B, Cc, n, m, _p(grad_out), _p(idx), _p(weight), _p(grad_points))


class _PointOps:
    @staticmethod
    def knnquery_cuda(m, nsample, xyz, new_xyz, offset, new_offset, idx, dist2):
        _L("pointops_full").ref_knnquery(m, nsample, _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx), _p(dist2))

    @staticmethod
    def furthestsampling_cuda(b, n_max, xyz, offset, new_offset, tmp, idx):
        _L("pointops_full").ref_furthestsampling(b, int(n_max), _p(xyz), _p(offset), _p(new_offset), _p(tmp), _p(idx))

    @staticmethod
    def furthestsampling_weights_cuda(b, n_max, xyz, offset, new_offset, weights, tmp, idx):
        _L("pointops").ref_furthestsampling_weights(b, int(n_max), _p(xyz), _p(offset), _p(new_offset), _p(weights), _p(tmp), _p(idx))

    @staticmethod
    def ballquery_cuda(m, radius, nsample, xyz, new_xyz, offset, new_offset, idx):
        _L("pointops_full").ref_ballquery(m, _f(radius), nsample, _p(xyz), _p(new_xyz), _p(offset), _p(new_offset), _p(idx))

    @staticmethod
    def grouping_forward_cuda(m, nsample, c, input, idx, output):
        _L("pointops_full").ref_grouping_forward(m, nsample, c, _p(input), _p(idx), _p(output))

    @staticmethod
    def grouping_backward_cuda(m, nsample, c, grad_output, idx, grad_input):
        _L("pointops_full").ref_grouping_backward(m, nsample, c, _p(grad_output), _p(idx), _p(grad_input))

    @staticmethod
    def interpolation_forward_cuda(n, c, k, input, idx, weight, output):
        _L("pointops_full").ref_interpolation_forward(n, c, k, _p(input), _p(idx), _p(weight), _p(output))

    @staticmethod
    def interpolation_backward_cuda(n, c, k, grad_output, idx, weight, grad_input):
        _L("pointops_full").ref_interpolation_backward(n, c, k, _p(grad_output), _p(idx), _p(weight), _p(grad_input))

    @staticmethod
    def subtraction_forward_cuda(n, nsample, c, input1, input2, idx, output):
        _L("pointops_full").ref_subtraction_forward(n, nsample, c, _p(input1), _p(input2), _p(idx), _p(output))

    @staticmethod
    def subtraction_backward_cuda(n, nsample, c, idx, grad_output, grad_input1, grad_input2):
        _L("pointops_full").ref_subtraction_backward(n, nsample, c, _p(idx), _p(grad_output), _p(grad_input1), _p(grad_input2))

    @staticmethod
    def aggregation_forward_cuda(n, nsample, c, w_c, input, position, weight, idx, output):
        _L("pointops_full").ref_aggregation_forward(n, nsample, c, w_c, _p(input), _p(position), _p(weight), _p(idx), _p(output))

    @staticmethod
    def aggregation_backward_cuda(n, nsample, c, w_c, input, position, weight, idx, grad_output,
                                  grad_input, grad_position, grad_weight):
        _L("pointops_full").ref_aggregation_backward(n, nsample, c, w_c, _p(input), _p(position), _p(weight), _p(idx),
                                                     _p(grad_output), _p(grad_input), _p(grad_position), _p(grad_weight))


class _PointNet2Ext:
    """Allocation rules of the reference's .cpp hosts (pointnet2/_ext_src/src/*.cpp): outputs are torch::zeros."""

    @staticmethod
    def gather_points(points, idx):
        B, Cc, N = points.shape
        out = torch.zeros((B, Cc, idx.shape[1]), device=points.device)
        _L("pn2ext").ref_gather_points(B, Cc, N, idx.shape[1], _p(points), _p(idx), _p(out))
        return out

    @staticmethod
    def gather_points_grad(grad_out, idx, n):
        B, Cc, M = grad_out.shape
        out = torch.zeros((B, Cc, n), device=grad_out.device)
        _L("pn2ext").ref_gather_points_grad(B, Cc, n, M, _p(grad_out), _p(idx), _p(out))
        return out

    @staticmethod
    def furthest_point_sampling(points, nsamples):
        B, N, _ = points.shape
        out = torch.zeros((B, nsamples), dtype=torch.int32, device=points.device)
        tmp = torch.full((B, N), 1e10, device=points.device)
        _L("pn2ext").ref_furthest_point_sampling(B, N, nsamples, _p(points), _p(tmp), _p(out))
        return out

    @staticmethod
    def three_nn(unknowns, knows):
        B, n, _ = unknowns.shape
        idx = torch.zeros((B, n, 3), dtype=torch.int32, device=unknowns.device)
        d2 = torch.zeros((B, n, 3), device=unknowns.device)
        _L("pn2ext").ref_three_nn(B, n, knows.shape[1], _p(unknowns), _p(knows), _p(d2), _p(idx))
        return [d2, idx]

    @staticmethod
    def three_interpolate(points, idx, weight):
        B, Cc, m = points.shape
        n = idx.shape[1]
        out = torch.zeros((B, Cc, n), device=points.device)
        _L("pn2ext").ref_three_interpolate(B, Cc, m, n, _p(points), _p(idx), _p(weight), _p(out))
        return out

    @staticmethod
    def three_interpolate_grad(grad_out, idx, weight, m):
        B, Cc, n = grad_out.shape
        out = torch.zeros((B, Cc, m), device=grad_out.device)
        _L("pn2ext").ref_three_interpolate_grad(B, Cc, n, m, _p(grad_out), _p(idx), _p(weight), _p(out))
        return out

    @staticmethod
    def ball_query(new_xyz, xyz, radius, nsample):
        B, M, _ = new_xyz.shape
        idx = torch.zeros((B, M, nsample), dtype=torch.int32, device=xyz.device)
        _L("pn2ext").ref_ball_query(B, xyz.shape[1], M, _f(radius), nsample, _p(new_xyz), _p(xyz), _p(idx))
        return idx

    @staticmethod
    def group_points(points, idx):
        B, Cc, N = points.shape
        _, M, ns = idx.shape
        out = torch.zeros((B, Cc, M, ns), device=points.device)
        _L("pn2ext").ref_group_points(B, Cc, N, M, ns, _p(points), _p(idx), _p(out))
        return out

    @staticmethod
    def group_points_grad(grad_out, idx, n):
        B, Cc, M, ns = grad_out.shape
        out = torch.zeros((B, Cc, n), device=grad_out.device)
        _L("pn2ext").ref_group_points_grad(B, Cc, n, M, ns, _p(grad_out), _p(idx), _p(out))
        return out


class KNN(torch.nn.Module):
    """`knn_cuda.KNN` on the reference side.  The package is not in the reference tree (parity
    unpinned), so its stand-in here is the reference's in-tree kNN kernel with the same contract (k nearest in
    ascending distance): pointops/src/knnquery/knnquery_cuda_kernel.cu through libref_pointops.so, batched by
    offsets exactly as pointops.knn does (pointops/functions/pointops.py:7-21).  Layouts as used at
    transformer.py:293 (transpose_mode=True) and :353 (False).  The float64 DEFINITION the project states for
    knn_cuda is `KNNDefinition` below (tests/test_knn_definition_gpu.py)."""

    def __init__(self, k, transpose_mode=False):
        super().__init__()
        self.k, self._t = k, transpose_mode

    @torch.no_grad()
    def forward(self, ref, query):
        if not self._t:
            ref, query = ref.transpose(1, 2), query.transpose(1, 2)
        ref, query = ref.contiguous().float(), query.contiguous().float()
        B, N, _ = ref.shape
        M = query.shape[1]
        off = torch.arange(1, B + 1, dtype=torch.int32, device=ref.device) * N
        noff = torch.arange(1, B + 1, dtype=torch.int32, device=ref.device) * M
        idx = torch.zeros((B * M, self.k), dtype=torch.int32, device=ref.device)
        d2 = torch.zeros((B * M, self.k), device=ref.device)
        _L("pointops").ref_knnquery(B * M, self.k, _p(ref.view(-1, 3)), _p(query.view(-1, 3)), _p(off), _p(noff), _p(idx), _p(d2))
        idx = (idx.view(B, M, self.k) - (off - N).view(B, 1, 1)).long()
        dist = torch.sqrt(d2).view(B, M, self.k)
        if not self._t:
            dist, idx = dist.transpose(1, 2).contiguous(), idx.transpose(1, 2).contiguous()
        return dist, idx


class KNNDefinition(torch.nn.Module):
    """The project's DEFINITION of the un-vendored `knn_cuda.KNN(k, transpose_mode)`: float64 brute force, ascending
    Euclidean distance, lowest index first on exact ties."""

    def __init__(self, k, transpose_mode=False):
        super().__init__()
        self.k, self._t = k, transpose_mode

    @torch.no_grad()
    def forward(self, ref, query):
        if not self._t:
            ref, query = ref.transpose(1, 2), query.transpose(1, 2)
        r, q = ref.double(), query.double()
        out_d, out_i = [], []
        for b in range(r.shape[0]):
            di, ii = [], []
            for s in range(0, q.shape[1], 2048):
                d = ((q[b, s:s + 2048, None, :] - r[b, None, :, :]) ** 2).sum(-1)      # exact differences in fp64
                dd, jj = torch.sort(d, dim=1, stable=True)                               # (distance, index) order
                di.append(dd[:, :self.k]); ii.append(jj[:, :self.k])
            out_d.append(torch.cat(di)); out_i.append(torch.cat(ii))
        dist = torch.sqrt(torch.stack(out_d))
        idx = torch.stack(out_i)
        if not self._t:
            dist, idx = dist.transpose(1, 2).contiguous(), idx.transpose(1, 2).contiguous()
        return dist, idx


class _LazyRefP2U:
    """`pointnet2_ops.pointnet2_utils` on the reference side = the in-tree twin pointnet2/pointnet2_utils.py
    (same semantics), which itself runs on the switched `pointnet2._ext`."""

    def __getattr__(self, name):
        import pointnet2.pointnet2_utils as p2u
        return getattr(p2u, name)


pointnet2_batch_cuda = _PointNet2Batch()
pointops_cuda = _PointOps()
pointnet2_ext = _PointNet2Ext()
pointnet2_ops_utils = _LazyRefP2U()
