"""Import-name shims so UNMODIFIED GeoT code runs on the sm_100a kernels.

    import geot_b200.compat as compat
    compat.install()            # before importing openpoints / pointops / pointnet2 / utils from the GeoT tree

installs, under the names the reference imports:
    openpoints.cpp.pointnet2_batch.pointnet2_cuda   <- backends.pointnet2_batch_cuda   (pointnet2_api.cpp:10-24)
    pointops_cuda                                   <- backends.pointops_cuda          (pointops_api.cpp)
    pointnet2._ext                                  <- backends.pointnet2_ext          (bindings.cpp:9-21)
    knn_cuda.KNN                                    <- KNN below                       (un-vendored dep, transformer.py:11)
    pointnet2_ops.pointnet2_utils                   <- geot_b200.pointnet2_utils       (un-vendored dep, train.py:39)
Nothing here computes anything; it is sys.modules plumbing.
"""
from __future__ import annotations

import sys
import types

import torch
import torch.nn as nn

from . import backends
from . import _lib
from . import functional as F_


class KNN(nn.Module):
    """knn_cuda.KNN(k, transpose_mode) as GeoT uses it (parity unpinned: the package is not in the reference tree).

    transpose_mode=True : ref (B,N,3), query (B,M,3) -> dist (B,M,k), idx (B,M,k)   (transformer.py:280,293)
    transpose_mode=False: ref (B,3,N), query (B,3,M) -> dist (B,k,M), idx (B,k,M)   (transformer.py:313,353)
    Ascending Euclidean distance, idx int64 into ref's point axis, no gradient."""

    def __init__(self, k, transpose_mode=False):
        super().__init__()
        self.k = k
        self._t = transpose_mode

    @torch.no_grad()
    def forward(self, ref, query):
        assert ref.size(0) == query.size(0), "ref.shape={} != query.shape={}".format(ref.shape, query.shape)
        if not self._t:
            ref, query = ref.transpose(1, 2), query.transpose(1, 2)
        ref = ref.contiguous().float()
        query = query.contiguous().float()
        B, N, D = ref.shape
        M = query.shape[1]
        assert D == 3, "the sm_100a kNN kernel is 3-D"
        idx, d2 = F_.knnquery(self.k, ref.view(-1, 3), query.view(-1, 3), None, None, n_per=N, m_per=M, flags=_lib.GEOT_KNN_TIES_BY_INDEX)
        base = (torch.arange(B, device=idx.device, dtype=torch.int32) * N).view(B, 1, 1)
        idx = (idx.view(B, M, self.k) - base).long()
        dist = torch.sqrt(d2).view(B, M, self.k)
        if not self._t:
            dist, idx = dist.transpose(1, 2).contiguous(), idx.transpose(1, 2).contiguous()
        return dist, idx


def install(force: bool = True):
    """Register the facades in sys.modules.  Safe to call twice."""
    def put(name, mod):
        if force or name not in sys.modules:
            sys.modules[name] = mod

    pn2b = types.ModuleType("openpoints.cpp.pointnet2_batch")
    pn2b.pointnet2_cuda = backends.pointnet2_batch_cuda
    put("openpoints.cpp.pointnet2_batch", pn2b)
    put("pointnet2_batch_cuda", backends.pointnet2_batch_cuda)
    put("pointops_cuda", backends.pointops_cuda)
    put("pointnet2._ext", backends.pointnet2_ext)
    knn_cuda = types.ModuleType("knn_cuda")
    knn_cuda.KNN = KNN
    put("knn_cuda", knn_cuda)
    from . import pointnet2_utils as p2u
    p2o = types.ModuleType("pointnet2_ops")
    p2o.pointnet2_utils = p2u
    put("pointnet2_ops", p2o)
    put("pointnet2_ops.pointnet2_utils", p2u)
