"""`knn_cuda.KNN` is not vendored in the reference tree (call sites transformer.py:280,293,313,353), so
no reference code can pin it.  This file states and tests the PROJECT'S DEFINITION of it: float64 brute force,
neighbours in ascending Euclidean distance, lowest index first on exact ties, idx int64 into ref's point axis, layouts
per `transpose_mode`.  geot_b200.compat.KNN must return exactly those indices wherever the float64 order is not decided
inside fp32 rounding (a margin test), and must agree with the reference's in-tree kNN kernel everywhere."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from geot_b200 import compat, synth  # noqa: E402


def _cloud(B, N, seed):
    return torch.from_numpy(synth.tooth_batch(B, N, first_cloud=seed)).cuda()


@pytest.mark.parametrize("k,transpose,Nr,Nq", [(32, True, 16000, 512), (4, False, 4096, 8192), (4, False, 512, 4096), (1, True, 100, 100)])
def test_compat_knn_matches_float64_definition(cuda, k, transpose, Nr, Nq):
    from tests import ref_backends as rb
    B = 3
    ref_pts = _cloud(B, max(Nr, Nq), 20)[:, :Nr].contiguous()
    qry_pts = _cloud(B, max(Nr, Nq), 40)[:, :Nq].contiguous()
    a, b = (ref_pts, qry_pts) if transpose else (ref_pts.transpose(1, 2).contiguous(), qry_pts.transpose(1, 2).contiguous())
    d_def, i_def = rb.KNNDefinition(k, transpose_mode=transpose)(a, b)
    d_got, i_got = compat.KNN(k, transpose_mode=transpose)(a, b)
    assert i_got.dtype == torch.int64 and i_got.shape == i_def.shape and d_got.shape == d_def.shape
    if not transpose:
        i_def, i_got, d_def, d_got = (t.transpose(1, 2) for t in (i_def, i_got, d_def, d_got))
    # rows whose k+1 nearest float64 distances are separated by more than fp32 rounding of d^2 (4 ulp): must be equal
    dd = torch.cdist(qry_pts.double(), ref_pts.double()) ** 2
    srt = torch.sort(dd, dim=2).values[:, :, :k + 1]
    gap = (srt[:, :, 1:] - srt[:, :, :-1]) if srt.shape[2] > 1 else torch.ones_like(srt)
    safe = (gap > 4 * 1.2e-7 * srt[:, :, -1:].clamp_min(1e-30)).all(dim=2)
    assert safe.float().mean() > 0.95, "the inputs are meant to be (almost) tie-free"
    assert torch.equal(i_got[safe], i_def[safe]), "compat.KNN differs from the float64 definition on well-separated rows"
    # everywhere: same neighbour SET up to near-ties, and distances within fp32 rounding
    torch.testing.assert_close(d_got.double(), d_def, rtol=2e-6, atol=1e-7)
    # and bit-equal to the reference's in-tree kNN kernel (pointops knnquery), the 'ref' side of tests/test_models_gpu.py
    d_ref, i_ref = rb.KNN(k, transpose_mode=transpose)(a, b)
    if not transpose:
        i_ref = i_ref.transpose(1, 2)
    assert torch.equal(i_got, i_ref)
