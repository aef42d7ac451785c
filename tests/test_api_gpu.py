"""The Python op API (the reference's own signatures) on top of the kernels: autograd contracts,
composites (QueryAndGroup, three_interpolation, pointops.knn/fps) and the reference's two own checks:
  - pointnet2/pointnet2_test.py:15-27   gradcheck of three_interpolate (fixed idx/weights, tol 1e-1)
  - subsample.py:171-185                gather_operation == torch.gather"""
import numpy as np
import pytest
import torch

from geot_b200 import layers, pointnet2_utils, pointops, synth
from geot_b200 import backends
from geot_b200 import functional as F

pytestmark = pytest.mark.gpu


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_reference_pointnet2_test_gradcheck(cuda):
    # restates pointnet2/pointnet2_test.py:15-27 (B=1, C=2, m=4; idx/weights fixed; atol=rtol=1e-1, fp32)
    feats = torch.randn(1, 2, 4, device="cuda", requires_grad=True)
    idx = torch.tensor([[[0, 1, 2], [1, 2, 3]]], dtype=torch.int32, device="cuda")
    weights = torch.tensor([[[1., 1., 1.], [2., 2., 2.]]], device="cuda")

    def interpolate_func(inputs):
        return pointnet2_utils.three_interpolate(inputs, idx, weights)

    assert torch.autograd.gradcheck(interpolate_func, feats, atol=1e-1, rtol=1e-1, eps=1e-2, nondet_tol=0.0)


def test_reference_subsample_smoke_gather_equals_torch_gather(cuda):
    # restates openpoints/models/layers/subsample.py:159-185 (FPS 10000 -> 4096, gather vs torch.gather)
    B, C, N = 2, 3, 10000
    points = torch.randn(B, N, 3, device="cuda")
    idx = layers.furthest_point_sample(points, 4096)
    a = torch.gather(points, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3))
    b = layers.gather_operation(points.transpose(1, 2).contiguous(), idx).transpose(1, 2).contiguous()
    assert torch.allclose(a, b) and idx.dtype == torch.int32
    for bi in range(B):
        assert len(torch.unique(idx[bi])) == 4096


def test_query_and_group_matches_torch_composition(cuda, oracle):
    xyz_h = synth.tooth_batch(2, 4096)
    xyz = cu(xyz_h)
    feat = torch.randn(2, 8, 4096, device="cuda", requires_grad=True)
    idx = layers.furthest_point_sample(xyz, 256)
    q = torch.gather(xyz, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    grouper = layers.create_grouper({"NAME": "ballquery", "radius": 0.1, "nsample": 32, "normalize_dp": True})
    dp, fj = grouper(q, xyz, feat)
    nbr = cu(oracle.ball_query(0.1, 32, xyz_h, q.cpu().numpy())).long()
    exp_dp = (torch.stack([xyz[b][nbr[b]] for b in range(2)]).permute(0, 3, 1, 2) - q.transpose(1, 2).unsqueeze(-1)) / 0.1
    exp_fj = layers.torch_grouping_operation(feat, nbr)
    assert torch.equal(dp, exp_dp) and torch.equal(fj, exp_fj)
    (fj * torch.arange(32, device="cuda")).sum().backward()
    g1 = feat.grad.clone(); feat.grad = None
    (exp_fj * torch.arange(32, device="cuda")).sum().backward()
    torch.testing.assert_close(g1, feat.grad, rtol=1e-5, atol=1e-5)
    # twin API (argument order xyz, new_xyz) returns the concatenated tensor
    nf = pointnet2_utils.QueryAndGroup(0.1, 32, use_xyz=True)(xyz, q, feat.detach())
    assert nf.shape == (2, 11, 256, 32) and torch.equal(nf[:, 3:], fj.detach())


@pytest.mark.parametrize("ns,relative,normalize", [(32, True, True), (7, True, False), (16, False, False), (5, True, True)])
def test_fused_xyz_grouping_equals_composite(cuda, ns, relative, normalize):
    """QueryAndGroup's xyz branch as one kernel (geot_group_xyz_relative) against the unfused composite
    (transpose copy + grouping_operation + subtract + divide), which is taken when a position tensor requires grad:
    bit-identical, for vectorised (ns % 4 == 0) and scalar neighbour counts."""
    xyz = cu(synth.tooth_batch(3, 3000))
    q = xyz[:, :200].contiguous()
    g = layers.QueryAndGroup(0.15, ns, relative_xyz=relative, normalize_dp=normalize)
    fused, _ = g(q, xyz, None)
    xyz_g = xyz.clone().requires_grad_(True)
    comp, _ = g(q, xyz_g, None)
    assert comp.grad_fn is not None and fused.grad_fn is None
    assert torch.equal(fused, comp.detach())
    nbr = layers.ball_query(0.15, ns, xyz, q)
    out = torch.empty(3, 3, 200, ns, device="cuda")
    F.group_xyz_relative(xyz, q, nbr, relative=relative, divisor=0.15 if (relative and normalize) else 0.0, out=out)
    assert torch.equal(out, fused)


@pytest.mark.parametrize("B,N,M,D", [(3, 5000, 777, 3), (2, 100, 100, 1), (1, 4096, 1, 7)])
def test_gather_rows_equals_transpose_gather_transpose(cuda, B, N, M, D):
    """Point-major gather (new_xyz = xyz[fps_idx]) against the reference's chain transpose -> gather_operation -> transpose."""
    src = torch.randn(B, N, D, device="cuda")
    idx = torch.randint(0, N, (B, M), device="cuda", dtype=torch.int32)
    got = F.gather_rows(src, idx)
    chain = layers.gather_operation(src.transpose(1, 2).contiguous(), idx).transpose(1, 2).contiguous()
    assert torch.equal(got, chain)
    assert torch.equal(got, torch.gather(src, 1, idx.long().unsqueeze(-1).expand(-1, -1, D)))


def test_pointops_fps_prefix_cache(cuda):
    """pointops.fps on the same tensor with decreasing targets (transformer.py:1037-1039) is served from the longest run:
    identical to fresh runs; an in-place write or another tensor invalidates the entry."""
    x = cu(synth.tooth_batch(2, 6000))
    a8 = pointops.fps(x, 2000)
    a4 = pointops.fps(x, 1000)
    a2 = pointops.fps(x, 500)
    assert torch.equal(a4, a8[:, :1000]) and torch.equal(a2, a8[:, :500])
    y = x.clone()
    assert torch.equal(pointops.fps(y, 1000), a4) and torch.equal(pointops.fps(x.clone(), 500), a2)  # fresh runs
    x.mul_(1.5)  # in-place edit: version counter moves, the cache must miss
    b4 = pointops.fps(x, 1000)
    assert torch.equal(b4, pointops.fps(x.clone(), 1000))
    assert torch.allclose(b4, a4 * 1.5)


def test_knn_group_and_knn_point(cuda):
    xyz = cu(synth.tooth_batch(2, 2048))
    q = xyz[:, :300].contiguous() + 1e-3
    d, i = layers.knn_point(9, q, xyz)
    # exact-difference cdist: the default matmul expansion loses ~1e-7 absolute on d2, i.e. 5e-5 on d near 1e-3
    dist = torch.cdist(q, xyz, compute_mode="donot_use_mm_for_euclid_dist")
    td, ti = dist.topk(9, dim=-1, largest=False, sorted=True)
    assert i.dtype == torch.int64 and torch.equal(i, ti)
    torch.testing.assert_close(d, td, rtol=1e-5, atol=1e-7)
    dp, fj = layers.KNNGroup(9)(q, xyz, xyz.transpose(1, 2).contiguous())
    assert dp.shape == (2, 3, 300, 9) and torch.equal(fj - q.transpose(1, 2).unsqueeze(-1), dp)
    d2, i2 = layers.KNN(9)(q, xyz)
    assert i2.dtype == torch.int32 and torch.equal(i2.long(), ti)


def test_three_interpolation_composite_and_grad(cuda):
    xyz = cu(synth.tooth_batch(2, 4096))
    known = xyz[:, ::8].contiguous()
    feat = torch.randn(2, 32, known.shape[1], device="cuda", requires_grad=True)
    out = layers.three_interpolation(xyz, known, feat)
    # known is a subset of xyz, so exact-zero distances occur: use the exact-difference cdist (the matmul
    # expansion turns those zeros into ~1e-4 and the 1/(d+1e-8) weights change completely)
    d = torch.cdist(xyz, known, compute_mode="donot_use_mm_for_euclid_dist")
    td, ti = d.topk(3, dim=-1, largest=False)
    w = 1.0 / (td + 1e-8); w = w / w.sum(-1, keepdim=True)
    exp = sum(torch.gather(feat, 2, ti[..., k].unsqueeze(1).expand(-1, 32, -1)) * w[..., k].unsqueeze(1) for k in range(3))
    torch.testing.assert_close(out, exp, rtol=1e-4, atol=1e-5)
    go = torch.randn_like(out)
    out.backward(go)
    g1 = feat.grad.clone(); feat.grad = None
    dist, idx = layers.three_nn(xyz, known)
    w = 1.0 / (dist + 1e-8); w = w / w.sum(-1, keepdim=True)
    exp = sum(torch.gather(feat, 2, idx[..., k].long().unsqueeze(1).expand(-1, 32, -1)) * w[..., k].unsqueeze(1) for k in range(3))
    exp.backward(go)
    torch.testing.assert_close(g1, feat.grad, rtol=1e-5, atol=1e-5)


def test_pointops_batched_helpers(cuda, oracle):
    xyz_h = synth.tooth_batch(3, 4000)
    xyz = cu(xyz_h)
    s = pointops.fps(xyz, 500)
    off = (np.arange(1, 4) * 4000).astype(np.int32); noff = (np.arange(1, 4) * 500).astype(np.int32)
    ei = oracle.furthestsampling(xyz_h.reshape(-1, 3), off, noff)
    assert np.array_equal(s.cpu().numpy(), xyz_h.reshape(-1, 3)[ei].reshape(3, 500, 3))
    i, d = pointops.knn(s, xyz, 5)
    assert i.dtype == torch.int64 and i.shape == (3, 500, 5) and int(i.max()) < 4000 and int(i.min()) >= 0
    assert torch.all(d[..., 0] == 0)  # every sample is one of the points
    w = torch.rand(3, 4000, device="cuda")
    sw = pointops.fps_weight(xyz, 100, w)
    ew = oracle.furthestsampling(xyz_h.reshape(-1, 3), off, (np.arange(1, 4) * 100).astype(np.int32), w.cpu().numpy().reshape(-1))
    assert np.array_equal(sw.cpu().numpy(), xyz_h.reshape(-1, 3)[ew].reshape(3, 100, 3))


def test_pointops_offset_autograd(cuda):
    n, c = 3000, 16
    xyz = cu(synth.tooth_cloud(n, 3))
    off = torch.tensor([1000, 3000], dtype=torch.int32, device="cuda")
    feat = torch.randn(n, c, device="cuda", requires_grad=True)
    out = pointops.queryandgroup(8, xyz, xyz, feat, None, off, off, use_xyz=True)
    assert out.shape == (n, 8, 3 + c)
    idx, _ = pointops.knnquery(8, xyz, xyz, off, off)
    assert torch.equal(out[..., 3:], feat[idx.long()])
    out[..., 3:].square().sum().backward()
    g = feat.grad.clone(); feat.grad = None
    feat[idx.long()].square().sum().backward()
    torch.testing.assert_close(g, feat.grad, rtol=1e-5, atol=1e-5)
    noff = torch.tensor([250, 750], dtype=torch.int32, device="cuda")
    sub = pointops.furthestsampling(xyz, off, noff)
    up = pointops.interpolation2(xyz[sub.long()].contiguous(), xyz, feat[sub.long()].detach().contiguous().requires_grad_(True), noff, off)
    assert up.shape == (n, c)


@pytest.mark.parametrize("method,normalize", [("knn", False), ("knn", True), ("ballquery", True)])
def test_pointops_querygroup_equals_row_gathers(cuda, method, normalize):
    """querygroup (pointops.py:106-147 role) against the fancy-indexing statement of the same thing, incl. the two
    normalize_dp branches, a caller-supplied idx, feat=None and the group-all (nsample=None) form."""
    n, c, ns, r = 3000, 8, 8, 0.2
    xyz = cu(synth.tooth_cloud(n, 5))
    off = torch.tensor([1200, 3000], dtype=torch.int32, device="cuda")
    feat = torch.randn(n, c, device="cuda", requires_grad=True)
    dp, fj = pointops.querygroup(ns, xyz, xyz, feat, off, off, radius=r, query_method=method, normalize_dp=normalize)
    idx = pointops.knnquery(ns, xyz, xyz, off, off)[0] if method == "knn" else pointops.ballquery(r, ns, xyz, xyz, off, off)
    rel = xyz[idx.long().flatten()].view(n, ns, 3) - xyz.unsqueeze(1)
    if normalize:
        rel = rel / ((rel.norm(dim=-1, keepdim=True).max(dim=1, keepdim=True)[0] + 1e-8) if method == "knn" else r)
    assert torch.equal(dp, rel) and torch.equal(fj, feat[idx.long().flatten()].view(n, ns, c))
    (fj * fj).sum().backward()
    g = feat.grad.clone(); feat.grad = None
    feat[idx.long().flatten()].square().sum().backward()
    torch.testing.assert_close(g, feat.grad, rtol=1e-5, atol=1e-5)
    dp2, fj2 = pointops.querygroup(ns, xyz, xyz, None, off, off, radius=r, query_method=method, normalize_dp=normalize, idx=idx)
    assert fj2 is None and torch.equal(dp2, dp)
    ga_xyz, ga_feat = pointops.querygroup(None, xyz.view(2, 1500, 3), None, feat.detach().view(2, c, 1500), off, off)
    assert ga_xyz.shape == (2, 3, 1, 1500) and ga_feat.shape == (2, c, 1, 1500)


def test_query_and_group_sample_uniformly(cuda):
    """pointnet2_utils.QueryAndGroup(sample_uniformly=True) (pointnet2_utils.py:291-373 role), vectorised: every row is
    its distinct ball-query ids ascending, then draws from exactly that set; the count is returned like the reference's."""
    xyz = cu(synth.tooth_batch(2, 2048))
    q = xyz[:, ::16].contiguous()
    ns = 24
    raw = pointnet2_utils.ball_query(0.12, ns, xyz, q)
    torch.manual_seed(5)
    grouper = pointnet2_utils.QueryAndGroup(0.12, ns, use_xyz=True, ret_grouped_xyz=True, sample_uniformly=True, ret_unique_cnt=True)
    nf, gxyz, cnt = grouper(xyz, q, None)
    assert nf.shape == (2, 3, 128, ns) and cnt.shape == (2, 128) and cnt.device.type == "cpu"
    idx, cnt2 = pointnet2_utils._resample_unique(raw)
    assert idx.dtype == torch.int32 and torch.equal(cnt2, cnt)
    raw_h, idx_h = raw.cpu(), idx.cpu()
    some_padding = False
    for b in range(2):
        for p in range(128):
            u = torch.unique(raw_h[b, p])
            k = u.numel()
            assert cnt[b, p] == k and torch.equal(idx_h[b, p, :k], u.int())
            assert set(idx_h[b, p, k:].tolist()) <= set(u.tolist())
            some_padding |= k < ns
    assert some_padding
    with pytest.raises(AssertionError):
        pointnet2_utils.QueryAndGroup(0.1, 8, ret_unique_cnt=True)


def test_facades_have_the_reference_callables(cuda):
    names = ["ball_query_wrapper", "group_points_wrapper", "group_points_grad_wrapper", "gather_points_wrapper",
             "gather_points_grad_wrapper", "furthest_point_sampling_wrapper", "three_nn_wrapper",
             "three_interpolate_wrapper", "three_interpolate_grad_wrapper"]  # pointnet2_api.cpp:10-24
    assert all(hasattr(backends.pointnet2_batch_cuda, n) for n in names)
    B, N, M = 2, 2048, 128
    xyz = cu(synth.tooth_batch(B, N))
    out = torch.zeros(B, M, dtype=torch.int32, device="cuda")
    temp = torch.full((B, N), 1e10, device="cuda")
    backends.pointnet2_batch_cuda.furthest_point_sampling_wrapper(B, N, M, xyz, temp, out)
    assert torch.equal(out, layers.furthest_point_sample(xyz, M))
    e = backends.pointnet2_ext.furthest_point_sampling(xyz, M)
    assert torch.equal(e, pointnet2_utils.furthest_point_sample(xyz, M))
    with pytest.raises(RuntimeError):
        backends.pointnet2_ext.gather_points(xyz.double(), out)
