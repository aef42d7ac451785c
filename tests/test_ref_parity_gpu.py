"""Strongest pin: the reference's OWN kernels (oracle/_ref, compiled unchanged from /root/reference for
sm_100) run on this B200 next to (a) the CPU oracle and (b) the geot_b200 kernels, on identical inputs.
This is what makes the oracle 'pinned' rather than a bare restatement.  Skipped if oracle/_ref was not built."""
import numpy as np
import pytest
import torch

from geot_b200 import _lib, synth
from geot_b200 import functional as F
from tests import refgpu

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refgpu.available(), reason="oracle/_ref not built")]


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def clouds(kind, B, N, seed=0):
    if kind == "tooth":
        return synth.tooth_batch(B, N, first_cloud=seed)
    return np.stack([synth.adversarial_cloud(N, 300 + seed + i, kind) for i in range(B)])


@pytest.mark.parametrize("kind,B,N,M", [("tooth", 2, 16384, 1024), ("quant", 2, 6000, 800), ("dup", 2, 3000, 1000),
                                         ("origin", 2, 9000, 2000), ("tooth", 2, 1000, 200), ("quant", 1, 777, 777)])
def test_fps_three_way(cuda, oracle, kind, B, N, M):
    xyz = clouds(kind, B, N)
    x = cu(xyz)
    for which, flags, cap, skip in [("pn2batch", 0, 1024, False), ("pn2ext", _lib.GEOT_FPS_CAP512 | _lib.GEOT_FPS_MAGSKIP, 512, True)]:
        ref = refgpu.fps(x, M, which).cpu().numpy()
        assert np.array_equal(oracle.furthest_point_sample(xyz, M, cap=cap, mag_skip=skip), ref), f"oracle vs reference {which}"
        assert np.array_equal(F.furthest_point_sampling(x, M, flags=flags).cpu().numpy(), ref), f"geot vs reference {which}"


@pytest.mark.parametrize("weighted", [False, True])
def test_fps_offset_three_way(cuda, oracle, weighted):
    sizes, ms = [5000, 1200, 16384, 333], [300, 100, 1024, 333]
    xyz = np.concatenate([synth.adversarial_cloud(n, 70 + i, "quant") for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    w = np.random.default_rng(1).random(xyz.shape[0]).astype(np.float32) if weighted else None
    ref = refgpu.furthestsampling(cu(xyz), cu(off), cu(noff), cu(w) if weighted else None).cpu().numpy()
    assert np.array_equal(oracle.furthestsampling(xyz, off, noff, w), ref)
    assert np.array_equal(F.furthestsampling(cu(xyz), cu(off), cu(noff), weights=cu(w) if weighted else None).cpu().numpy(), ref)


@pytest.mark.parametrize("kind,B,N,M,r,ns", [("tooth", 2, 16384, 1024, 0.1, 32), ("quant", 2, 5000, 300, 0.125, 48),
                                              ("tooth", 2, 4099, 257, 0.2, 64), ("dup", 1, 2000, 100, 0.05, 16)])
def test_ball_query_three_way(cuda, oracle, kind, B, N, M, r, ns):
    xyz = clouds(kind, B, N)
    q = xyz[:, np.random.default_rng(2).choice(N, M, replace=False)].copy()
    q[:, -1] += 9.0
    for which in ("pn2batch", "pn2ext"):
        ref = refgpu.ball_query(r, ns, cu(xyz), cu(q), which).cpu().numpy()
        assert np.array_equal(oracle.ball_query(r, ns, xyz, q), ref)
        assert np.array_equal(F.ball_query(cu(q), cu(xyz), r, ns).cpu().numpy(), ref)


@pytest.mark.parametrize("kind", ["tooth", "dup", "quant"])
@pytest.mark.parametrize("k", [1, 3, 16, 32, 64, 130])
def test_knn_three_way(cuda, oracle, kind, k):
    sizes, ms = [4000, 900, 2500], [300, 200, 257]
    xyz = np.concatenate([clouds(kind, 1, n, seed=i)[0] for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    starts = np.concatenate([[0], off[:-1]])
    q = np.concatenate([xyz[s:s + m] for s, m in zip(starts, ms)]).astype(np.float32)
    ri, rd2 = refgpu.knnquery(k, cu(xyz), cu(q), cu(off), cu(noff))
    ei, _, ed2 = oracle.knnquery(k, xyz, q, off, noff)
    assert np.array_equal(ei, ri.cpu().numpy()) and np.array_equal(ed2, rd2.cpu().numpy()), "oracle vs reference"
    gi, gd2 = F.knnquery(k, cu(xyz), cu(q), cu(off), cu(noff))
    assert torch.equal(gd2, rd2) and torch.equal(gi, ri), "geot vs reference"


@pytest.mark.parametrize("kind,B,n,m", [("tooth", 2, 4096, 512), ("quant", 2, 3000, 700), ("dup", 1, 1000, 300), ("tooth", 1, 50, 2)])
def test_three_nn_three_way(cuda, oracle, kind, B, n, m):
    unknown = clouds(kind, B, n)
    known = unknown[:, np.random.default_rng(8).choice(n, m, replace=False)].copy()
    rd2, ri = refgpu.three_nn(cu(unknown), cu(known))
    _, ei, ed2 = oracle.three_nn(unknown, known)
    assert np.array_equal(ei, ri.cpu().numpy()) and np.array_equal(ed2, rd2.cpu().numpy())
    gd2, gi = F.three_nn(cu(unknown), cu(known))
    assert torch.equal(gi, ri) and torch.equal(gd2, rd2)


def test_group_gather_interpolate_three_way(cuda, oracle):
    rng = np.random.default_rng(12)
    B, C, N, M, ns = 2, 19, 4096, 256, 32
    feat = rng.standard_normal((B, C, N)).astype(np.float32)
    idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
    f, i = cu(feat), cu(idx)
    ref = refgpu.group_points(f, i)
    assert torch.equal(F.group_points(f, i), ref) and np.array_equal(oracle.grouping_operation(feat, idx), ref.cpu().numpy())
    g = cu(rng.standard_normal((B, C, M, ns)).astype(np.float32))
    rg = refgpu.group_points_grad(g, i, N)  # float atomics: order varies, 1e-5 contract
    torch.testing.assert_close(F.group_points_grad(g, F.ScatterPlan(i, B, N), N), rg, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(oracle.grouping_operation_grad(g.cpu().numpy(), idx, N), rg.cpu().numpy(), rtol=1e-5, atol=1e-5)
    i1 = i[:, :, 0].contiguous()
    assert torch.equal(F.gather_points(f, i1), refgpu.gather_points(f, i1))
    m, n = 512, 4096
    pts = cu(rng.standard_normal((B, C, m)).astype(np.float32))
    i3 = cu(rng.integers(0, m, (B, n, 3)).astype(np.int32))
    w = torch.rand((B, n, 3), device="cuda")
    w = (w / w.sum(-1, keepdim=True)).contiguous()
    for which in ("pn2batch", "pn2ext"):
        ro = refgpu.three_interpolate(pts, i3, w, which)
        torch.testing.assert_close(F.three_interpolate(pts, i3, w), ro, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(oracle.three_interpolate(pts.cpu().numpy(), i3.cpu().numpy(), w.cpu().numpy()), ro.cpu().numpy(), rtol=1e-5, atol=1e-6)
    go = cu(rng.standard_normal((B, C, n)).astype(np.float32))
    rgi = refgpu.three_interpolate_grad(go, i3, w, m)
    torch.testing.assert_close(F.three_interpolate_grad(go, w, F.ScatterPlan(i3, B, m), m), rgi, rtol=1e-5, atol=1e-5)


def test_offset_ops_three_way(cuda, oracle):
    rng = np.random.default_rng(13)
    n, m, ns, c, k = 3000, 1200, 16, 32, 3
    inp = cu(rng.standard_normal((n, c)).astype(np.float32))
    idx = cu(rng.integers(0, n, (m, ns)).astype(np.int32))
    assert torch.equal(F.grouping_forward(inp, idx), refgpu.grouping_forward(inp, idx))
    g = cu(rng.standard_normal((m, ns, c)).astype(np.float32))
    torch.testing.assert_close(F.grouping_backward(g, F.ScatterPlan(idx, 1, n), n), refgpu.grouping_backward(g, idx, n), rtol=1e-5, atol=1e-5)
    ik = cu(rng.integers(0, n, (m, k)).astype(np.int32))
    w = torch.rand((m, k), device="cuda")
    torch.testing.assert_close(F.interpolation_forward(inp, ik, w), refgpu.interpolation_forward(inp, ik, w), rtol=1e-5, atol=1e-6)
    go = cu(rng.standard_normal((m, c)).astype(np.float32))
    torch.testing.assert_close(F.interpolation_backward(go, w, F.ScatterPlan(ik, 1, n), n, k), refgpu.interpolation_backward(go, ik, w, n), rtol=1e-5, atol=1e-5)
    idn = cu(rng.integers(0, n, (n, ns)).astype(np.int32))
    in2 = cu(rng.standard_normal((n, c)).astype(np.float32))
    assert torch.equal(F.subtraction_forward(inp, in2, idn), refgpu.subtraction_forward(inp, in2, idn))
    gs = cu(rng.standard_normal((n, ns, c)).astype(np.float32))
    plan = F.ScatterPlan(idn, 1, n)
    g1, g2 = F.subtraction_backward(gs, plan)
    r1, r2 = refgpu.subtraction_backward(idn, gs)
    torch.testing.assert_close(g1, r1, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(g2, r2, rtol=1e-5, atol=1e-5)
    pos = cu(rng.standard_normal((n, ns, c)).astype(np.float32))
    wt = cu(rng.standard_normal((n, ns, 8)).astype(np.float32))
    torch.testing.assert_close(F.aggregation_forward(inp, pos, wt, idn), refgpu.aggregation_forward(inp, pos, wt, idn), rtol=1e-4, atol=1e-5)
    ga = cu(rng.standard_normal((n, c)).astype(np.float32))
    gi, gp, gw = F.aggregation_backward(inp, pos, wt, idn, ga, plan)
    ri, rp, rw = refgpu.aggregation_backward(inp, pos, wt, idn, ga)
    torch.testing.assert_close(gi, ri, rtol=1e-4, atol=1e-5)
    assert torch.equal(gp, rp)
    torch.testing.assert_close(gw, rw, rtol=1e-4, atol=1e-5)
    q = inp[:, :3].contiguous()
    off = cu(np.array([1000, 3000], np.int32)); noff = cu(np.array([400, 1200], np.int32))
    qq = q[:1200].contiguous()
    assert torch.equal(F.ballquery_offset(0.8, 24, q, qq, off, noff), refgpu.ballquery_offset(0.8, 24, q, qq, off, noff))
