"""Strongest pin: the reference's OWN kernels (oracle/_ref, compiled unchanged from /root/reference for
sm_100) run on this B200 next to (a) the CPU oracle and (b) the geot_b200 kernels, on identical inputs.
This is what makes the oracle 'pinned' rather than a bare restatement.  On a CUDA box a missing oracle/_ref is a
FAILURE, not a skip: a build that loses the reference kernels must not turn the strongest pin into a silent pass."""
import numpy as np
import pytest
import torch

from geot_b200 import _lib, synth
from geot_b200 import functional as F
from tests import refgpu

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _reference_kernels_present(cuda):
    missing = [n for n in ("pn2batch", "pn2ext", "pointops", "pointops_full") if not refgpu.available(n)]
    if missing:
        pytest.fail(f"oracle/_ref/libref_{{{','.join(missing)}}}.so missing on a CUDA box: build them where /root/reference exists "
                    "(make -C oracle ref) - they travel with the snapshot")


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
    # sums of signed products: within 1e-5 of the sum of the magnitudes of the terms (tests/tolerance.py), the
    # magnitudes from the CPU oracle evaluated on |inputs|
    from tests.tolerance import assert_sum_close
    h = lambda t: np.abs(t.cpu().numpy())
    ga = cu(rng.standard_normal((n, c)).astype(np.float32))
    idn_h = idn.cpu().numpy()
    assert_sum_close(F.aggregation_forward(inp, pos, wt, idn), refgpu.aggregation_forward(inp, pos, wt, idn),
                     oracle.aggregation_forward(h(inp), h(pos), h(wt), idn_h), what="aggregation_forward vs reference kernel")
    gi, gp, gw = F.aggregation_backward(inp, pos, wt, idn, ga, plan)
    ri, rp, rw = refgpu.aggregation_backward(inp, pos, wt, idn, ga)
    ti, _, tw = oracle.aggregation_backward(h(inp), h(pos), h(wt), idn_h, h(ga))
    assert_sum_close(gi, ri, ti, what="aggregation grad_input vs reference kernel")
    assert torch.equal(gp, rp)
    assert_sum_close(gw, rw, tw, what="aggregation grad_weight vs reference kernel")
    q = inp[:, :3].contiguous()
    off = cu(np.array([1000, 3000], np.int32)); noff = cu(np.array([400, 1200], np.int32))
    qq = q[:1200].contiguous()
    assert torch.equal(F.ballquery_offset(0.8, 24, q, qq, off, noff), refgpu.ballquery_offset(0.8, 24, q, qq, off, noff))


def test_knn_three_way_cfg2_size(cuda):
    """BASELINE config 2's first level: self-kNN, k=16, 8 segments of 16384 points (the grid path) against the
    reference kernel itself - indices bit-equal, squared distances bit-equal."""
    B, N, k = 8, 16384, 16
    xyz = cu(synth.tooth_batch(B, N).reshape(B * N, 3))
    off = cu((np.arange(1, B + 1) * N).astype(np.int32))
    ri, rd = refgpu.knnquery(k, xyz, xyz, off, off)
    gi, gd = F.knnquery(k, xyz, xyz, off, off)
    assert torch.equal(gi, ri) and torch.equal(gd, rd)
    # cross-resolution query of the same level (centres = every 4th point), as the transition-down does
    q = xyz.view(B, N, 3)[:, ::4].reshape(-1, 3).contiguous()
    noff = cu((np.arange(1, B + 1) * (N // 4)).astype(np.int32))
    ri, rd = refgpu.knnquery(k, xyz, q, off, noff)
    gi, gd = F.knnquery(k, xyz, q, off, noff)
    assert torch.equal(gi, ri) and torch.equal(gd, rd)


@pytest.mark.parametrize("radius,ns", [(0.05, 16), (0.1, 32), (0.2, 64)])
def test_ball_query_and_grouping_three_way_cfg3_stage1(cuda, oracle, radius, ns):
    """BASELINE config 3, stage 1 shapes: M=4096 FPS centres of 16384 points, multi-radius ball query (ns 16/32/64),
    grouping fwd + bwd of C=9 channels - ours vs the reference kernels (and the CPU oracle for
    the indices)."""
    B, N, M, C = 2, 16384, 4096, 9
    xyz_h = synth.tooth_batch(B, N, first_cloud=5)
    xyz = cu(xyz_h)
    fidx = F.furthest_point_sampling(xyz, M)
    assert torch.equal(fidx, refgpu.fps(xyz, M))
    new_xyz = F.gather_rows(xyz, fidx)
    ref = refgpu.ball_query(radius, ns, xyz, new_xyz)
    got = F.ball_query(new_xyz, xyz, radius, ns)
    assert torch.equal(got, ref)
    assert np.array_equal(got.cpu().numpy(), oracle.ball_query(radius, ns, xyz_h, new_xyz.cpu().numpy()))
    feat = torch.randn((B, C, N), device="cuda", generator=torch.Generator(device="cuda").manual_seed(ns))
    assert torch.equal(F.group_points(feat, got), refgpu.group_points(feat, got))
    go = torch.randn((B, C, M, ns), device="cuda", generator=torch.Generator(device="cuda").manual_seed(ns + 1))
    ours = F.group_points_grad(go, F.grouping_plan(got, B, N), N)
    from tests.tolerance import assert_sum_close
    terms = refgpu.group_points_grad(go.abs(), got, N)
    assert_sum_close(ours, refgpu.group_points_grad(go, got, N), terms, what="group_points_grad vs reference atomics")
