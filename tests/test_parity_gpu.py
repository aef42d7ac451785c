"""Parity proper: the sm_100a kernels, called through the C ABI (geot_b200.functional -> libgeot_b200.so),
against the CPU oracle on the same seeded inputs.  Bit-exact for indices / copies; 1e-5 relative for
interpolated values and gradients (north_star tolerance)."""
import numpy as np
import pytest
import torch

from geot_b200 import _lib, synth
from geot_b200 import functional as F
from tests.tolerance import assert_sum_close

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-6


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def clouds(kind, B, N, seed=0):
    if kind == "tooth":
        return synth.tooth_batch(B, N, first_cloud=seed)
    return np.stack([synth.adversarial_cloud(N, 100 + seed + i, kind) for i in range(B)])


FPS_CASES = [("tooth", 2, 16384, 1024), ("tooth", 3, 4096, 512), ("tooth", 2, 1000, 100), ("tooth", 2, 5000, 300),
             ("tooth", 1, 16000, 512), ("quant", 2, 6000, 700), ("dup", 2, 3000, 900), ("origin", 2, 9000, 400),
             ("tooth", 2, 37, 37), ("tooth", 1, 1, 1), ("quant", 1, 12289, 64)]


@pytest.mark.parametrize("kind,B,N,M", FPS_CASES)
@pytest.mark.parametrize("variant", ["batch", "ext"])
def test_fps_batched(cuda, oracle, kind, B, N, M, variant):
    xyz = clouds(kind, B, N)
    flags = 0 if variant == "batch" else (_lib.GEOT_FPS_CAP512 | _lib.GEOT_FPS_MAGSKIP)
    exp = oracle.furthest_point_sample(xyz, M, cap=1024 if variant == "batch" else 512, mag_skip=variant == "ext")
    got = F.furthest_point_sampling(cu(xyz), M, flags=flags).cpu().numpy()
    assert np.array_equal(got, exp)
    got = F.furthest_point_sampling(cu(xyz), M, flags=flags | _lib.GEOT_FPS_FORCE_GENERIC).cpu().numpy()
    assert np.array_equal(got, exp), "generic fallback kernel"
    got = F.furthest_point_sampling(cu(xyz), M, flags=flags | _lib.GEOT_FPS_NO_PRUNE).cpu().numpy()
    assert np.array_equal(got, exp), "un-pruned one-CTA kernel"
    got = F.furthest_point_sampling(cu(xyz), M, flags=flags | _lib.GEOT_FPS_SPEC).cpu().numpy()
    assert np.array_equal(got, exp), "pruned kernel, manager guessing one round ahead"
    got = F.furthest_point_sampling(cu(xyz), M, flags=flags & ~_lib.GEOT_FPS_SPEC).cpu().numpy()
    assert np.array_equal(got, exp), "pruned kernel, manager in lock-step"


def test_ball_query_prebuilt_and_fps_ball_query(cuda):
    """GEOT_BQ_BUILD_ONLY + GEOT_BQ_PREBUILT (the binning run ahead of the queries) and F.fps_ball_query (the same with
    the binning on a side stream while FPS runs) give the indices of the plain calls; a workspace that still holds
    another batch's bins is rebuilt, not reused; small clouds (no binned path) pass through."""
    ws = None
    for (B, N, M, r, ns, first, flags) in [(3, 16384, 512, 0.1, 32, 0, 0), (2, 8192, 300, 0.05, 16, 7, _lib.GEOT_BQ_FORCE_GRID),
                                           (3, 16384, 512, 0.2, 64, 11, 0), (2, 2000, 100, 0.2, 16, 3, 0), (2, 3000, 64, 0.3, 8, 5, _lib.GEOT_BQ_FORCE_GRID)]:
        xyz = cu(synth.tooth_batch(B, N, first_cloud=first))
        idx0 = F.furthest_point_sampling(xyz, M)
        cen0 = F.gather_rows(xyz, idx0)
        exp = F.ball_query(cen0, xyz, r, ns, flags=flags)
        ws = F.ball_query_prepare(xyz, r, ns, workspace=ws, flags=flags)
        got = F.ball_query(cen0, xyz, r, ns, workspace=ws, flags=flags | _lib.GEOT_BQ_PREBUILT)
        assert torch.equal(got, exp), (B, N, M, r, ns)
        idx1, cen1, nbr1 = F.fps_ball_query(xyz, M, r, ns, bq_flags=flags, workspace=ws)
        assert torch.equal(idx1, idx0) and torch.equal(cen1, cen0) and torch.equal(nbr1, exp), (B, N, M, r, ns)
    with pytest.raises(_lib.GeotError):
        F.ball_query(cen0, xyz, r, ns, workspace=ws, flags=_lib.GEOT_BQ_PREBUILT | _lib.GEOT_BQ_BUILD_ONLY)


def test_fps_large_n_uses_fallback(cuda, oracle):
    xyz = clouds("tooth", 2, 20000)
    assert np.array_equal(F.furthest_point_sampling(cu(xyz), 200).cpu().numpy(), oracle.furthest_point_sample(xyz, 200))


def test_fps_more_samples_than_points(cuda, oracle):
    xyz = clouds("tooth", 2, 50)
    assert np.array_equal(F.furthest_point_sampling(cu(xyz), 80).cpu().numpy(), oracle.furthest_point_sample(xyz, 80))


@pytest.mark.parametrize("weighted", [False, True])
def test_fps_offset_ragged(cuda, oracle, weighted):
    sizes, ms = [5000, 1200, 16384, 333], [300, 100, 1024, 333]
    xyz = np.concatenate([synth.tooth_cloud(n, 50 + i) for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    w = np.random.default_rng(1).random(xyz.shape[0]).astype(np.float32) if weighted else None
    if weighted:
        w[::7] = 0.0  # exercises max(w, 1e-12)
    exp = oracle.furthestsampling(xyz, off, noff, w)
    got = F.furthestsampling(cu(xyz), cu(off), cu(noff), weights=cu(w) if weighted else None).cpu().numpy()
    assert np.array_equal(got, exp)


def test_group_points_row_staged_vs_direct(cuda, oracle):
    """E >= 2N takes the row-staged kernel (cp.async.bulk + smem gather); geot_group_points_config(1) forces the
    direct L2 gather.  Both must equal the oracle bit for bit."""
    rng = np.random.default_rng(3)
    for (B, C, N, M, ns) in [(2, 5, 4096, 512, 32), (1, 3, 16384, 1024, 32), (2, 7, 1024, 256, 16), (1, 2, 50000, 4096, 32)]:
        feat = rng.standard_normal((B, C, N)).astype(np.float32)
        idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
        exp = oracle.grouping_operation(feat, idx)
        try:
            for direct in (0, 1):
                _lib.lib().geot_group_points_config(direct)
                got = F.group_points(cu(feat), cu(idx)).cpu().numpy()
                assert np.array_equal(got, exp), f"direct={direct} shape={(B, C, N, M, ns)}"
        finally:
            _lib.lib().geot_group_points_config(0)


def test_group_points_grad_plan_resident_vs_direct(cuda, oracle):
    """E <= 65535 takes the plan-resident backward (16-bit plan + gradient row in shared memory; one CTA per SM over an
    equal share of the (cloud, channel) rows, the next row prefetched into L2, for plans in pieces the row summed in two
    halves as they land); modes 9-13 select the variants one by one (9 = the round-2 kernel: a grid of CTAs, whole rows,
    no prefetch), mode 1 the direct kernel.  All sum in the same plan order, so they must agree bit for bit; they match
    the oracle's double sum to 1e-5.
    Random lists (dense targets: the general path), ball-query lists (few non-empty targets: the compact task list) and
    a row length that is not a multiple of 8 (the halves differ in length)."""
    rng = np.random.default_rng(4)
    for (B, C, N, M, ns, ball) in [(2, 6, 4096, 512, 32, 0), (1, 5, 16384, 1024, 32, 0), (2, 3, 1000, 250, 16, 0), (2, 3, 1000, 251, 12, 0),
                                   (2, 9, 16384, 1024, 32, 0.1), (3, 4, 4096, 509, 20, 0.3),
                                   (2, 3, 20000, 700, 32, 0), (2, 5, 20000, 700, 32, 0.1), (1, 1, 64, 16, 4, 0)]:  # > 16 targets per thread; one row
        if ball:
            xyz = clouds("tooth", B, N)
            idx = F.ball_query(cu(xyz[:, rng.choice(N, M, replace=False)].copy()), cu(xyz), ball, ns).cpu().numpy()
        else:
            idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
            idx[:, : M // 8] = idx[:, : M // 8, :1]  # padded rows: long segments for a few targets
        g = rng.standard_normal((B, C, M, ns)).astype(np.float32)
        exp = oracle.grouping_operation_grad(g, idx, N)
        plan = F.ScatterPlan(cu(idx), B, N)
        res = []
        try:
            for mode in (0, 9, 10, 11, 12, 13, 1):
                _lib.lib().geot_group_points_config(mode)
                res.append(F.group_points_grad(cu(g), plan, N))
        finally:
            _lib.lib().geot_group_points_config(0)
        assert all(torch.equal(r, res[-1]) for r in res[:-1]), (B, C, N, M, ns, ball)
        assert_sum_close(res[0], exp, oracle.grouping_operation_grad(np.abs(g), idx, N), what="group_points_grad")


BALL_CASES = [("tooth", 2, 16384, 1024, 0.1, 32), ("tooth", 2, 4099, 257, 0.2, 64), ("tooth", 3, 1000, 64, 0.05, 16),
              ("quant", 2, 5000, 300, 0.125, 48), ("dup", 1, 2000, 100, 0.1, 8), ("tooth", 1, 30, 5, 0.5, 40)]


@pytest.mark.parametrize("kind,B,N,M,r,ns", BALL_CASES)
def test_ball_query(cuda, oracle, kind, B, N, M, r, ns):
    xyz = clouds(kind, B, N)
    q = xyz[:, np.random.default_rng(2).choice(N, M, replace=False)].copy()
    q[:, -1] += 9.0  # an empty ball -> all-zero row
    exp = oracle.ball_query(r, ns, xyz, q)
    out = torch.full((B, M, ns), -7, dtype=torch.int32, device="cuda")  # every slot must be written
    got = F.ball_query(cu(q), cu(xyz), r, ns, out=out).cpu().numpy()
    assert np.array_equal(got, exp)
    assert not exp[:, -1].any()
    out.fill_(-7)
    got = F.ball_query(cu(q), cu(xyz), r, ns, out=out, use_grid=False).cpu().numpy()
    assert np.array_equal(got, exp), "index-order scan"
    out.fill_(-7)
    got = F.ball_query(cu(q), cu(xyz), r, ns, out=out, flags=_lib.GEOT_BQ_FORCE_GRID).cpu().numpy()
    assert np.array_equal(got, exp), "grid path"


@pytest.mark.parametrize("r", [0.003, 0.02, 0.1, 0.4, 3.0])
def test_ball_query_radius_sweep_and_outside_queries(cuda, oracle, r):
    """tiny radius (grid capped at 32 cells per axis), huge radius (one cell), queries outside the cloud's box."""
    xyz = clouds("tooth", 2, 8192)
    rng = np.random.default_rng(5)
    q = np.concatenate([xyz[:, :200], xyz[:, 200:300] + rng.normal(0, r, (2, 100, 3)).astype(np.float32),
                        rng.uniform(-2, 2, (2, 100, 3)).astype(np.float32)], 1)
    exp = oracle.ball_query(r, 32, xyz, q)
    for flags in (0, _lib.GEOT_BQ_FORCE_GRID):
        got = F.ball_query(cu(q), cu(xyz), r, 32, flags=flags).cpu().numpy()
        assert np.array_equal(got, exp), f"flags={flags}"


@pytest.mark.parametrize("B,N,M,r,ns", [(2, 16384, 700, 0.1, 32), (1, 16384, 512, 0.05, 16), (2, 20000, 300, 0.2, 64), (1, 65536, 256, 0.05, 8),
                                        (2, 4100, 200, 0.3, 5), (1, 2047, 100, 0.2, 32)])
def test_ball_query_index_levels(cuda, oracle, B, N, M, r, ns):
    """The grid path bins the points in 1, 2, 4 or 8 index levels and a query stops after the first level that fills its
    row: every admissible level count (GEOT_BQ_LEVELS; the kernel caps it by the cell budget and N >= 2048 per level), the
    automatic choice and the index-order scan give the same rows; sparse queries walk all levels, N not a multiple of
    the level size, N above the register-cached build's 16384."""
    xyz = clouds("tooth", B, N)
    rng = np.random.default_rng(N + ns)
    q = np.concatenate([xyz[:, rng.choice(N, M - 20, replace=False)], rng.uniform(-1, 1, (B, 20, 3)).astype(np.float32)], 1)
    scan = F.ball_query(cu(q), cu(xyz), r, ns, use_grid=False)
    if B * M * N <= 40_000_000:
        assert np.array_equal(scan.cpu().numpy(), oracle.ball_query(r, ns, xyz, q))
    assert torch.equal(F.ball_query(cu(q), cu(xyz), r, ns), scan)
    for lev in (1, 2, 4, 8):
        out = torch.full((B, M, ns), -7, dtype=torch.int32, device="cuda")
        F.ball_query(cu(q), cu(xyz), r, ns, out=out, flags=_lib.GEOT_BQ_FORCE_GRID | _lib.GEOT_BQ_LEVELS(lev))
        assert torch.equal(out, scan), f"levels={lev}: {(out != scan).sum().item()} entries differ"


def test_fps_understated_n_max_marks_the_segment(cuda, oracle):
    """geot_furthestsampling sizes its kernels from the caller's n_max (pointops_cuda.furthestsampling_cuda takes it as an
    argument).  A segment longer than promised cannot be sampled by the kernel that was picked: its indices come back as
    -1 instead of uninitialised memory; the other segments are unaffected."""
    sizes, ms = [3000, 9000], [50, 64]
    xyz = np.concatenate([synth.tooth_cloud(n, 5 + i) for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    good = F.furthestsampling(cu(xyz), cu(off), cu(noff))
    assert np.array_equal(good.cpu().numpy(), oracle.furthestsampling(xyz, off, noff))
    for flags in (0, _lib.GEOT_FPS_NO_PRUNE):
        out = torch.full((114,), 7, dtype=torch.int32, device="cuda")
        F.furthestsampling(cu(xyz), cu(off), cu(noff), n_max=4096, m_total=114, out=out, flags=flags)
        assert torch.equal(out[:50], good[:50]) and bool((out[50:] == -1).all())


@pytest.mark.parametrize("kind", ["dup", "quant", "tooth"])
@pytest.mark.parametrize("k", [1, 5, 16, 33, 64])
def test_knnquery_ties_by_index(cuda, kind, k):
    """GEOT_KNN_TIES_BY_INDEX (knn_point / knn_cuda.KNN semantics: no heap replay): rows ascending in (distance, index);
    the same rows from the grid and the brute-force kernels (visiting order does not matter); the k best are a prefix
    of the k+7 best; the distances are the ones the reference's heap order returns (only the order of equal ones and the
    choice among equals at the boundary may differ) - on clouds with many exact ties (duplicates, quantised coordinates)."""
    B, N, M = 2, 6000, 700
    xyz = torch.from_numpy(clouds(kind, B, N, seed=3)).cuda().view(-1, 3).contiguous()
    q = xyz.view(B, N, 3)[:, :M].reshape(-1, 3).contiguous()
    T = _lib.GEOT_KNN_TIES_BY_INDEX
    idx, d2 = F.knnquery(k, xyz, q, None, None, n_per=N, m_per=M, flags=T)
    idx_b, d2_b = F.knnquery(k, xyz, q, None, None, n_per=N, m_per=M, flags=T | _lib.GEOT_KNN_NO_GRID)
    assert torch.equal(idx, idx_b) and torch.equal(d2, d2_b)
    same = d2[:, 1:] == d2[:, :-1]
    assert bool((d2[:, 1:] >= d2[:, :-1]).all()) and bool((idx[:, 1:] > idx[:, :-1])[same].all())
    if kind != "tooth" and k >= 5:
        assert int(same.sum()) > 0
    idx7, d27 = F.knnquery(k + 7, xyz, q, None, None, n_per=N, m_per=M, flags=T)
    assert torch.equal(idx7[:, :k], idx) and torch.equal(d27[:, :k], d2)
    _, d2_heap = F.knnquery(k, xyz, q, None, None, n_per=N, m_per=M)
    assert torch.equal(d2_heap, d2)
    # the distances are exact: recomputed from the returned indices in double they agree to fp32 rounding
    p, c = xyz[idx.long()].double(), q.double().unsqueeze(1)
    assert float(((p - c).square().sum(-1) - d2.double()).abs().max()) <= 1e-6 * float(d2.max()) + 1e-12


def test_ball_query_offset(cuda, oracle):
    sizes, ms = [3000, 700, 5001], [100, 33, 250]
    xyz = np.concatenate([synth.tooth_cloud(n, 60 + i) for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    starts = np.concatenate([[0], off[:-1]])
    q = np.concatenate([xyz[s:s + m] for s, m in zip(starts, ms)])
    exp = oracle.ballquery(0.15, 24, xyz, q, off, noff)
    got = F.ballquery_offset(0.15, 24, cu(xyz), cu(q), cu(off), cu(noff)).cpu().numpy()
    assert np.array_equal(got, exp)


@pytest.mark.parametrize("kind", ["tooth", "dup", "quant"])
@pytest.mark.parametrize("k", [1, 3, 16, 31, 32, 64, 100, 130])
def test_knnquery_offset(cuda, oracle, kind, k):
    sizes, ms = [4000, 900, 2500, 40], [300, 200, 257, 40]
    xyz = np.concatenate([clouds(kind, 1, n, seed=i)[0] for i, n in enumerate(sizes)])
    off, noff = np.cumsum(sizes).astype(np.int32), np.cumsum(ms).astype(np.int32)
    starts = np.concatenate([[0], off[:-1]])
    q = np.concatenate([xyz[s:s + m] + (0 if kind != "tooth" else 1e-3) for s, m in zip(starts, ms)]).astype(np.float32)
    ei, _, ed2 = oracle.knnquery(k, xyz, q, off, noff)
    for flags in (0, _lib.GEOT_KNN_NO_GRID):  # segments 0 and 2 take the grid path unless it is switched off
        gi, gd2 = F.knnquery(k, cu(xyz), cu(q), cu(off), cu(noff), flags=flags)
        assert np.array_equal(gd2.cpu().numpy(), ed2)
        assert np.array_equal(gi.cpu().numpy(), ei)


@pytest.mark.parametrize("kind", ["tooth", "uniform", "dup", "quant"])
@pytest.mark.parametrize("k", [1, 8, 16, 33, 100])
def test_knnquery_grid_vs_scan(cuda, kind, k):
    """Grid path (segments >= 2048 points) against the brute-force scan of the same library at sizes the CPU
    oracle does not finish quickly: self-queries (distance 0), queries off the surface and far outside the
    bounding box (they fall through both rings and are redone exactly), ragged segment sizes."""
    rng = np.random.default_rng(k)
    sizes = [16384, 2048, 7001, 1500]
    if kind == "uniform":
        xyz = np.concatenate([rng.random((n, 3), dtype=np.float32) * np.float32(1 + i) for i, n in enumerate(sizes)])
    else:
        xyz = np.concatenate([clouds(kind, 1, n, seed=10 + i)[0] for i, n in enumerate(sizes)])
    off = np.cumsum(sizes).astype(np.int32)
    q = xyz.copy()
    q[::7] += rng.normal(0, 0.05, (len(q[::7]), 3)).astype(np.float32)
    q[5::101] += np.float32(3.0)
    q[9::211] *= np.float32(-20.0)
    a_i, a_d = F.knnquery(k, cu(xyz), cu(q), cu(off), cu(off))
    b_i, b_d = F.knnquery(k, cu(xyz), cu(q), cu(off), cu(off), flags=_lib.GEOT_KNN_NO_GRID)
    assert torch.equal(a_d, b_d)
    assert torch.equal(a_i, b_i)


def test_knnquery_grid_degenerate_clouds(cuda, oracle):
    """All points identical / on a line / on a plane: the grid collapses to 1 cell in the flat axes."""
    n, k = 3000, 9
    rng = np.random.default_rng(0)
    same = np.tile(np.float32([0.3, -1.0, 2.0]), (n, 1))
    line = np.zeros((n, 3), np.float32); line[:, 0] = rng.random(n, dtype=np.float32)
    plane = rng.random((n, 3), dtype=np.float32); plane[:, 2] = 0.5
    for xyz in (same, line, plane):
        off = np.array([n], np.int32)
        q = xyz[:200] + np.float32(1e-3)
        noff = np.array([200], np.int32)
        ei, _, ed2 = oracle.knnquery(k, xyz, q, off, noff)
        gi, gd2 = F.knnquery(k, cu(xyz), cu(q), cu(off), cu(noff))
        assert np.array_equal(gd2.cpu().numpy(), ed2) and np.array_equal(gi.cpu().numpy(), ei)


def test_knnquery_batched_layout(cuda, oracle):
    b, n, m, k = 3, 2048, 512, 16
    xyz = clouds("tooth", b, n)
    q = xyz[:, :m] + np.float32(2e-3)
    off = (np.arange(1, b + 1) * n).astype(np.int32)
    noff = (np.arange(1, b + 1) * m).astype(np.int32)
    ei, _, ed2 = oracle.knnquery(k, xyz.reshape(-1, 3), q.reshape(-1, 3), off, noff)
    gi, gd2 = F.knnquery(k, cu(xyz.reshape(-1, 3)), cu(q.reshape(-1, 3)), None, None, n_per=n, m_per=m)
    assert np.array_equal(gi.cpu().numpy(), ei) and np.array_equal(gd2.cpu().numpy(), ed2)


GROUP_CASES = [(2, 3, 16384, 1024, 32), (2, 64, 4096, 256, 32), (1, 5, 1000, 33, 7), (3, 17, 500, 64, 16), (1, 1, 10, 1, 1)]


@pytest.mark.parametrize("B,C,N,M,ns", GROUP_CASES)
def test_group_points_fwd_bwd(cuda, oracle, B, C, N, M, ns):
    rng = np.random.default_rng(5)
    feat = rng.standard_normal((B, C, N)).astype(np.float32)
    idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
    idx[:, 0, :] = idx[:, 0, :1]  # a padded row: one target hit ns times
    out = F.group_points(cu(feat), cu(idx)).cpu().numpy()
    assert np.array_equal(out, oracle.grouping_operation(feat, idx))
    g = rng.standard_normal((B, C, M, ns)).astype(np.float32)
    plan = F.ScatterPlan(cu(idx), B, N)
    got = F.group_points_grad(cu(g), plan, N)
    np.testing.assert_allclose(got.cpu().numpy(), oracle.grouping_operation_grad(g, idx, N), rtol=RTOL, atol=ATOL)
    again = F.group_points_grad(cu(g), F.ScatterPlan(cu(idx), B, N), N)
    assert torch.equal(got, again), "backward must be bit-reproducible"


def test_scatter_plan_is_stable_csr(cuda):
    rng = np.random.default_rng(6)
    B, N, E = 3, 777, 5000
    idx = rng.integers(0, N, (B, E)).astype(np.int32)
    idx[0, :2000] = 5  # one very long segment
    plan = F.ScatterPlan(cu(idx), B, N)
    rp, pm = plan.row_ptr().cpu().numpy(), plan.perm().cpu().numpy()
    for b in range(B):
        order = np.argsort(idx[b], kind="stable")
        assert np.array_equal(pm[b], order.astype(np.int32))
        assert np.array_equal(rp[b], np.concatenate([[0], np.cumsum(np.bincount(idx[b], minlength=N))]))


@pytest.mark.parametrize("B,C,N,M,ns", [(2, 5, 4096, 1024, 64), (2, 9, 16384, 4096, 32), (1, 4, 8192, 1536, 32), (3, 3, 2048, 2048, 48)])
def test_group_points_grad_pieces(cuda, oracle, B, C, N, M, ns):
    """Neighbour lists too long for one shared-memory pass (M*ns > 65535 or row + plan > 227 KB) are reduced piece
    by piece (geot_scatter_plan_pieces / geot_group_points_grad_pieces): against the oracle, against the
    single-plan direct kernel, and bit-for-bit against the piece-ordered direct kernel (same summation order)."""
    xyz = clouds("tooth", B, N)
    q = xyz[:, :M]
    nbr = F.ball_query(cu(q), cu(xyz), 0.2, ns)
    E = M * ns
    Q = int(_lib.lib().geot_scatter_plan_pieces(N, E))
    assert Q > 1, (N, E, Q)
    g = torch.randn(B, C, M, ns, device="cuda")
    plan_q = F.grouping_plan(nbr, B, N)
    assert plan_q.pieces == Q
    got = F.group_points_grad(g, plan_q, N)
    single = F.group_points_grad(g, F.ScatterPlan(nbr, B, N), N)
    exp = oracle.grouping_operation_grad(g.cpu().numpy(), nbr.cpu().numpy(), N)
    terms = oracle.grouping_operation_grad(np.abs(g.cpu().numpy()), nbr.cpu().numpy(), N)
    assert_sum_close(got, exp, terms, what="group_points_grad in pieces")
    assert_sum_close(single, exp, terms, what="group_points_grad single plan")
    _lib.lib().geot_group_points_config(1)
    try:
        direct = F.group_points_grad(g, plan_q, N)
    finally:
        _lib.lib().geot_group_points_config(0)
    assert torch.equal(got, direct)
    assert torch.equal(got, F.group_points_grad(g, plan_q, N))
    # through autograd (layers.GroupingOperation picks the pieces plan itself)
    from geot_b200 import layers
    feat = torch.randn(B, C, N, device="cuda", requires_grad=True)
    layers.grouping_operation(feat, nbr).backward(g)
    assert torch.equal(feat.grad, got)


def test_scatter_plan_large_universe_heavy_segments(cuda):
    """E > 2^20: the CTA bitmap sort of long segments walks the source universe in two chunks."""
    rng = np.random.default_rng(7)
    B, N, E = 1, 3000, (1 << 20) + 70001
    idx = rng.integers(0, N, (B, E)).astype(np.int32)
    idx[0, rng.random(E) < 0.2] = 0      # ~2e5 sources on target 0
    idx[0, -5000:] = N - 1               # a long segment made of the highest sources
    plan = F.ScatterPlan(cu(idx), B, N)
    rp, pm = plan.row_ptr().cpu().numpy(), plan.perm().cpu().numpy()
    assert np.array_equal(pm[0], np.argsort(idx[0], kind="stable").astype(np.int32))
    assert np.array_equal(rp[0], np.concatenate([[0], np.cumsum(np.bincount(idx[0], minlength=N))]))


@pytest.mark.parametrize("B,C,N,M", [(2, 3, 16384, 1024), (2, 67, 999, 130), (1, 4, 8, 8)])
def test_gather_points_fwd_bwd(cuda, oracle, B, C, N, M):
    rng = np.random.default_rng(7)
    feat = rng.standard_normal((B, C, N)).astype(np.float32)
    idx = rng.integers(0, N, (B, M)).astype(np.int32)
    assert np.array_equal(F.gather_points(cu(feat), cu(idx)).cpu().numpy(), oracle.gather_operation(feat, idx))
    g = rng.standard_normal((B, C, M)).astype(np.float32)
    got = F.gather_points_grad(cu(g), F.ScatterPlan(cu(idx), B, N), N).cpu().numpy()
    np.testing.assert_allclose(got, oracle.gather_operation_grad(g, idx, N), rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("kind,B,n,m", [("tooth", 2, 4096, 512), ("tooth", 2, 16000, 8192), ("quant", 2, 3000, 700),
                                         ("dup", 1, 1000, 300), ("tooth", 2, 100, 2), ("tooth", 1, 513, 1),
                                         ("tooth", 1, 30000, 16000), ("quant", 1, 21000, 20000)])  # known cloud searched in 2 / 3 parts
def test_three_nn(cuda, oracle, kind, B, n, m):
    unknown = clouds(kind, B, n)
    known = unknown[:, np.random.default_rng(8).choice(n, m, replace=False)].copy() if m <= n else clouds(kind, B, m, 3)
    _, ei, ed2 = oracle.three_nn(unknown, known)
    gd2, gi = F.three_nn(cu(unknown), cu(known))
    assert np.array_equal(gi.cpu().numpy(), ei)
    assert np.array_equal(gd2.cpu().numpy(), ed2)


@pytest.mark.parametrize("kind", ["tooth", "uniform", "dup", "quant", "plane", "same"])
@pytest.mark.parametrize("B,n,m", [(3, 16384, 1024), (2, 5000, 64), (2, 3001, 8192), (1, 2048, 4097), (2, 700, 333),
                                   (2, 4000, 16000), (1, 2500, 27649)])  # 2 and 4 parts: ties across parts resolved by GLOBAL index
def test_three_nn_grid_vs_scan(cuda, kind, B, n, m):
    """Grid search (64 <= m <= 8192) against the brute-force kernel of the same library (exact on its own, see
    test_three_nn): duplicates and quantised coordinates (ties resolved by index), unknown points far outside the
    known cloud, flat and single-point known clouds."""
    rng = np.random.default_rng(n + m)
    if kind == "uniform":
        known = rng.random((B, m, 3), dtype=np.float32)
        unknown = rng.random((B, n, 3), dtype=np.float32) * np.float32(1.2) - np.float32(0.1)
    elif kind == "plane":
        known = rng.random((B, m, 3), dtype=np.float32); known[..., 1] = np.float32(0.25)
        unknown = rng.random((B, n, 3), dtype=np.float32)
    elif kind == "same":
        known = np.tile(np.float32([0.5, 0.5, 0.5]), (B, m, 1))
        unknown = rng.random((B, n, 3), dtype=np.float32)
    else:
        big = clouds(kind, B, max(n, m))
        known = np.ascontiguousarray(big[:, rng.permutation(max(n, m))[:m]])
        unknown = np.ascontiguousarray(big[:, :n])
    unknown[:, 3::97] += np.float32(7.0)
    unknown[:, 5::131] *= np.float32(-3.0)
    _lib.lib().geot_group_points_config(2)  # grid even where the work-size heuristic would pick the brute-force kernel
    try:
        gd, gi = F.three_nn(cu(unknown), cu(known))
        _lib.lib().geot_group_points_config(1)
        sd, si = F.three_nn(cu(unknown), cu(known))
    finally:
        _lib.lib().geot_group_points_config(0)
    dd, di = F.three_nn(cu(unknown), cu(known))  # default dispatch
    assert torch.equal(di, si) and torch.equal(dd, sd)
    assert torch.equal(gi, si)
    assert torch.equal(gd, sd)


@pytest.mark.parametrize("B,C,m,n", [(2, 384, 512, 4096), (2, 16, 8192, 16000), (1, 3, 5, 9)])
def test_three_interpolate_fwd_bwd(cuda, oracle, B, C, m, n):
    rng = np.random.default_rng(9)
    feat = rng.standard_normal((B, C, m)).astype(np.float32)
    idx = rng.integers(0, m, (B, n, 3)).astype(np.int32)
    w = rng.random((B, n, 3)).astype(np.float32)
    w /= w.sum(-1, keepdims=True)
    out = F.three_interpolate(cu(feat), cu(idx), cu(w)).cpu().numpy()
    np.testing.assert_allclose(out, oracle.three_interpolate(feat, idx, w), rtol=RTOL, atol=ATOL)
    g = rng.standard_normal((B, C, n)).astype(np.float32)
    got = F.three_interpolate_grad(cu(g), cu(w), F.ScatterPlan(cu(idx), B, m), m)
    np.testing.assert_allclose(got.cpu().numpy(), oracle.three_interpolate_grad(g, idx, w, m), rtol=RTOL, atol=2e-6)
    assert torch.equal(got, F.three_interpolate_grad(cu(g), cu(w), F.ScatterPlan(cu(idx), B, m), m))


@pytest.mark.parametrize("B,C,m,n", [(3, 64, 1024, 16384), (2, 70, 512, 4096), (2, 5, 4096, 8192), (1, 4, 100, 1024), (2, 9, 333, 2000),
                                     (1, 33, 16000, 100000), (2, 3, 64, 512), (2, 16, 1024, 4098),
                                     (40, 24, 256, 2048), (1, 4, 64, 4), (5, 37, 1024, 6148), (64, 64, 1024, 16384), (1, 1100, 128, 512)])
def test_three_interpolate_interleaved_vs_rows_vs_direct(cuda, oracle, B, C, m, n):
    """The persistent channel-interleaved forward kernel (default when n % 4 == 0, C >= 4, m <= 4096, m % 4 == 0) against
    its grid-per-pair form (mode 4), the plain row-staged kernel (mode 3), the direct kernel (mode 1) and the oracle:
    partial last channel group (C % 4 != 0), partial last CTA group (C % G != 0), m not a multiple of 4, shapes that
    fall back (m > 4096, C < 4, n % 4 != 0), more (cloud, channel group) pairs than CTAs, fewer work items than CTAs,
    a CTA's run starting and ending inside a pair.  Same FMA order in all kernels: identical bits."""
    g = torch.Generator(device="cuda").manual_seed(C * 7 + m)
    feat = torch.randn(B, C, m, device="cuda", generator=g)
    idx = torch.randint(0, m, (B, n, 3), device="cuda", generator=g, dtype=torch.int32)
    w = torch.rand(B, n, 3, device="cuda", generator=g)
    w = (w / w.sum(-1, keepdim=True)).contiguous()
    got = F.three_interpolate(feat, idx, w)
    res = []
    try:
        for mode in (2, 4, 3, 1):     # 2: the persistent kernel whatever the launch size
            _lib.lib().geot_group_points_config(mode)
            res.append(F.three_interpolate(feat, idx, w))
    finally:
        _lib.lib().geot_group_points_config(0)
    assert all(torch.equal(got, r) for r in res)
    assert torch.equal(got, F.three_interpolate(feat, idx, w))
    if B * C * n <= 4_000_000:
        exp = oracle.three_interpolate(feat.cpu().numpy(), idx.cpu().numpy(), w.cpu().numpy())
        np.testing.assert_allclose(got.cpu().numpy(), exp, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("B,C,m,n", [(3, 5, 1024, 16384), (2, 64, 700, 4096), (1, 13, 2048, 6002), (2, 7, 4096, 12288),
                                     (2, 4, 1024, 30000), (1, 9, 6000, 8192), (2, 3, 512, 4096), (4, 16, 100, 20000), (1, 6, 8192, 16000),
                                     (8, 200, 1024, 5000), (3, 17, 1, 2048), (2, 33, 1000, 1028)])
def test_three_interpolate_grad_rows_vs_direct(cuda, oracle, B, C, m, n):
    """The backward kernels - streamed (thread per target, source tiles through shared memory; mode 8 forces it, default for
    m <= 1024 when the launch fills the machine), row-resident (mode 5: gradient rows in shared memory, packed {source,
    weight} edges, lane groups per target) and direct L2 gather (mode 1) - against the oracle, on real three_nn indices:
    skewed segment lengths, targets that no point selects (half of the known points are moved away), channel counts that
    leave a partial last CTA, a last source tile shorter than the others, m > 1024 (no streamed kernel), rows too long
    for more than one per CTA, few edges per target, n % 4 != 0 (direct kernel).  The streamed kernel sums a target in
    plan order like the direct one: identical bits."""
    unknown = clouds("tooth", B, n)
    known = np.ascontiguousarray(unknown[:, :m])
    known[:, m // 2:] += np.float32(5.0)  # half of the known points are far away: empty segments
    d2, idx = F.three_nn(cu(unknown), cu(known))
    w = 1.0 / (d2 + 1e-8)
    w = (w / w.sum(-1, keepdim=True)).contiguous()
    g = torch.randn(B, C, n, device="cuda")
    plan = F.ScatterPlan(idx, B, m)
    got = F.three_interpolate_grad(g, w, plan, m)
    res = {}
    try:
        for name, mode in (("direct", 1), ("row-resident", 5), ("streamed", 8)):
            _lib.lib().geot_group_points_config(mode)
            res[name] = F.three_interpolate_grad(g, w, plan, m)
    finally:
        _lib.lib().geot_group_points_config(0)
    exp = oracle.three_interpolate_grad(g.cpu().numpy(), idx.cpu().numpy(), w.cpu().numpy(), m)
    terms = oracle.three_interpolate_grad(np.abs(g.cpu().numpy()), idx.cpu().numpy(), w.cpu().numpy(), m)
    for name, r in res.items():
        assert_sum_close(r, exp, terms, what="three_interpolate_grad " + name)
    assert_sum_close(got, exp, terms, what="three_interpolate_grad default")
    if m <= 1024 and n % 4 == 0 and n >= 1024:
        assert torch.equal(res["streamed"], res["direct"]), "streamed == direct, bit for bit"
    assert torch.equal(got, F.three_interpolate_grad(g, w, plan, m)), "run-to-run identical"
    assert torch.equal(got, F.three_interpolate_grad(g, w, F.ScatterPlan(idx, B, m), m)), "a fresh plan gives the same bits"


def _check_aggregation(oracle, inp, pos, wt, idn, ga, plan):
    """aggregation fwd / bwd against the oracle's double sums, each within 1e-5 of the sum of the magnitudes of its
    terms (the oracle evaluated on |inputs|): out = sum_s (inp[idx]+pos)*w, grad_input = scatter of ga*w,
    grad_position = ga*w (one product: exact), grad_weight = sum_c ga*(inp[idx]+pos)."""
    a = np.abs
    assert_sum_close(F.aggregation_forward(cu(inp), cu(pos), cu(wt), cu(idn)), oracle.aggregation_forward(inp, pos, wt, idn),
                     oracle.aggregation_forward(a(inp), a(pos), a(wt), idn), what="aggregation_forward")
    gi, gp, gw = F.aggregation_backward(cu(inp), cu(pos), cu(wt), cu(idn), cu(ga), plan)
    ei, ep, ew = oracle.aggregation_backward(inp, pos, wt, idn, ga)
    ti, _, tw = oracle.aggregation_backward(a(inp), a(pos), a(wt), idn, a(ga))
    assert_sum_close(gi, ei, ti, what="aggregation grad_input")
    assert np.array_equal(gp.cpu().numpy(), ep)
    assert_sum_close(gw, ew, tw, what="aggregation grad_weight")
    return gi, gp, gw


@pytest.mark.parametrize("c", [32, 7])
def test_offset_channel_last_ops(cuda, oracle, c):
    rng = np.random.default_rng(10)
    n, m, ns, k = 3000, 1200, 16, 3
    inp = rng.standard_normal((n, c)).astype(np.float32)
    idx = rng.integers(0, n, (m, ns)).astype(np.int32)
    assert np.array_equal(F.grouping_forward(cu(inp), cu(idx)).cpu().numpy(), oracle.grouping(inp, idx))
    g = rng.standard_normal((m, ns, c)).astype(np.float32)
    got = F.grouping_backward(cu(g), F.ScatterPlan(cu(idx), 1, n), n).cpu().numpy()
    np.testing.assert_allclose(got, oracle.grouping_grad(g, idx, n), rtol=RTOL, atol=ATOL)
    ik = rng.integers(0, n, (m, k)).astype(np.int32)
    w = rng.random((m, k)).astype(np.float32)
    np.testing.assert_allclose(F.interpolation_forward(cu(inp), cu(ik), cu(w)).cpu().numpy(),
                               oracle.interpolation_forward(inp, ik, w), rtol=RTOL, atol=ATOL)
    go = rng.standard_normal((m, c)).astype(np.float32)
    got = F.interpolation_backward(cu(go), cu(w), F.ScatterPlan(cu(ik), 1, n), n, k).cpu().numpy()
    np.testing.assert_allclose(got, oracle.interpolation_backward(go, ik, w, n), rtol=RTOL, atol=ATOL)
    # subtraction / aggregation
    idn = rng.integers(0, n, (n, ns)).astype(np.int32)
    in2 = rng.standard_normal((n, c)).astype(np.float32)
    assert np.array_equal(F.subtraction_forward(cu(inp), cu(in2), cu(idn)).cpu().numpy(), oracle.subtraction_forward(inp, in2, idn))
    gs = rng.standard_normal((n, ns, c)).astype(np.float32)
    plan = F.ScatterPlan(cu(idn), 1, n)
    g1, g2 = F.subtraction_backward(cu(gs), plan)
    e1, e2 = oracle.subtraction_backward(idn, gs)
    np.testing.assert_allclose(g1.cpu().numpy(), e1, rtol=RTOL, atol=5e-6)
    np.testing.assert_allclose(g2.cpu().numpy(), e2, rtol=RTOL, atol=5e-6)
    if c % 4 == 0:
        wc = c // 4
        pos = rng.standard_normal((n, ns, c)).astype(np.float32)
        wt = rng.standard_normal((n, ns, wc)).astype(np.float32)
        _check_aggregation(oracle, inp, pos, wt, idn, rng.standard_normal((n, c)).astype(np.float32), plan)


@pytest.mark.parametrize("c,wc", [(64, 8), (32, 4), (48, 12), (6, 3), (128, 16), (256, 32)])
def test_aggregation_vector_and_scalar_paths(cuda, oracle, c, wc):
    """aggregation fwd/bwd (openpoints/cpp/pointops/src/aggregation) for channel counts that take the float4 kernels
    (c, w_c multiples of 4; the fused grad_position + grad_weight kernel needs c/4 and w_c/4 to be powers of two, c/4 <= 32)
    and for those that fall back to the scalar ones."""
    rng = np.random.default_rng(c)
    n, ns = 2500, 16
    inp = rng.standard_normal((n, c)).astype(np.float32)
    idn = rng.integers(0, n, (n, ns)).astype(np.int32)
    pos = rng.standard_normal((n, ns, c)).astype(np.float32)
    wt = rng.standard_normal((n, ns, wc)).astype(np.float32)
    ga = rng.standard_normal((n, c)).astype(np.float32)
    plan = F.ScatterPlan(cu(idn), 1, n)
    gi, gp, gw = _check_aggregation(oracle, inp, pos, wt, idn, ga, plan)
    gi2, gp2, gw2 = F.aggregation_backward(cu(inp), cu(pos), cu(wt), cu(idn), cu(ga), plan)
    assert torch.equal(gi, gi2) and torch.equal(gw, gw2)


@pytest.mark.parametrize("B,N,E_shape", [(4, 16384, (1024, 32)), (2, 4096, (1024, 64)), (3, 1000, (250, 16)), (2, 8192, (16000, 3)),
                                         (1, 37, (5, 8)), (2, 40000, (4096, 32)), (1, 2048, (16384, 16)),
                                         # few targets, long lists: the chunked counting-sort builder (E >= 24 N, N <= 2048, E <= 65535)
                                         (3, 1024, (16384, 3)), (2, 256, (4096, 3)), (1, 2047, (16383, 3)), (2, 100, (1000, 7)), (1, 1, (65535, 1)), (2, 2048, (21845, 3))])
def test_scatter_plan_fused_builder_equals_multi_kernel_builder(cuda, B, N, E_shape):
    """The one-CTA-per-cloud fused builder (shared-memory counters, in-place segment sort) against the multi-kernel
    builder (mode 2 of the thread-local override forces it): identical row_ptr and perm - ascending sources per target -
    including long segments (first-hit padding, all-zero rows) and out-of-range indices, which both skip."""
    rng = np.random.default_rng(N + B)
    M, ns = E_shape
    idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
    idx[:, : max(M // 8, 1)] = idx[:, : max(M // 8, 1), :1]      # padded rows -> a few long segments
    idx[:, -max(M // 16, 1):] = 0                                  # empty balls -> one very long segment at target 0
    idx[0, 0, 0] = -1
    idx[-1, -1, -1] = N + 5
    t = cu(idx)
    fused = F.ScatterPlan(t, B, N)
    _lib.lib().geot_group_points_config(2)
    try:
        multi = F.ScatterPlan(t, B, N)
    finally:
        _lib.lib().geot_group_points_config(0)
    assert torch.equal(fused.row_ptr(), multi.row_ptr())
    total = fused.row_ptr()[:, -1]
    for b in range(B):
        n_b = int(total[b])
        assert torch.equal(fused.perm()[b, :n_b], multi.perm()[b, :n_b])
    rp, pm = fused.row_ptr().cpu().numpy(), fused.perm().cpu().numpy()
    flat = idx.reshape(B, -1)
    for b in range(B):
        for k in (0, int(flat[b, 5]) if 0 <= flat[b, 5] < N else 1, N - 1):
            seg = pm[b, rp[b, k]:rp[b, k + 1]]
            assert np.array_equal(seg, np.nonzero(flat[b] == k)[0])
