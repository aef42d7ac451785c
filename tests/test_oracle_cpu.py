"""The CPU oracle against independent brute-force statements of the op semantics (numpy, float64 where
order-free) and against the FPS tie rule derived from the reference's shared-memory tree.  The golden
fixtures produced by the reference's own kernels on a B200 are checked in test_golden.py."""
import numpy as np
import pytest

from geot_b200 import synth


def _d2_ref(a, b):
    """fma(dz,dz,fma(dx,dx,dy*dy)) emulated in float64->float32 steps (exact: fp32 products fit in fp64)."""
    d = (a.astype(np.float32) - b.astype(np.float32)).astype(np.float32)
    dx, dy, dz = (d[..., i].astype(np.float64) for i in range(3))
    t = (dy * dy).astype(np.float32).astype(np.float64)
    t = (dx * dx + t).astype(np.float32).astype(np.float64)
    return (dz * dz + t).astype(np.float32)


def test_opt_n_threads(oracle):
    assert oracle.opt_n_threads(16384) == 1024 and oracle.opt_n_threads(16000) == 1024
    assert oracle.opt_n_threads(1000) == 512 and oracle.opt_n_threads(16000, 512) == 512
    assert oracle.opt_n_threads(7) == 4 and oracle.opt_n_threads(1) == 1
    for p in range(0, 15):
        assert oracle.opt_n_threads(1 << p, 1 << 20) == 1 << p  # exact at powers of two (glibc log)


def _fps_rule(xyz, M, BS, skip=None):
    """max distance, then min (bitrev(k mod BS), k div BS) — the closed form of the reference's tree."""
    N = xyz.shape[0]
    lb = BS.bit_length() - 1
    k = np.arange(N)
    br = np.array([int(format(t, f"0{lb}b")[::-1], 2) if lb else 0 for t in range(BS)])
    rank = br[k % BS] * ((N + BS - 1) // BS) + k // BS
    mind = np.full(N, 1e10, np.float32)
    live = np.ones(N, bool) if skip is None else ~skip
    out = [0]
    for _ in range(1, M):
        d = _d2_ref(xyz, xyz[out[-1]][None])
        mind = np.where(live, np.minimum(d, mind), mind)
        cand = np.where(live, mind, -1.0)
        best = cand.max()
        tie = np.flatnonzero(cand == best)
        out.append(int(tie[np.argmin(rank[tie])]))
    return np.array(out, np.int32)


@pytest.mark.parametrize("kind,N,M", [("plain", 2000, 64), ("quant", 3000, 200), ("dup", 1500, 300), ("quant", 700, 700)])
def test_fps_tie_rule_closed_form(oracle, kind, N, M):
    xyz = synth.tooth_cloud(N, 7) if kind == "plain" else synth.adversarial_cloud(N, 11, kind)
    got = oracle.furthest_point_sample(xyz[None], M)[0]
    assert np.array_equal(got, _fps_rule(xyz, M, oracle.opt_n_threads(N)))


def test_fps_magskip_variant(oracle):
    xyz = synth.adversarial_cloud(2500, 5, "origin")
    skip = (_d2_ref(xyz, np.zeros((1, 3), np.float32)).astype(np.float64) <= 1e-3)
    assert skip.sum() > 3
    M = 1800  # deep enough that the plain variant samples the near-origin points the _ext variant skips
    got = oracle.furthest_point_sample(xyz[None], M, cap=512, mag_skip=True)[0]
    assert np.array_equal(got, _fps_rule(xyz, M, 512, skip))
    assert not np.array_equal(got, oracle.furthest_point_sample(xyz[None], M)[0])
    assert not skip[got[1:]].any()


def test_fps_offset_matches_batched(oracle):
    xyz = synth.tooth_batch(3, 1024)
    a = oracle.furthest_point_sample(xyz, 100)
    off = np.array([1024, 2048, 3072], np.int32)
    b = oracle.furthestsampling(xyz.reshape(-1, 3), off, np.array([100, 200, 300], np.int32))
    assert np.array_equal(a + (off - 1024)[:, None], b.reshape(3, 100))


def test_ball_query_bruteforce(oracle):
    xyz = synth.tooth_batch(2, 3000)
    q = xyz[:, ::50].copy()
    r, ns = 0.12, 16
    got = oracle.ball_query(r, ns, xyz, q)
    r2 = np.float32(r) * np.float32(r)
    for b in range(2):
        for j in range(q.shape[1]):
            hits = np.flatnonzero(_d2_ref(xyz[b], q[b, j][None]) < r2)[:ns]
            exp = np.zeros(ns, np.int32)
            if len(hits):
                exp[:] = hits[0]
                exp[:len(hits)] = hits
            assert np.array_equal(got[b, j], exp)
    far = oracle.ball_query(0.01, 8, xyz, q + 5.0)
    assert not far.any()


def test_three_nn_and_knn_bruteforce(oracle):
    xyz = synth.tooth_batch(2, 2048)
    known = xyz[:, :300].copy()
    dist, idx, d2 = oracle.three_nn(xyz, known)
    for b in range(2):
        D = _d2_ref(xyz[b][:, None, :], known[b][None, :, :])
        order = np.argsort(D, axis=1, kind="stable")[:, :3]
        assert np.array_equal(idx[b], order.astype(np.int32))
        assert np.array_equal(d2[b], np.take_along_axis(D, order, 1))
    off = np.array([2048, 4096], np.int32)
    noff = np.array([300, 600], np.int32)
    ki, kd, kd2 = oracle.knnquery(8, xyz.reshape(-1, 3), known.reshape(-1, 3), off, noff)
    for b in range(2):
        D = _d2_ref(known[b][:, None, :], xyz[b][None, :, :])
        order = np.argsort(D, axis=1, kind="stable")[:, :8]
        if all(len(np.unique(D[i, order[i]])) == 8 for i in range(300)):
            assert np.array_equal(ki[b * 300:(b + 1) * 300], (order + b * 2048).astype(np.int32))
    # short segment: unfilled slots are (1e10, segment start)
    ki, _, kd2 = oracle.knnquery(5, xyz[0, :3], xyz[0, :2], np.array([3], np.int32), np.array([2], np.int32))
    assert (kd2[:, 3:] == np.float32(1e10)).all() and (ki[:, 3:] == 0).all()
    d, i3, d2 = oracle.three_nn(xyz[:, :4], xyz[:, :2])
    assert np.isinf(d2[..., 2]).all() and (i3[..., 2] == 0).all()


def test_group_interp_forward_backward_bruteforce(oracle):
    rng = np.random.default_rng(3)
    B, C, N, M, ns = 2, 5, 64, 10, 7
    feat = rng.standard_normal((B, C, N)).astype(np.float32)
    idx = rng.integers(0, N, (B, M, ns)).astype(np.int32)
    out = oracle.grouping_operation(feat, idx)
    exp = np.stack([feat[b][:, idx[b]] for b in range(B)])
    assert np.array_equal(out, exp)
    g = rng.standard_normal(out.shape).astype(np.float32)
    gi = oracle.grouping_operation_grad(g, idx, N)
    e = np.zeros((B, C, N))
    for b in range(B):
        for c in range(C):
            np.add.at(e[b, c], idx[b].ravel(), g[b, c].ravel().astype(np.float64))
    np.testing.assert_allclose(gi, e, rtol=1e-6, atol=1e-6)
    i3 = rng.integers(0, N, (B, 20, 3)).astype(np.int32)
    w = rng.random((B, 20, 3)).astype(np.float32)
    o = oracle.three_interpolate(feat, i3, w)
    e = np.stack([(feat[b][:, i3[b]] * w[b][None]).sum(-1) for b in range(B)])
    np.testing.assert_allclose(o, e, rtol=1e-5, atol=1e-6)
    go = rng.standard_normal(o.shape).astype(np.float32)
    gp = oracle.three_interpolate_grad(go, i3, w, N)
    e = np.zeros((B, C, N))
    for b in range(B):
        for c in range(C):
            np.add.at(e[b, c], i3[b].ravel(), (go[b, c][:, None] * w[b]).ravel().astype(np.float64))
    np.testing.assert_allclose(gp, e, rtol=1e-5, atol=1e-6)
    assert np.array_equal(oracle.gather_operation(feat, idx[:, :, 0]), np.stack([feat[b][:, idx[b, :, 0]] for b in range(B)]))


def test_offset_ops_bruteforce(oracle):
    rng = np.random.default_rng(4)
    n, m, ns, c, wc = 50, 20, 6, 8, 4
    inp = rng.standard_normal((n, c)).astype(np.float32)
    idx = rng.integers(0, n, (m, ns)).astype(np.int32)
    assert np.array_equal(oracle.grouping(inp, idx), inp[idx])
    g = rng.standard_normal((m, ns, c)).astype(np.float32)
    e = np.zeros((n, c)); np.add.at(e, idx.ravel(), g.reshape(-1, c).astype(np.float64))
    np.testing.assert_allclose(oracle.grouping_grad(g, idx, n), e, rtol=1e-6, atol=1e-6)
    w = rng.random((m, ns)).astype(np.float32)
    np.testing.assert_allclose(oracle.interpolation_forward(inp, idx, w), (inp[idx] * w[..., None]).sum(1), rtol=1e-5, atol=1e-6)
    idn = rng.integers(0, n, (n, ns)).astype(np.int32)
    in2 = rng.standard_normal((n, c)).astype(np.float32)
    assert np.array_equal(oracle.subtraction_forward(inp, in2, idn), inp[:, None, :] - in2[idn])
    pos = rng.standard_normal((n, ns, c)).astype(np.float32)
    wt = rng.standard_normal((n, ns, wc)).astype(np.float32)
    exp = ((inp[idn] + pos) * np.tile(wt, (1, 1, c // wc))).sum(1)
    np.testing.assert_allclose(oracle.aggregation_forward(inp, pos, wt, idn), exp, rtol=1e-5, atol=1e-5)
