"""oracle/cpu_ref.c against golden OUTPUTS OF THE REFERENCE'S OWN CUDA KERNELS (tests/golden/pointops_golden.npz,
made on a B200 by tests/golden/make_pointops_golden.py from oracle/_ref).  Indices bit-exact; values 1e-5."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from make_pointops_golden import CASES, clouds, features  # noqa: E402  (input generators only; no GPU import at module level)

G = np.load(os.path.join(ROOT, "tests", "golden", "pointops_golden.npz"))


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_oracle_matches_reference_kernels(oracle, ci):
    kind, B, N, M = CASES[ci]
    t = f"c{ci}"
    xyz = clouds(kind, B, N)
    fb = oracle.furthest_point_sample(xyz, M)
    assert np.array_equal(fb, G[f"{t}_fps_batch"])
    assert np.array_equal(oracle.furthest_point_sample(xyz, M, cap=512, mag_skip=True), G[f"{t}_fps_ext"])
    centres = np.take_along_axis(xyz, fb[..., None].astype(np.int64).repeat(3, 2), 1)
    r, ns = (0.1, 32) if kind == "tooth" else (0.125, 24)
    bq = oracle.ball_query(r, ns, xyz, centres)
    assert np.array_equal(bq, G[f"{t}_ball"])
    _, i3, d2 = oracle.three_nn(xyz, centres)
    assert np.array_equal(i3, G[f"{t}_nn_idx"]) and np.array_equal(d2, G[f"{t}_nn_d2"])
    feat = features((B, 5, N), 10 + ci)
    assert np.array_equal(oracle.grouping_operation(feat, bq), G[f"{t}_group"])
    go = features((B, 5, M, ns), 20 + ci)
    np.testing.assert_allclose(oracle.grouping_operation_grad(go, bq, N), G[f"{t}_group_grad"], rtol=1e-5, atol=1e-5)
    assert np.array_equal(oracle.gather_operation(feat, fb), G[f"{t}_gather"])
    w = features((B, N, 3), 30 + ci)
    w = np.exp(w - w.max(-1, keepdims=True)); w = (w / w.sum(-1, keepdims=True)).astype(np.float32)
    kf = features((B, 5, M), 40 + ci)
    np.testing.assert_allclose(oracle.three_interpolate(kf, i3, w), G[f"{t}_interp"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(oracle.three_interpolate_grad(features((B, 5, N), 50 + ci), i3, w, M), G[f"{t}_interp_grad"], rtol=1e-5, atol=1e-5)
    flat = xyz.reshape(-1, 3)
    offset = np.array([N * (i + 1) for i in range(B)], np.int32)
    new_offset = np.array([M * (i + 1) for i in range(B)], np.int32)
    fo = oracle.furthestsampling(flat, offset, new_offset)
    assert np.array_equal(fo, G[f"{t}_fps_offset"])
    wts = (np.random.default_rng(60 + ci).random(B * N).astype(np.float32) + np.float32(0.05)).astype(np.float32)
    assert np.array_equal(oracle.furthestsampling(flat, offset, new_offset, wts), G[f"{t}_fps_offset_w"])
    q = np.ascontiguousarray(flat[fo.astype(np.int64)])
    for k in (3, 16):
        ki, _, kd = oracle.knnquery(k, flat, q, offset, new_offset)
        assert np.array_equal(ki, G[f"{t}_knn{k}_idx"]) and np.array_equal(kd, G[f"{t}_knn{k}_d2"])
    assert np.array_equal(oracle.ballquery(r, ns, flat, q, offset, new_offset), G[f"{t}_ball_offset"])
