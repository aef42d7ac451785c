"""Validation upsampling + metrics on the GPU (through the C ABI: geot_interp_weights, geot_interp_argmax,
geot_seg_counts) against (1) the golden fixture produced by the reference's own get_pred_whole / get_seg_metrics and
(2) oracle/validation_ref.py on seeded inputs, plus size-independent properties at full mesh sizes.
Labels and counts are integer work: bit-exact, except that an arg-max over fp32 scores may differ where the two best
scores are closer than the 1e-5 relative tolerance of the north_star (checked explicitly); weights within 1e-6."""
import os

import numpy as np
import pytest
import torch

from geot_b200 import functional as F_
from geot_b200 import layers, synth, validation as VAL
from oracle import validation_ref as V

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "validation_golden.npz"))
CASES = {"C17": 2, "C33": 3}


def t(name, dev="cuda"):
    return torch.from_numpy(G[name]).to(dev)


def labels_equal_up_to_near_ties(pred, ref_pred, ref_scores, tol=1e-5):
    """pred, ref_pred (n) int64; ref_scores (C,n): every disagreement must be a near tie of the oracle's scores."""
    pred, ref_pred = pred.cpu().reshape(-1), ref_pred.cpu().reshape(-1)
    bad = (pred != ref_pred).nonzero().reshape(-1)
    if bad.numel() == 0:
        return 0
    s = ref_scores[:, bad]
    gap = (s.max(dim=0).values - s.gather(0, pred[bad].unsqueeze(0)).squeeze(0)).abs()
    assert float(gap.max()) <= tol, f"{bad.numel()} labels differ, largest score gap {float(gap.max())}"
    return int(bad.numel())


@pytest.mark.parametrize("tag", list(CASES))
def test_golden_get_pred_whole_and_metrics(cuda, tag):
    B = CASES[tag]
    C = G[f"{tag}_logits"].shape[1]
    whole = [t(f"{tag}_whole{b}", "cpu") for b in range(B)]  # the reference keeps the meshes on the host
    center = [t(f"{tag}_center{b}", "cpu") for b in range(B)]
    scale = [t(f"{tag}_scale{b}", "cpu") for b in range(B)]
    preds = VAL.get_pred_whole(t(f"{tag}_logits"), t(f"{tag}_points"), whole, center, scale)
    _, ref_scores = V.get_pred_whole(t(f"{tag}_logits", "cpu"), t(f"{tag}_points", "cpu"), whole, center, scale)
    differing = 0
    for b in range(B):
        assert preds[b].shape == (1, whole[b].shape[0]) and preds[b].dtype == torch.int64
        differing += labels_equal_up_to_near_ties(preds[b], t(f"{tag}_pred{b}", "cpu"), ref_scores[b][0])
    assert differing <= 5  # softmax runs on the GPU here and on the CPU in the fixture: last-bit differences only
    # metrics on the FIXTURE's predictions: integer counts -> bit-exact fp32 results
    acc, miou, mdsc = VAL.get_seg_metrics([t(f"{tag}_pred{b}") for b in range(B)], [t(f"{tag}_label{b}", "cpu") for b in range(B)], C)
    assert np.array_equal(np.array([float(a) for a in acc], np.float32), G[f"{tag}_acc"])
    assert np.array_equal(np.array(miou, np.float32), G[f"{tag}_miou"])
    assert np.array_equal(np.array(mdsc, np.float32), G[f"{tag}_mdsc"])
    # and with the class count inferred from the data
    acc2, miou2, _ = VAL.get_seg_metrics([t(f"{tag}_pred{b}") for b in range(B)], [t(f"{tag}_label{b}") for b in range(B)])
    assert np.array_equal(np.array(miou2, np.float32), G[f"{tag}_miou"])


def test_interp_weights_vs_torch_expression(cuda):
    g = torch.Generator(device="cuda").manual_seed(3)
    d2 = torch.rand(4, 5001, 3, device="cuda", generator=g) * 0.01
    d2[0, :7] = 0.0          # coincident points: 1 / 1e-8
    d2[1, 3, 1:] = float("inf")  # three_nn with m < 3 leaves inf in the missing slots
    w, dist = F_.interp_weights(d2, want_dist=True)
    assert torch.equal(dist, torch.sqrt(d2))
    ref = V.interp_weights(torch.sqrt(d2).cpu())
    torch.testing.assert_close(w.cpu(), ref, rtol=1e-6, atol=1e-12)
    torch.testing.assert_close(w.sum(-1), torch.ones_like(w[..., 0]), rtol=1e-6, atol=0)
    assert w[1, 3].tolist() == [1.0, 0.0, 0.0]
    assert F_.interp_weights(d2[:, :0]).shape == (4, 0, 3)


def test_three_interpolation_composite_matches_the_unfused_chain(cuda):
    g = torch.Generator(device="cuda").manual_seed(5)
    known = torch.rand(3, 700, 3, device="cuda", generator=g)
    unknown = torch.rand(3, 2900, 3, device="cuda", generator=g)
    feat = torch.randn(3, 24, 700, device="cuda", generator=g).requires_grad_(True)
    out = layers.three_interpolation(unknown, known, feat)
    dist, idx = layers.three_nn(unknown, known)
    dist_recip = 1.0 / (dist + 1e-8)
    weight = dist_recip / torch.sum(dist_recip, dim=2, keepdim=True)
    ref = layers.three_interpolate(feat, idx, weight)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)  # weights may differ in the last bit (order of the 3-term sum)
    go = torch.randn_like(out)
    g1, = torch.autograd.grad(out, feat, go, retain_graph=True)
    g2, = torch.autograd.grad(ref, feat, go)
    torch.testing.assert_close(g1, g2, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("B,C,m,n", [(1, 17, 16000, 100003), (2, 33, 4096, 20000), (1, 1, 5, 9), (3, 6, 2, 64)])
def test_interp_argmax_equals_interpolate_then_argmax(cuda, B, C, m, n):
    """the fused kernel blends with three_interpolate's expression: identical scores, identical labels (bit-exact)."""
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + C)
    known = torch.rand(B, m, 3, device="cuda", generator=g)
    unknown = torch.rand(B, n, 3, device="cuda", generator=g)
    prob = torch.softmax(torch.randn(B, C, m, device="cuda", generator=g) * 3, dim=1)
    d2, idx = F_.three_nn(unknown, known)
    w = F_.interp_weights(d2)
    if m < 3:
        w = torch.nan_to_num(w, nan=0.0)
    pred, score = F_.interp_argmax(prob, idx, w, want_score=True)
    full = F_.three_interpolate(prob, idx, w)
    assert torch.equal(score, full.max(dim=1).values)
    # first maximum, computed without relying on torch.argmax's tie order
    classes = torch.arange(C, device="cuda").view(1, C, 1).expand_as(full)
    first = torch.where(full == full.max(dim=1, keepdim=True).values, classes, C).min(dim=1).values
    assert torch.equal(pred, first)
    assert torch.equal(VAL.upsample_labels(prob, known, unknown), pred) or m < 3


def test_interp_argmax_vs_oracle(cuda):
    g = torch.Generator().manual_seed(11)
    B, C, m, n = 2, 33, 3000, 7001
    known, unknown = torch.rand(B, m, 3, generator=g), torch.rand(B, n, 3, generator=g)
    unknown[0, :100] = known[0, :100]
    prob = torch.softmax(torch.randn(B, C, m, generator=g) * 3, dim=1)
    scores, idx, weight = V.upsample_scores(prob, known, unknown)
    pred = VAL.upsample_labels(prob.cuda(), known.cuda(), unknown.cuda())
    for b in range(B):
        labels_equal_up_to_near_ties(pred[b], scores[b].argmax(dim=0), scores[b])
    # same neighbours and weights in: bit-exact labels out
    pred2 = F_.interp_argmax(prob.cuda(), idx.cuda(), weight.contiguous().cuda())
    assert torch.equal(pred2.cpu(), scores.argmax(dim=1))


@pytest.mark.parametrize("n,C", [(0, 3), (1, 2), (777, 17), (250000, 33), (1 << 21, 33)])
def test_seg_counts_vs_bincount_and_oracle(cuda, n, C):
    g = torch.Generator(device="cuda").manual_seed(n + C)
    label = torch.randint(0, C, (n,), device="cuda", generator=g)
    pred = torch.where(torch.rand(n, device="cuda", generator=g) < 0.7, label, torch.randint(0, C, (n,), device="cuda", generator=g))
    cnt = F_.seg_counts(pred, label, C)
    assert torch.equal(cnt[:C], torch.bincount(pred[pred == label], minlength=C))
    assert torch.equal(cnt[C:2 * C], torch.bincount(pred, minlength=C))
    assert torch.equal(cnt[2 * C:3 * C], torch.bincount(label, minlength=C))
    assert int(cnt[3 * C]) == int((pred == label).sum()) and int(cnt[3 * C + 1]) == 0
    assert int(cnt[C:2 * C].sum()) == n  # checksum of checksums
    if 0 < n <= 250000:
        acc, miou, mdsc = VAL.seg_metrics_one(pred, label, C)
        racc, rmiou, rmdsc = V.seg_metrics_one(pred.cpu(), label.cpu())
        assert float(acc) == float(racc)
        assert np.array_equal(np.float32(miou), np.float32(rmiou), equal_nan=True)
        assert np.array_equal(np.float32(mdsc), np.float32(rmdsc), equal_nan=True)


def test_seg_metrics_edge_cases(cuda):
    acc, miou, mdsc = VAL.seg_metrics_one(torch.tensor([0, 1, 0], device="cuda"), torch.tensor([0, 0, 0], device="cuda"))
    assert float(acc) == pytest.approx(2 / 3) and np.isnan(miou) and np.isnan(mdsc)
    acc, miou, mdsc = VAL.seg_metrics_one(torch.tensor([1, 1, 0, 0], device="cuda"), torch.tensor([1, 1, 2, 2], device="cuda"))
    assert float(miou) == pytest.approx(0.5) and float(mdsc) == pytest.approx(0.5)
    with pytest.raises(ValueError):
        VAL.seg_metrics_one(torch.tensor([5, 1], device="cuda"), torch.tensor([1, 1], device="cuda"), num_classes=3)
    from geot_b200._lib import GeotError
    with pytest.raises(GeotError):
        VAL.seg_metrics_one(torch.tensor([1, 1]), torch.tensor([1, 1]), num_classes=3)  # CPU tensors: no fallback
