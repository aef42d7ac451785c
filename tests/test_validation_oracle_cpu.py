"""oracle/validation_ref.py against the fixture produced by the REFERENCE'S OWN get_pred_whole / get_seg_metrics
(tests/golden/make_validation_golden.py ran the two function sources of examples/segmentation/train.py on CPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import validation_ref as V

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "validation_golden.npz"))
CASES = {"C17": 2, "C33": 3}


def t(name):
    return torch.from_numpy(G[name])


@pytest.mark.parametrize("tag", list(CASES))
def test_get_pred_whole(tag):
    B = CASES[tag]
    preds, _ = V.get_pred_whole(t(f"{tag}_logits"), t(f"{tag}_points"), [t(f"{tag}_whole{b}") for b in range(B)],
                                [t(f"{tag}_center{b}") for b in range(B)], [t(f"{tag}_scale{b}") for b in range(B)])
    for b in range(B):
        assert torch.equal(preds[b], t(f"{tag}_pred{b}"))  # same arithmetic, same machine: bit-exact labels


@pytest.mark.parametrize("tag", list(CASES))
def test_get_seg_metrics(tag):
    B = CASES[tag]
    acc, miou, mdsc = V.get_seg_metrics([t(f"{tag}_pred{b}") for b in range(B)], [t(f"{tag}_label{b}") for b in range(B)])
    assert np.array_equal(np.array([float(a) for a in acc], np.float32), G[f"{tag}_acc"])
    assert np.array_equal(np.array(miou, np.float32), G[f"{tag}_miou"])
    assert np.array_equal(np.array(mdsc, np.float32), G[f"{tag}_mdsc"])


def test_metrics_edge_cases():
    # only background in the ground truth: no class enters the means (the reference's np.array([]).mean() is nan)
    acc, miou, mdsc = V.seg_metrics_one(torch.tensor([0, 1, 0]), torch.tensor([0, 0, 0]))
    assert float(acc) == pytest.approx(2 / 3) and np.isnan(miou) and np.isnan(mdsc)
    # a class the prediction never produces counts with IoU 0
    acc, miou, mdsc = V.seg_metrics_one(torch.tensor([1, 1, 0, 0]), torch.tensor([1, 1, 2, 2]))
    assert float(miou) == pytest.approx(0.5) and float(mdsc) == pytest.approx(0.5)


def _counts(pred, label, C):
    """what geot_seg_counts returns, computed with numpy (host logic test: no GPU here)."""
    p, l = pred.numpy().reshape(-1), label.numpy().reshape(-1)
    inter = np.bincount(p[p == l], minlength=C)
    return torch.from_numpy(np.concatenate([inter, np.bincount(p, minlength=C), np.bincount(l, minlength=C),
                                            [int((p == l).sum()), 0]]).astype(np.int64))


@pytest.mark.parametrize("tag", list(CASES))
def test_product_host_metrics_from_counts(tag):
    """geot_b200.validation.metrics_from_counts (the host half of the product's get_seg_metrics) on the fixture."""
    from geot_b200 import validation as VAL
    C = G[f"{tag}_logits"].shape[1]
    for b in range(CASES[tag]):
        pred, label = t(f"{tag}_pred{b}"), t(f"{tag}_label{b}")
        acc, miou, mdsc = VAL.metrics_from_counts(_counts(pred, label, C), C, label.numel())
        assert np.float32(float(acc)) == G[f"{tag}_acc"][b]
        assert np.float32(miou) == G[f"{tag}_miou"][b] and np.float32(mdsc) == G[f"{tag}_mdsc"][b]
    bad = _counts(t(f"{tag}_pred0"), t(f"{tag}_label0"), C)
    bad[3 * C + 1] = 4
    with pytest.raises(ValueError):
        VAL.metrics_from_counts(bad, C, 10)
