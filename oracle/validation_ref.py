"""oracle/validation_ref.py — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

CPU restatement of the reference's validation upsampling and metrics (examples/segmentation/train.py:781-832):
the 3-nearest-neighbour search and blend go through oracle/cpu_ref.c (the restated kernels, pinned by the reference's
own kernels run on a B200), the weights and the metrics follow the reference's torch / numpy expressions.
Pinned against the reference's OWN function sources run on CPU by tests/golden/make_validation_golden.py
(fixture tests/golden/validation_golden.npz).  `pt_utils` at train.py:39 is the un-vendored `pointnet2_ops`; its
semantics are those of the in-tree twin pointnet2/pointnet2_utils.py:117-203, which cpu_ref.c restates.

Paths are relative to the reference root.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import cpu_ref


def interp_weights(dist):
    """train.py:794-796 = upsampling.py:98-100: dist (…,3) fp32 tensor of (non-squared) distances -> weights."""
    dist_recip = 1.0 / (dist + 1e-8)
    norm = torch.sum(dist_recip, dim=-1, keepdim=True)
    return dist_recip / norm


def upsample_scores(prob, points, points_whole):
    """train.py:793-797 for one batch: prob (B,C,m), points (B,m,3), points_whole (B,n,3) ->
    (scores (B,C,n), idx (B,n,3), weight (B,n,3)) as torch CPU tensors."""
    dist, idx, _ = cpu_ref.three_nn(points_whole.numpy(), points.numpy())
    weight = interp_weights(torch.from_numpy(dist))
    scores = cpu_ref.three_interpolate(prob.contiguous().numpy(), idx, weight.contiguous().numpy())
    return torch.from_numpy(scores), torch.from_numpy(idx), weight


def get_pred_whole(logits, points, points_whole, center, scale):
    """train.py:781-800 -> list of (1, n_b) int64 predictions (and the score tensors, for near-tie checks)."""
    prob = F.softmax(logits, dim=1)
    preds, scores = [], []
    for b in range(prob.shape[0]):
        point = points[b].unsqueeze(0) * scale[b].unsqueeze(0) + center[b].unsqueeze(0)
        s, _, _ = upsample_scores(prob[b].unsqueeze(0), point.contiguous(), points_whole[b].unsqueeze(0).contiguous())
        preds.append(s.argmax(dim=1))
        scores.append(s)
    return preds, scores


def seg_metrics_one(pred, label):
    """train.py:808-826 for one cloud: integer tensors (n) -> (acc 0-dim fp32 tensor, miou, mdsc numpy fp32)."""
    pred, label = pred.reshape(-1), label.reshape(-1)
    iou, dsc = [], []
    for jcls in torch.unique(label).numpy():
        if jcls == 0:
            continue
        both = torch.logical_and(pred == jcls, label == jcls).sum()
        either = torch.logical_or(pred == jcls, label == jcls).sum()
        iou.append((both / either).float())
        dsc.append(2 * iou[-1] / (1 + iou[-1]))
    acc = (pred == label).sum() / label.shape[0]
    with np.errstate(all="ignore"):
        miou = np.array(iou, dtype=np.float32).mean() if iou else np.float32("nan")
        mdsc = np.array(dsc, dtype=np.float32).mean() if dsc else np.float32("nan")
    return acc, miou, mdsc


def get_seg_metrics(preds_whole, labels_whole):
    """train.py:803-832."""
    out = [seg_metrics_one(p, l) for p, l in zip(preds_whole, labels_whole)]
    return [o[0] for o in out], [o[1] for o in out], [o[2] for o in out]
