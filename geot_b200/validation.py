"""geot_b200.validation — the inference side of the hot path: 16000-point predictions -> full-mesh labels -> metrics.

Mirrors examples/segmentation/train.py:781-832 (`get_pred_whole`, `get_seg_metrics`; same names, arguments and
return values) over geot_b200/csrc/upsample.cu.  The reference runs, per cloud, three_nn + four torch kernels for
the weights + three_interpolate into a (1,C,n_whole) tensor + argmax, then moves everything to the CPU and loops over
the classes in Python.  Here a cloud is three launches (three_nn, weights, blend + arg-max; the upsampled score tensor
never exists) and its metrics come from one integer histogram kernel and one small device->host copy.  CUDA only."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import functional as F_


@torch.no_grad()
def upsample_labels(prob, points, points_whole):
    """prob (B,C,m) per-class scores at the sampled points, points (B,m,3), points_whole (B,n,3) -> labels (B,n) int64:
    arg-max over classes of the inverse-distance blend of the 3 nearest sampled points (train.py:793-798)."""
    dist2, idx = F_.three_nn(points_whole.contiguous().float(), points.contiguous().float())
    weight = F_.interp_weights(dist2)
    return F_.interp_argmax(prob.contiguous().float(), idx, weight)


@torch.no_grad()
def get_pred_whole(logits, points, points_whole, center, scale):
    """train.py:781-800.  logits (B,C,N), points (B,N,3) normalised sample, points_whole: per-cloud (n_b,3) mesh vertices,
    center / scale: per-cloud de-normalisation -> list of (1, n_b) int64 predictions."""
    dev = logits.device
    logits = F.softmax(logits, dim=1)  # B C N
    preds_whole = []
    for index in range(logits.shape[0]):
        logit = logits[index].unsqueeze(0).contiguous()
        point = points[index].unsqueeze(0).contiguous()
        s = scale[index].to(dev).unsqueeze(0).contiguous()
        c = center[index].to(dev).unsqueeze(0).contiguous()
        point_whole = points_whole[index].to(dev).unsqueeze(0).contiguous()
        point = point * s + c
        preds_whole.append(upsample_labels(logit, point, point_whole))
    return preds_whole


def metrics_from_counts(cnt, num_classes, n):
    """Host half of get_seg_metrics: cnt = the (3C+2) int64 CPU tensor geot_seg_counts fills, n = number of vertices ->
    (acc, miou, mdsc) exactly as train.py:808-826 computes them (classes = labels present in the ground truth except 0;
    IoU and DSC in fp32; numpy fp32 means; nan when no class qualifies)."""
    C = num_classes
    if int(cnt[3 * C + 1]) != 0:
        raise ValueError(f"get_seg_metrics: {int(cnt[3 * C + 1])} vertices carry a class id outside [0, {C})")
    inter, pcnt, lcnt, correct = cnt[:C], cnt[C:2 * C], cnt[2 * C:3 * C], cnt[3 * C]
    iou, dsc = [], []
    for jcls in range(1, C):
        if int(lcnt[jcls]) == 0:  # torch.unique(label_whole): only the classes the ground truth holds
            continue
        jcls_and = inter[jcls]
        jcls_or = pcnt[jcls] + lcnt[jcls] - inter[jcls]
        iou.append((jcls_and / jcls_or).float())
        dsc.append((2 * iou[-1] / (1 + iou[-1])))
    acc = correct / n
    with np.errstate(all="ignore"):
        miou = np.array(iou, dtype=np.float32).mean() if iou else np.float32("nan")
        mdsc = np.array(dsc, dtype=np.float32).mean() if dsc else np.float32("nan")
    return acc, miou, mdsc


@torch.no_grad()
def seg_metrics_one(pred, label, num_classes=None):
    """One cloud: pred, label (n) integer tensors on the GPU -> (acc, miou, mdsc); one histogram kernel, one copy of
    3C+2 integers to the host, then metrics_from_counts."""
    pred = pred.reshape(-1).long().contiguous()
    label = label.reshape(-1).long().contiguous()
    if num_classes is None:
        num_classes = int(torch.maximum(pred.max(), label.max()).item()) + 1 if pred.numel() else 1
    cnt = F_.seg_counts(pred, label, num_classes).cpu()
    return metrics_from_counts(cnt, num_classes, label.numel())


@torch.no_grad()
def get_seg_metrics(preds_whole, labels_whole, num_classes=None):
    """train.py:803-832: lists of per-cloud predictions / labels -> (acc_list, miou_list, mdsc_list)."""
    acc_list, miou_list, mdsc_list = [], [], []
    for index in range(len(preds_whole)):
        pred_whole = preds_whole[index].detach()
        label_whole = labels_whole[index].to(pred_whole.device)
        acc, miou, mdsc = seg_metrics_one(pred_whole, label_whole, num_classes)
        acc_list.append(acc)
        miou_list.append(miou)
        mdsc_list.append(mdsc)
    return acc_list, miou_list, mdsc_list
