#!/usr/bin/env python
"""Generates tests/golden/validation_golden.npz by running the REFERENCE'S OWN `get_pred_whole` and `get_seg_metrics`
(examples/segmentation/train.py:781-832; the two function sources are cut from the file and exec'd as they are) on CPU,
in the authoring container (needs /root/reference; never run in tests or on the GPU box).  `.cuda()` is patched to the
identity; `pt_utils` (the un-vendored pointnet2_ops, train.py:39) is bound to the CPU restatement of the in-tree twin's
kernels (oracle/cpu_ref.c, itself pinned by the reference's kernels run on a B200).

    python tests/golden/make_validation_golden.py
"""
import ast
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = os.environ.get("GEOT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import cpu_ref  # noqa: E402

torch.Tensor.cuda = lambda self, *a, **k: self


def function_source(path, name):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            return ast.unparse(node)
    raise KeyError(name)


class _PtUtils:  # pointnet2_ops.pointnet2_utils surface used by train.py:792-797
    @staticmethod
    def three_nn(unknown, known):
        dist, idx, _ = cpu_ref.three_nn(unknown.numpy(), known.numpy())
        return torch.from_numpy(dist), torch.from_numpy(idx)

    @staticmethod
    def three_interpolate(features, idx, weight):
        return torch.from_numpy(cpu_ref.three_interpolate(features.numpy(), idx.numpy(), weight.contiguous().numpy()))


ns = {"torch": torch, "F": F, "np": np, "pt_utils": _PtUtils}
src = os.path.join(REF, "examples/segmentation/train.py")
for fn in ("get_pred_whole", "get_seg_metrics"):
    exec(compile(function_source(src, fn), "<reference train.py>", "exec"), ns)

out = {}
g = torch.Generator().manual_seed(1609)
for tag, (B, C, N, n_whole) in {"C17": (2, 17, 600, (1500, 2100)), "C33": (3, 33, 400, (900, 1300, 700))}.items():
    points = torch.rand(B, N, 3, generator=g) * 2 - 1
    center = [torch.rand(3, generator=g) * 10 for _ in range(B)]
    scale = [torch.rand(1, generator=g) * 20 + 5 for _ in range(B)]
    whole = [(torch.rand(n, 3, generator=g) * 2 - 1) * scale[b] + center[b] for b, n in enumerate(n_whole)]
    whole[0][:N] = points[0] * scale[0] + center[0]  # vertices that coincide with a sampled point: dist 0, weight 1/1e-8
    labels_pts = torch.randint(0, C, (B, N), generator=g)
    logits = torch.randn(B, C, N, generator=g) + 4.0 * F.one_hot(labels_pts, C).permute(0, 2, 1).float()
    preds = ns["get_pred_whole"](logits, points, whole, center, scale)
    # ground truth = the prediction with 20 % of the vertices relabelled; cloud 1 never holds class 3
    labels = []
    for b, p in enumerate(preds):
        l = p.squeeze(0).clone()
        flip = torch.rand(l.shape, generator=g) < 0.2
        l[flip] = torch.randint(0, C, (int(flip.sum()),), generator=g)
        if b == 1:
            l[l == 3] = 0
        labels.append(l)
    acc, miou, mdsc = ns["get_seg_metrics"](preds, labels)
    out[f"{tag}_logits"] = logits.numpy()
    out[f"{tag}_points"] = points.numpy()
    for b in range(B):
        out[f"{tag}_whole{b}"] = whole[b].numpy()
        out[f"{tag}_center{b}"] = center[b].numpy()
        out[f"{tag}_scale{b}"] = scale[b].numpy()
        out[f"{tag}_pred{b}"] = preds[b].numpy()
        out[f"{tag}_label{b}"] = labels[b].numpy()
    out[f"{tag}_acc"] = np.array([float(a) for a in acc], np.float32)
    out[f"{tag}_miou"] = np.array(miou, np.float32)
    out[f"{tag}_mdsc"] = np.array(mdsc, np.float32)
np.savez_compressed(os.path.join(HERE, "validation_golden.npz"), **out)
print("wrote", os.path.join(HERE, "validation_golden.npz"), len(out), "arrays",
      {k: out[k] for k in out if k.endswith(("acc", "miou", "mdsc"))})
