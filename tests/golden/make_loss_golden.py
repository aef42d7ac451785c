#!/usr/bin/env python
"""Generates tests/golden/loss_golden.npz by running the REFERENCE'S OWN `Poly1FocalLoss_U_corr`
(openpoints/loss/build.py:799-892; the class source is cut from the file and exec'd as is, only the registry
decorator is dropped) and tests/golden/loss_t_golden.npz from `Poly1FocalLoss_U_T` (build.py:564-688; `.cuda()` patched
to the identity) on CPU, in the authoring container (needs /root/reference; never run in tests or on the GPU box).

    python tests/golden/make_loss_golden.py
"""
import ast
import os
from copy import deepcopy

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

REF = os.environ.get("GEOT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def class_source(path, name):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == name:
            node.decorator_list = []  # @LOSS.register_module(): registry plumbing
            return ast.unparse(node)
    raise KeyError(name)


ns = {"torch": torch, "nn": nn, "F": F, "deepcopy": deepcopy}
exec(compile(class_source(os.path.join(REF, "openpoints/loss/build.py"), "Poly1FocalLoss_U_corr"), "<reference build.py>", "exec"), ns)
Loss = ns["Poly1FocalLoss_U_corr"]

out = {}
g = torch.Generator().manual_seed(2024)
for tag, (B, C, N) in {"C17": (2, 17, 257), "C33": (3, 33, 130), "C5": (1, 5, 64)}.items():
    logits = (torch.randn(B, C, N, generator=g) * 3).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), generator=g)
    conf = torch.rand(B, N, generator=g)
    conf[0, :5] = 0.95  # exactly on the threshold: ge() keeps them
    for variant, kw in {"thr": dict(thresh=0.95), "low": dict(thresh=0.3), "mask": dict(thresh=0.95, mask=(torch.rand(B, N, generator=g) > 0.5))}.items():
        crit = Loss()  # epsilon=1.0, alpha=0.25, gamma=2.0 (defaults used by cfgs/tooth_semi/*_ntm.yaml)
        logits.grad = None
        loss = crit(logits, labels, conf, **kw)
        loss.backward()
        out[f"{tag}_{variant}_loss"] = loss.detach().numpy()
        out[f"{tag}_{variant}_grad"] = logits.grad.detach().numpy().copy()
        if "mask" in kw:
            out[f"{tag}_{variant}_mask"] = kw["mask"].numpy()
        out[f"{tag}_{variant}_thresh"] = np.float32(kw["thresh"])
    # non-default hyper-parameters
    crit = Loss(epsilon=0.5, alpha=-1.0, gamma=1.5)
    logits.grad = None
    loss = crit(logits, labels, conf, thresh=0.5)
    loss.backward()
    out[f"{tag}_hp_loss"] = loss.detach().numpy()
    out[f"{tag}_hp_grad"] = logits.grad.detach().numpy().copy()
    out[f"{tag}_logits"] = logits.detach().numpy()
    out[f"{tag}_labels"] = labels.numpy()
    out[f"{tag}_conf"] = conf.numpy()
np.savez_compressed(os.path.join(HERE, "loss_golden.npz"), **out)
print("wrote", os.path.join(HERE, "loss_golden.npz"), len(out), "arrays")

# ---------------------------------------------------------------- Poly1FocalLoss_U_T (build.py:564-688)
torch.Tensor.cuda = lambda self, *a, **k: self
exec(compile(class_source(os.path.join(REF, "openpoints/loss/build.py"), "Poly1FocalLoss_U_T"), "<reference build.py>", "exec"), ns)
LossT = ns["Poly1FocalLoss_U_T"]
out = {}
g = torch.Generator().manual_seed(2025)
for tag, (B, C, N) in {"C17": (2, 17, 257), "C33": (3, 33, 130), "C5": (1, 5, 64)}.items():
    logits = (torch.randn(B, C, N, generator=g) * 3).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), generator=g)
    conf = torch.rand(B, N, generator=g)
    conf[0, :5] = 0.95
    pred_t = torch.softmax(torch.randn(B, C, N, generator=g) * 2, dim=1).requires_grad_(True)  # corrected prediction
    T = torch.eye(C)
    for variant, hp, kw in [("thr", {}, dict(thresh=0.95)), ("low", {}, dict(thresh=0.3)),
                            ("mask", {}, dict(thresh=0.95, mask=(torch.rand(B, N, generator=g) > 0.5))),
                            ("hp", dict(epsilon=0.5, alpha=-1.0, gamma=1.5), dict(thresh=0.5))]:
        crit = LossT(**hp)
        logits.grad = None; pred_t.grad = None
        loss = crit(logits, labels, conf, T, pred_t, **kw)
        loss.backward()
        out[f"{tag}_{variant}_loss"] = loss.detach().numpy()
        out[f"{tag}_{variant}_grad"] = logits.grad.detach().numpy().copy()
        out[f"{tag}_{variant}_grad_t"] = pred_t.grad.detach().numpy().copy()
        if "mask" in kw:
            out[f"{tag}_{variant}_mask"] = kw["mask"].numpy()
        out[f"{tag}_{variant}_thresh"] = np.float32(kw["thresh"])
    out[f"{tag}_logits"] = logits.detach().numpy()
    out[f"{tag}_labels"] = labels.numpy()
    out[f"{tag}_conf"] = conf.numpy()
    out[f"{tag}_pred_t"] = pred_t.detach().numpy()
np.savez_compressed(os.path.join(HERE, "loss_t_golden.npz"), **out)
print("wrote", os.path.join(HERE, "loss_t_golden.npz"), len(out), "arrays")

# ---------------------------------------------------------------- Poly1FocalLoss_U_T_v1 (build.py:690-795; class-level T^T.p)
exec(compile(class_source(os.path.join(REF, "openpoints/loss/build.py"), "Poly1FocalLoss_U_T_v1"), "<reference build.py>", "exec"), ns)
LossV1 = ns["Poly1FocalLoss_U_T_v1"]
out = {}
g = torch.Generator().manual_seed(2026)
for tag, (B, C, N) in {"C17": (2, 17, 257), "C33": (3, 33, 130), "C5": (1, 5, 64)}.items():
    logits = (torch.randn(B, C, N, generator=g) * 3).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), generator=g)
    conf = torch.rand(B, N, generator=g)
    conf[0, :5] = 0.95
    pred_u = torch.softmax(torch.randn(B, C, N, generator=g) * 2, dim=1).requires_grad_(True)
    T = torch.nn.functional.normalize(torch.rand(C, C, generator=g) + torch.eye(C) * 4, p=1, dim=1)
    delta_T = (torch.randn(C, C, generator=g) * 0.01).requires_grad_(True)
    for variant, hp, kw in [("thr", {}, dict(thresh=0.95)), ("low", {}, dict(thresh=0.3)),
                            ("mask", {}, dict(thresh=0.95, mask=(torch.rand(B, N, generator=g) > 0.5))),
                            ("hp", dict(epsilon=0.5, alpha=-1.0, gamma=1.5), dict(thresh=0.5))]:
        crit = LossV1(**hp)
        logits.grad = None; pred_u.grad = None; delta_T.grad = None
        loss, dT = crit(logits, labels, conf, T, pred_u, delta_T, **kw)
        assert dT is delta_T
        loss.backward()
        out[f"{tag}_{variant}_loss"] = loss.detach().numpy()
        out[f"{tag}_{variant}_grad"] = logits.grad.detach().numpy().copy()
        out[f"{tag}_{variant}_grad_u"] = pred_u.grad.detach().numpy().copy()
        out[f"{tag}_{variant}_grad_dT"] = delta_T.grad.detach().numpy().copy()
        if "mask" in kw:
            out[f"{tag}_{variant}_mask"] = kw["mask"].numpy()
        out[f"{tag}_{variant}_thresh"] = np.float32(kw["thresh"])
    out[f"{tag}_logits"] = logits.detach().numpy()
    out[f"{tag}_labels"] = labels.numpy()
    out[f"{tag}_conf"] = conf.numpy()
    out[f"{tag}_pred_u"] = pred_u.detach().numpy()
    out[f"{tag}_T"] = T.numpy()
    out[f"{tag}_delta_T"] = delta_T.detach().numpy()
np.savez_compressed(os.path.join(HERE, "loss_tv1_golden.npz"), **out)
print("wrote", os.path.join(HERE, "loss_tv1_golden.npz"), len(out), "arrays")
