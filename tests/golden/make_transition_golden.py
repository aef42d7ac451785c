#!/usr/bin/env python
"""Generates tests/golden/transition_*.npz by running the REFERENCE'S OWN Python for the transition-matrix
path on CPU, in the authoring container (needs /root/reference; never run on the GPU box or in tests).

  sig_t_mean / sig_t           openpoints/models/backbone/transformer.py:1071-1131   (class source exec'd as is)
  threeD_space_loss,
  feature_space_loss,
  Idenyity_loss                utils/insT_loss.py                                     (module source exec'd as is)
  knn_point                    openpoints/models/layers/knn.py:7-20                   (module source exec'd as is)
  train.py:505-557             inline code in train_one_epoch: the statement range is cut from the file and
                               exec'd in a namespace holding its free variables (no edits to the statements)

The reference hard-codes `.cuda()`; on this GPU-less container `torch.Tensor.cuda` / `nn.Module.cuda` are
patched to identity for the duration of the run — the arithmetic is untouched.

    python tests/golden/make_transition_golden.py
"""
import ast
import os
import sys
import textwrap
import types

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

REF = os.environ.get("GEOT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

torch.Tensor.cuda = lambda self, *a, **k: self
nn.Module.cuda = lambda self, *a, **k: self


def class_source(path, names):
    src = open(path).read()
    tree = ast.parse(src)
    out = {}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name in names:
            node.decorator_list = []  # @MODELS.register_module() is registry plumbing, not arithmetic
            out[node.name] = ast.unparse(node)
    return out


def load_reference():
    ns = {"torch": torch, "nn": nn, "F": F}
    for name, code in class_source(os.path.join(REF, "openpoints/models/backbone/transformer.py"), {"sig_t_mean", "sig_t"}).items():
        exec(compile(code, f"<reference transformer.py:{name}>", "exec"), ns)
    knn_ns = {}
    exec(compile(open(os.path.join(REF, "openpoints/models/layers/knn.py")).read(), "<reference knn.py>", "exec"), knn_ns)
    # utils/insT_loss.py imports two modules that need CUDA extensions; give it the two names it uses.
    layers = types.ModuleType("openpoints.models.layers")
    layers.knn_point = knn_ns["knn_point"]
    layers.torch_grouping_operation = None  # imported, unused
    for m in ("openpoints", "openpoints.models"):
        sys.modules.setdefault(m, types.ModuleType(m))
    sys.modules["openpoints.models.layers"] = layers
    po_f = types.ModuleType("pointops.functions")
    po_f.pointops = types.ModuleType("pointops.functions.pointops")
    sys.modules.setdefault("pointops", types.ModuleType("pointops"))
    sys.modules["pointops.functions"] = po_f
    loss_ns = {}
    exec(compile(open(os.path.join(REF, "utils/insT_loss.py")).read(), "<reference insT_loss.py>", "exec"), loss_ns)
    return ns, knn_ns, loss_ns


def train_block(first, last):
    """Lines [first, last] of examples/segmentation/train.py, dedented."""
    lines = open(os.path.join(REF, "examples/segmentation/train.py")).read().split("\n")
    return textwrap.dedent("\n".join(lines[first - 1:last]))


def softmax_field(g, B, C, N, sharp=3.0):
    return F.softmax(torch.from_numpy(g.standard_normal((B, C, N)).astype(np.float32)) * sharp, dim=1)


def main():
    ns, knn_ns, loss_ns = load_reference()
    gaussian_src = train_block(835, 836)
    out = {}
    for C, B, N, k in [(17, 2, 64, 7), (33, 2, 16, 5), (5, 1, 40, 3)]:
        g = np.random.default_rng(1609 + C)
        tag = f"C{C}"
        torch.manual_seed(C)
        # ---- sig_t_mean: reference module, reference init (nn.Linear default), reference forward ----
        mod = ns["sig_t_mean"](C)
        W = torch.stack([mod.fc[kk].weight.detach() for kk in range(C)])  # (C, C, 2C)
        x = softmax_field(g, B, C, N).requires_grad_(True)
        cm = F.normalize(torch.from_numpy(g.random((C, C)).astype(np.float32)) + 0.01, p=1, dim=1)
        insT = mod(x, cm)
        go = torch.from_numpy(g.standard_normal(tuple(insT.shape)).astype(np.float32))
        gW = torch.autograd.grad(insT, [mod.fc[kk].weight for kk in range(C)] + [x], go, retain_graph=True)
        out.update({f"{tag}_W": W, f"{tag}_x": x.detach(), f"{tag}_cm": cm, f"{tag}_insT": insT.detach(), f"{tag}_go": go,
                    f"{tag}_gW": torch.stack(gW[:C]), f"{tag}_gx": gW[C]})
        # ---- sig_t ----
        mod2 = ns["sig_t"](C)
        out[f"{tag}_sigt_W"] = mod2.fc.weight.detach()
        out[f"{tag}_sigt_T"] = mod2(x.detach()).detach()
        # ---- train.py:549-552 per-point application, statements exec'd as they are ----
        ema_t_corr = F.normalize(torch.from_numpy(g.random((C, C)).astype(np.float32)) + torch.eye(C) * 3, p=1, dim=1)
        pred_u_strong = torch.from_numpy(g.standard_normal((B, C, N)).astype(np.float32) * 2).requires_grad_(True)
        insT_l = insT.detach().clone().requires_grad_(True)
        env = {"torch": torch, "F": F, "cfg": types.SimpleNamespace(lambma=0.9, num_classes=C), "ema_t_corr": ema_t_corr,
               "insT": insT_l, "pred_u_strong": pred_u_strong, "batch_size_u": B, "num_point": N}
        exec(train_block(549, 552), env)
        corr = env["pred_u_strong_corr"]
        go2 = torch.from_numpy(g.standard_normal((B, C, N)).astype(np.float32))
        g_log, g_ins = torch.autograd.grad(corr, [pred_u_strong, insT_l], go2)
        out.update({f"{tag}_ema_t_corr": ema_t_corr, f"{tag}_logits": pred_u_strong.detach(), f"{tag}_corr": corr.detach(),
                    f"{tag}_go2": go2, f"{tag}_g_logits": g_log, f"{tag}_g_insT_apply": g_ins})
        # ---- manifold losses: reference modules ----
        pos = torch.from_numpy(g.random((B, N, 3)).astype(np.float32))
        labels = torch.from_numpy(g.integers(0, 3, (B, N)))
        insT_m = insT.detach().clone().requires_grad_(True)
        _, idx3 = knn_ns["knn_point"](k + 1, pos, pos)
        l3 = loss_ns["threeD_space_loss"](k=k, sigma=0.5, num_classes=C)(pos, labels, insT_m)
        g3, = torch.autograd.grad(l3, insT_m)
        logit_f = pred_u_strong.detach()
        _, idxf = knn_ns["knn_point"](k + 1, logit_f.permute(0, 2, 1).contiguous(), logit_f.permute(0, 2, 1).contiguous())
        lf = loss_ns["feature_space_loss"](k=k, sigma=2.0, num_classes=C)(logit_f, labels, insT_m)
        gf, = torch.autograd.grad(lf, insT_m)
        Id = torch.eye(C).unsqueeze(0)
        li = loss_ns["Idenyity_loss"]()(insT_m, Id)
        out.update({f"{tag}_pos": pos, f"{tag}_labels": labels, f"{tag}_k": torch.tensor(k), f"{tag}_idx3": idx3[:, :, 1:],
                    f"{tag}_loss3": l3.detach(), f"{tag}_g_insT_loss3": g3, f"{tag}_idxf": idxf[:, :, 1:], f"{tag}_lossf": lf.detach(),
                    f"{tag}_g_insT_lossf": gf, f"{tag}_loss_id": li.detach()})
        # ---- class-level statistics: train.py:505-545,556-557 exec'd as they are (C=17: LABEL_PROJ train.py:48) ----
        if C == 17:
            pred_u = softmax_field(g, B, C, N, sharp=1.5)
            ema_t = F.normalize(torch.from_numpy(g.random((C, C)).astype(np.float32)) + torch.eye(C) * 5, p=1, dim=1)
            sigma = torch.full((C,), 0.4) + torch.from_numpy(g.random(C).astype(np.float32)) * 0.2
            env = {"torch": torch, "F": F}
            exec(gaussian_src, env)
            exec(train_block(48, 48), env)  # LABEL_PROJ
            env.update({"cfg": types.SimpleNamespace(num_classes=C, filter_outlier=False, geo_lambma=0.999, ema_t_decay=0.999),
                        "pred_u": pred_u, "sigma": sigma, "ema_t": ema_t.clone(), "batch_size_u": B, "num_point": N})
            exec(train_block(505, 545), env)
            exec(train_block(556, 557), env)
            out.update({f"{tag}_pred_u": pred_u, f"{tag}_ema_t_in": ema_t, f"{tag}_sigma": sigma, f"{tag}_class_T": env["class_T"],
                        f"{tag}_ema_t_corr_out": env["ema_t_corr"], f"{tag}_ema_t_out": env["ema_t"],
                        f"{tag}_label_proj": torch.tensor(env["LABEL_PROJ"])})
            # cal_mean_feature inner loop, train.py:884-895
            probs = pred_u.permute(0, 2, 1).contiguous().view(B * N, C)
            target = torch.from_numpy(g.integers(0, C - 2, (B * N,)))
            env = {"torch": torch, "c": C, "logits": probs, "target": target, "cm": torch.zeros((C, C)), "c_num": torch.zeros((C,))}
            for _ in range(2):
                exec(train_block(887, 895), env)
            out.update({f"{tag}_mf_probs": probs, f"{tag}_mf_target": target, f"{tag}_mf_cm": env["cm"], f"{tag}_mf_cnum": env["c_num"]})
    path = os.path.join(HERE, "transition_golden.npz")
    np.savez_compressed(path, **{k: (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
