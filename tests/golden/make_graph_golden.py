#!/usr/bin/env python
"""Generates tests/golden/graph_golden.npz by running the REFERENCE'S OWN DGCNN_Propagation.get_graph_feature
(openpoints/models/backbone/transformer.py:343-364; the method source is cut from the class and exec'd as it is, with a
`self` that carries k and a CPU knn - knn_cuda is not in the reference tree) on CPU, in the authoring container
(needs /root/reference; never run in tests or on the GPU box).

    python tests/golden/make_graph_golden.py
"""
import ast
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("GEOT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import graph_ref  # noqa: E402


def method_source(path, cls, name):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for sub in node.body:
                if isinstance(sub, ast.FunctionDef) and sub.name == name:
                    sub.decorator_list = []
                    return ast.unparse(sub)
    raise KeyError(name)


ns = {"torch": torch}
exec(compile(method_source(os.path.join(REF, "openpoints/models/backbone/transformer.py"), "DGCNN_Propagation", "get_graph_feature"),
             "<reference transformer.py>", "exec"), ns)
out = {}
g = torch.Generator().manual_seed(343)
for tag, (B, C, Nk, Nq, k) in {"k4": (2, 24, 96, 200, 4), "k4self": (3, 7, 130, 130, 4), "k16": (1, 5, 64, 33, 16)}.items():
    coor_k = torch.rand(B, 3, Nk, generator=g)
    coor_q = coor_k.clone() if Nq == Nk else torch.rand(B, 3, Nq, generator=g)
    x_k = torch.randn(B, C, Nk, generator=g).requires_grad_(True)
    x_q = torch.randn(B, C, Nq, generator=g).requires_grad_(True)
    me = types.SimpleNamespace(k=k, knn=lambda ck, cq, k=k: (None, graph_ref.knn_idx(k, ck, cq)))
    feat = ns["get_graph_feature"](me, coor_q, x_q, coor_k, x_k)
    go = torch.randn(feat.shape, generator=g)
    gk, gq = torch.autograd.grad(feat, [x_k, x_q], go)
    for name, v in dict(coor_k=coor_k, coor_q=coor_q, x_k=x_k, x_q=x_q, feat=feat, go=go, gk=gk, gq=gq).items():
        out[f"{tag}_{name}"] = v.detach().numpy()
    out[f"{tag}_k"] = np.int32(k)
np.savez_compressed(os.path.join(HERE, "graph_golden.npz"), **out)
print("wrote", os.path.join(HERE, "graph_golden.npz"), len(out), "arrays", {t: out[f"{t}_feat"].shape for t in ("k4", "k4self", "k16")})
