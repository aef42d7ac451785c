"""oracle/graph_ref.py against the fixture produced by the REFERENCE'S OWN DGCNN_Propagation.get_graph_feature
(tests/golden/make_graph_golden.py ran the method source of transformer.py:343-364 on CPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import graph_ref as GR

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "graph_golden.npz"))
CASES = ["k4", "k4self", "k16"]


def t(name):
    return torch.from_numpy(G[name])


@pytest.mark.parametrize("tag", CASES)
def test_get_graph_feature_and_grads(tag):
    k = int(G[f"{tag}_k"])
    x_k, x_q = t(f"{tag}_x_k").requires_grad_(True), t(f"{tag}_x_q").requires_grad_(True)
    idx = GR.knn_idx(k, t(f"{tag}_coor_k"), t(f"{tag}_coor_q"))
    feat = GR.get_graph_feature(x_q, x_k, idx)
    assert torch.equal(feat, t(f"{tag}_feat"))
    gk, gq = torch.autograd.grad(feat, [x_k, x_q], t(f"{tag}_go"))
    torch.testing.assert_close(gk, t(f"{tag}_gk"), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(gq, t(f"{tag}_gq"), rtol=1e-6, atol=1e-6)
