"""DGCNN edge features on the GPU (geot_graph_feature / geot_graph_feature_grad_q through the C ABI) against the golden
fixture of the reference method and against oracle/graph_ref.py at GeoT's sizes.  The forward is one fp32 subtraction
per element: bit-exact; gradients within 1e-5 relative (north_star), deterministic run to run."""
import os

import numpy as np
import pytest
import torch

from geot_b200 import functional as F_
from geot_b200 import layers
from oracle import graph_ref as GR

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "graph_golden.npz"))


def t(name, dev="cuda"):
    return torch.from_numpy(G[name]).to(dev)


def rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("tag", ["k4", "k4self", "k16"])
def test_golden_get_graph_feature(cuda, tag):
    k = int(G[f"{tag}_k"])
    x_k, x_q = t(f"{tag}_x_k").requires_grad_(True), t(f"{tag}_x_q").requires_grad_(True)
    feat = layers.get_graph_feature(t(f"{tag}_coor_q"), x_q, t(f"{tag}_coor_k"), x_k, k=k)
    assert torch.equal(feat.cpu(), t(f"{tag}_feat", "cpu"))  # same neighbours (random clouds: no ties), same subtraction
    gk, gq = torch.autograd.grad(feat, [x_k, x_q], t(f"{tag}_go"))
    assert rel(gk, t(f"{tag}_gk", "cpu")) < 1e-5
    assert rel(gq, t(f"{tag}_gq", "cpu")) < 1e-5


@pytest.mark.parametrize("B,C,Nk,Nq,k", [(6, 384, 512, 4096, 4), (2, 64, 4096, 8192, 4), (1, 16, 30000, 1000, 4),
                                         (2, 8, 1001, 777, 5), (1, 3, 8, 5, 8), (2, 4, 50000, 64, 4)])
def test_graph_feature_vs_oracle_expression(cuda, B, C, Nk, Nq, k):
    """staged and direct kernels, k = 4 and generic k, unaligned rows; against the reference expression run on the GPU."""
    g = torch.Generator(device="cuda").manual_seed(Nk + Nq)
    x_k = torch.randn(B, C, Nk, device="cuda", generator=g).requires_grad_(True)
    x_q = torch.randn(B, C, Nq, device="cuda", generator=g).requires_grad_(True)
    idx = torch.randint(0, Nk, (B, Nq, k), device="cuda", generator=g, dtype=torch.int32)
    idx[:, : min(Nq, 7)] = 0  # heavy collisions on one target
    out = layers.GraphFeature.apply(x_k, x_q, idx)
    ref = GR.get_graph_feature(x_q, x_k, idx.long().transpose(1, 2).contiguous())
    assert out.shape == (B, 2 * C, Nq, k) and torch.equal(out, ref)
    go = torch.randn(out.shape, device="cuda", generator=g)
    gk, gq = torch.autograd.grad(out, [x_k, x_q], go)
    rk, rq = torch.autograd.grad(ref, [x_k, x_q], go)
    assert rel(gk, rk) < 1e-5 and rel(gq, rq) < 1e-5
    gk2, gq2 = torch.autograd.grad(layers.GraphFeature.apply(x_k, x_q, idx), [x_k, x_q], go)
    assert torch.equal(gk, gk2) and torch.equal(gq, gq2)  # no float atomics: run-to-run identical bits


def test_knn_neighbours_match_compat_knn(cuda):
    from geot_b200 import compat
    g = torch.Generator(device="cuda").manual_seed(4)
    coor_k = torch.rand(3, 3, 2048, device="cuda", generator=g)
    coor_q = torch.rand(3, 3, 4096, device="cuda", generator=g)
    idx = layers.knn_neighbours(4, coor_q, coor_k)
    _, ref = compat.KNN(k=4, transpose_mode=False)(coor_k, coor_q)  # (B,k,Nq) int64
    assert torch.equal(idx.long(), ref.transpose(1, 2))
    oracle = GR.knn_idx(4, coor_k.cpu(), coor_q.cpu())
    assert torch.equal(ref.cpu(), oracle)


@pytest.mark.parametrize("B,N,G,M", [(6, 16000, 512, 32), (2, 1000, 37, 5)])
def test_group_fps_knn_patches(cuda, B, N, G, M):
    """layers.Group (transformer.py:274-303) against the reference's expressions built from the same library's FPS and
    KNN front ends: centres = xyz[fps_idx], patches = xyz.view(B*N, 3)[idx + idx_base] - centre; one subtraction per
    element: bit-exact."""
    from geot_b200 import compat, pointnet2_utils as pu, synth
    xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
    nb, center, idx = layers.Group(G, M)(xyz)
    fps_idx = pu.furthest_point_sample(xyz, G)
    center_ref = pu.gather_operation(xyz.transpose(1, 2).contiguous(), fps_idx).transpose(1, 2).contiguous()
    _, kidx = compat.KNN(k=M, transpose_mode=True)(xyz, center_ref)  # B G M
    flat = (kidx + torch.arange(0, B, device="cuda").view(-1, 1, 1) * N).view(-1)
    nb_ref = xyz.view(B * N, -1)[flat, :].view(B, G, M, 3).contiguous() - center_ref.unsqueeze(2)
    assert torch.equal(center, center_ref)
    assert idx.dtype == torch.int64 and torch.equal(idx, flat)
    assert nb.shape == (B, G, M, 3) and torch.equal(nb, nb_ref)
