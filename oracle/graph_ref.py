"""oracle/graph_ref.py — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Torch restatement of DGCNN_Propagation.get_graph_feature (openpoints/models/backbone/transformer.py:343-364) with the
neighbour search through oracle/cpu_ref.c's kNN (knn_cuda is un-vendored: parity of the neighbour ORDER is unpinned,
SURVEY §8c; ascending distance is what upstream documents).  Pinned against the reference method's own source run on
CPU by tests/golden/make_graph_golden.py (fixture tests/golden/graph_golden.npz).  Paths relative to the reference root.
"""
from __future__ import annotations

import numpy as np
import torch

from . import cpu_ref


def knn_idx(k, coor_k, coor_q):
    """knn_cuda.KNN(k, transpose_mode=False)(coor_k, coor_q): (B,3,Nk), (B,3,Nq) -> idx (B,k,Nq) int64, cloud-local."""
    B, _, Nk = coor_k.shape
    Nq = coor_q.shape[2]
    ref = coor_k.transpose(1, 2).contiguous().view(-1, 3).numpy()
    qry = coor_q.transpose(1, 2).contiguous().view(-1, 3).numpy()
    off = (np.arange(1, B + 1) * Nk).astype(np.int32)
    noff = (np.arange(1, B + 1) * Nq).astype(np.int32)
    idx = cpu_ref.knnquery(k, ref, qry, off, noff)[0]
    idx = torch.from_numpy(idx.astype(np.int64)).view(B, Nq, k) - (torch.arange(B) * Nk).view(B, 1, 1)
    return idx.transpose(1, 2).contiguous()


def get_graph_feature(x_q, x_k, idx):
    """transformer.py:355-363 given idx (B,k,Nq) int64 cloud-local: -> (B,2C,Nq,k)."""
    B, C, Nk = x_k.shape
    Nq = x_q.shape[2]
    k = idx.shape[1]
    flat = (idx + (torch.arange(B, device=idx.device) * Nk).view(-1, 1, 1)).view(-1)
    rows = x_k.transpose(2, 1).contiguous().view(B * Nk, -1)[flat, :]
    feature = rows.view(B, k, Nq, C).permute(0, 3, 2, 1).contiguous()
    xq = x_q.view(B, C, Nq, 1).expand(-1, -1, -1, k)
    return torch.cat((feature - xq, xq), dim=1)
