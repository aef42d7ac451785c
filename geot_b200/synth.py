"""Synthetic "tooth-scan-shaped" clouds (numpy, CPU) for tests and bench.py.

No Teeth3DS data exists on the build or GPU boxes (cfgs/tooth_semi/default.yaml:4,18 point at an
unmounted path), so the workload is generated: a dental arch (parabola centre-line in the xy plane),
a gingiva band along it (label 0) and 16 super-ellipsoid crowns at equal arc-length slots (labels
1..16 for the lower jaw, 17..32 for the upper), surface-sampled, jittered, SHUFFLED (the dataset's
np.random.choice order, tooth_dataset.py:134-137, matters to ball_query's first-nsample rule) and then
centred and scaled to the unit sphere exactly like pc_norm (tooth_dataset.py:108-114).
Seed convention: 1609 (cfgs/tooth_semi/transformer_finetune_fixmatch_ntm.yaml:67) + cloud index.
"""
from __future__ import annotations

import numpy as np

BASE_SEED = 1609


def _arch(t, width=60.0, depth=45.0):
    """Centre-line of the arch for t in [-1, 1]: x = w/2 * t, y = depth * (1 - t^2)."""
    return np.stack([0.5 * width * t, depth * (1.0 - t * t), np.zeros_like(t)], -1)


def tooth_cloud(n_points=16384, seed=BASE_SEED, jaw=0, return_labels=False):
    """One cloud: (n_points, 3) float32 in the unit sphere (+ labels (n_points,) int32)."""
    rng = np.random.default_rng(seed)
    n_gum = int(n_points * 0.35)
    n_teeth = n_points - n_gum
    # gingiva: band around the centre-line, |t| <= 1.05, height in [-9, 0] mm, both flanks
    t = rng.uniform(-1.05, 1.05, n_gum)
    c = _arch(t)
    tang = np.stack([np.full_like(t, 30.0), -90.0 * t, np.zeros_like(t)], -1)
    tang /= np.linalg.norm(tang, axis=1, keepdims=True)
    nrm = np.stack([-tang[:, 1], tang[:, 0], np.zeros_like(t)], -1)
    side = rng.choice([-1.0, 1.0], n_gum)
    h = rng.uniform(-9.0, 0.0, n_gum)
    bulge = 5.5 + 1.5 * np.cos(h / 9.0 * np.pi)
    gum = c + nrm * (side * bulge)[:, None]
    gum[:, 2] = h
    lab_gum = np.zeros(n_gum, np.int32)
    # teeth: 16 super-ellipsoid crowns at equal slots; molars wider than incisors
    slot = rng.integers(0, 16, n_teeth)
    tc = -0.94 + (slot + 0.5) * (1.88 / 16.0)
    centre = _arch(tc)
    central = 1.0 - np.abs(tc)  # 1 at incisors, 0 at molars
    rx = 3.0 + 2.3 * (1 - central)
    ry = 2.6 + 2.6 * (1 - central)
    rz = 4.5 + 1.0 * central
    u = rng.uniform(0, 2 * np.pi, n_teeth)
    v = np.arccos(rng.uniform(-0.2, 1.0, n_teeth))  # mostly the crown's upper part
    e = 0.7
    sgn = lambda a: np.sign(a) * np.abs(a) ** e
    px = rx * sgn(np.cos(u)) * np.abs(np.sin(v)) ** e
    py = ry * sgn(np.sin(u)) * np.abs(np.sin(v)) ** e
    pz = rz * sgn(np.cos(v))
    ang = np.arctan2(-90.0 * tc, 30.0)
    ca, sa = np.cos(ang), np.sin(ang)
    teeth = centre + np.stack([ca * px - sa * py, sa * px + ca * py, pz + 2.0], -1)
    lab_teeth = (slot + 1 + (16 if jaw else 0)).astype(np.int32)
    pts = np.concatenate([gum, teeth], 0)
    lab = np.concatenate([lab_gum, lab_teeth], 0)
    pts = pts + rng.normal(0.0, 0.05, pts.shape)
    if jaw:
        pts[:, 2] = -pts[:, 2]
    perm = rng.permutation(n_points)
    pts, lab = pts[perm], lab[perm]
    pts = pts - pts.mean(0)
    pts = pts / np.max(np.sqrt((pts ** 2).sum(1)))
    pts = pts.astype(np.float32)
    return (pts, lab) if return_labels else pts


def tooth_batch(B, n_points=16384, first_cloud=0, return_labels=False):
    """(B, n_points, 3) float32; cloud i uses seed 1609 + first_cloud + i and jaw = i % 2."""
    out = [tooth_cloud(n_points, BASE_SEED + first_cloud + i, (first_cloud + i) % 2, True) for i in range(B)]
    xyz = np.stack([o[0] for o in out], 0)
    if return_labels:
        return xyz, np.stack([o[1] for o in out], 0)
    return xyz


def adversarial_cloud(n_points, seed, kind):
    """Parity edge cases: 'dup' = sampled with replacement (duplicate points),
    'quant' = coordinates quantised to 2^-6 (exact distance ties), 'origin' = a few points at/near 0
    (mag-skip of pointnet2/_ext)."""
    rng = np.random.default_rng(seed)
    pts = tooth_cloud(n_points, seed)
    if kind == "dup":
        pts = pts[rng.integers(0, max(n_points // 3, 1), n_points)]
    elif kind == "quant":
        pts = np.round(pts * 64.0) / 64.0
    elif kind == "origin":
        k = max(n_points // 50, 1)
        pts[rng.choice(n_points, k, replace=False)] = rng.normal(0, 0.01, (k, 3)).astype(np.float32)
        pts[0] = 0.0
    else:
        raise ValueError(kind)
    return np.ascontiguousarray(pts, dtype=np.float32)
