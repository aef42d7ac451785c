#!/usr/bin/env python
"""Generates tests/golden/pointops_golden.npz ON THE GPU BOX by running the REFERENCE'S OWN CUDA kernels
(oracle/_ref/*.so: the reference's .cu files compiled unchanged, oracle/ref_build/Makefile) on seeded inputs.
Only outputs are stored; inputs are regenerated from geot_b200.synth by (kind, N, seed).

    gpurun -- 'python tests/golden/make_pointops_golden.py gpurun_out/pointops_golden.npz'
then copy the file into tests/golden/.  tests/test_oracle_golden_cpu.py checks oracle/cpu_ref.c against it.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from geot_b200 import synth  # noqa: E402
from tests import refgpu     # noqa: E402

CASES = [("tooth", 2, 2048, 256), ("quant", 2, 1500, 300), ("dup", 1, 1200, 400), ("origin", 2, 1000, 128)]


def clouds(kind, B, N, seed=0):
    if kind == "tooth":
        return synth.tooth_batch(B, N, first_cloud=seed)
    return np.stack([synth.adversarial_cloud(N, 300 + seed + i, kind) for i in range(B)])


def features(shape, seed):
    return np.random.default_rng(seed).standard_normal(shape).astype(np.float32)


def main(path):
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = {}
    for ci, (kind, B, N, M) in enumerate(CASES):
        t = f"c{ci}"
        xyz = clouds(kind, B, N)
        x = cu(xyz)
        fb = refgpu.fps(x, M, "pn2batch")
        fe = refgpu.fps(x, M, "pn2ext")
        out[f"{t}_fps_batch"], out[f"{t}_fps_ext"] = fb.cpu().numpy(), fe.cpu().numpy()
        centres = torch.gather(x, 1, fb.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
        r, ns = (0.1, 32) if kind == "tooth" else (0.125, 24)
        bq = refgpu.ball_query(r, ns, x, centres)
        out[f"{t}_ball"] = bq.cpu().numpy()
        d2, i3 = refgpu.three_nn(x, centres)
        out[f"{t}_nn_d2"], out[f"{t}_nn_idx"] = d2.cpu().numpy(), i3.cpu().numpy()
        feat = cu(features((B, 5, N), 10 + ci))
        out[f"{t}_group"] = refgpu.group_points(feat, bq).cpu().numpy()
        go = cu(features((B, 5, M, ns), 20 + ci))
        out[f"{t}_group_grad"] = refgpu.group_points_grad(go, bq, N).cpu().numpy()
        out[f"{t}_gather"] = refgpu.gather_points(feat, fb).cpu().numpy()
        w = torch.softmax(cu(features((B, N, 3), 30 + ci)), -1).contiguous()
        kf = cu(features((B, 5, M), 40 + ci))
        out[f"{t}_interp"] = refgpu.three_interpolate(kf, i3, w).cpu().numpy()
        out[f"{t}_interp_grad"] = refgpu.three_interpolate_grad(cu(features((B, 5, N), 50 + ci)), i3, w, M).cpu().numpy()
        # offset format
        flat = cu(xyz.reshape(-1, 3))
        offset = torch.tensor([N * (i + 1) for i in range(B)], dtype=torch.int32).cuda()
        new_offset = torch.tensor([M * (i + 1) for i in range(B)], dtype=torch.int32).cuda()
        fo = refgpu.furthestsampling(flat, offset, new_offset)
        out[f"{t}_fps_offset"] = fo.cpu().numpy()
        wts = cu(np.random.default_rng(60 + ci).random(B * N).astype(np.float32) + 0.05)
        out[f"{t}_fps_offset_w"] = refgpu.furthestsampling(flat, offset, new_offset, wts).cpu().numpy()
        q = flat[fo.long()].contiguous()
        for k in (3, 16):
            ki, kd = refgpu.knnquery(k, flat, q, offset, new_offset)
            out[f"{t}_knn{k}_idx"], out[f"{t}_knn{k}_d2"] = ki.cpu().numpy(), kd.cpu().numpy()
        out[f"{t}_ball_offset"] = refgpu.ballquery_offset(r, ns, flat, q, offset, new_offset).cpu().numpy()
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", len(out), "arrays")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "pointops_golden.npz"))
