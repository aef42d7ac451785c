#!/usr/bin/env python
"""Dev tool (GPU box): launch ONE op of the bench workload a few times so ncu can capture it.
usage: python tools/prof_one.py {fps|fps_noprune|ball|group|group3|groupxyz|groupbwd|plan|plan3|threenn|tibwd|ti|ti_old|gf|upsample|knn|knntie|sigt|apply|manifold} [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from geot_b200 import _lib, synth
from geot_b200 import functional as F

what = sys.argv[1]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
N, M, NS, C = 16384, 1024, 32, 64
xyz = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
feat = torch.randn(B, C, N, device="cuda")
idx = F.furthest_point_sampling(xyz, M)
xyz_t = xyz.transpose(1, 2).contiguous()
centres = F.gather_points(xyz_t, idx).transpose(1, 2).contiguous()
nbr = F.ball_query(centres, xyz, 0.1, NS)
if what == "groupbwd":
    go = torch.randn(B, C, M, NS, device="cuda")
    plan = F.ScatterPlan(nbr, B, N)
if what in ("threenn", "tibwd", "ti", "ti_old", "plan3"):
    d2, i3 = F.three_nn(xyz, centres)
    w = 1.0 / (d2 + 1e-8); w = (w / w.sum(-1, keepdim=True)).contiguous()
    go3 = torch.randn(B, C, N, device="cuda")
    plan3 = F.ScatterPlan(i3, B, M)
if what == "knntie":  # self-kNN k=33 on the cfg4 clouds under the reference's heap semantics: one tied query is replayed
    tx = torch.from_numpy(synth.tooth_batch(2, 16000)).cuda().view(-1, 3).contiguous()
if what == "knn":
    kb = 8
    kx = xyz[:kb].reshape(-1, 3).contiguous()
    koff = torch.arange(1, kb + 1, device="cuda", dtype=torch.int32) * N
if what in ("sigt", "manifold", "apply"):
    from geot_b200 import transition as T
    Ct, Bt, Nt = 33, 2, 16000
    logits = torch.randn(Bt, Ct, Nt, device="cuda") * 3
    pt = logits.softmax(1)
    cmt = torch.rand(Ct, Ct, device="cuda"); cmt = cmt / cmt.sum(1, keepdim=True)
    modt = T.sig_t_mean(Ct).cuda()
    got = torch.randn(Bt * Nt, Ct, Ct, device="cuda")
    post = xyz[:Bt, :Nt].contiguous()
    labt = logits.argmax(1)
    insTt = modt(pt, cmt).detach().requires_grad_(True)
    losst = T.threeD_space_loss(k=32, sigma=1.0, num_classes=Ct)
    idxt = T._neighbours(32, post)
    emat = torch.rand(Ct, Ct, device="cuda"); emat = emat / emat.sum(1, keepdim=True)
    lgt = logits.clone().requires_grad_(True)
    go2t = torch.randn(Bt, Ct, Nt, device="cuda")
if what == "gf":  # DGCNN edge features at cfg4's last propagation level
    gx_k = torch.randn(6, 384, 4096, device="cuda"); gx_q = torch.randn(6, 384, 8192, device="cuda")
    gidx = torch.randint(0, 4096, (6, 8192, 4), device="cuda", dtype=torch.int32)
if what == "upsample":  # validation: 16000 sampled points -> 100000 mesh vertices, 33 classes
    from geot_b200 import validation as VAL
    upts = xyz[:1, :16000].contiguous()
    umesh = (upts[:, torch.randint(0, 16000, (100000,), device="cuda")] + 0.002 * torch.randn(1, 100000, 3, device="cuda")).contiguous()
    uprob = torch.softmax(torch.randn(1, 33, 16000, device="cuda") * 3, dim=1)
torch.cuda.synchronize()
for _ in range(3):
    if what == "fps": F.furthest_point_sampling(xyz, M)
    elif what == "fps_noprune": F.furthest_point_sampling(xyz, M, flags=_lib.GEOT_FPS_NO_PRUNE)
    elif what == "ball": F.ball_query(centres, xyz, 0.1, NS)
    elif what == "groupxyz": F.group_xyz_relative(xyz, centres, nbr, relative=True, divisor=0.1)
    elif what == "group": F.group_points(feat, nbr)
    elif what == "group3": F.group_points(xyz_t, nbr)
    elif what == "groupbwd": F.group_points_grad(go, plan, N)
    elif what == "plan": F.ScatterPlan(nbr, B, N)
    elif what == "plan3": F.ScatterPlan(i3, B, M)
    elif what == "threenn": F.three_nn(xyz, centres)
    elif what == "tibwd": F.three_interpolate_grad(go3, w, plan3, M)
    elif what == "gf": F.graph_feature(gx_k, gx_q, gidx)
    elif what == "upsample": VAL.upsample_labels(uprob, upts, umesh)
    elif what in ("ti", "ti_old"):  # (B, C, M) -> (B, C, N); ti_old = without the channel-interleaved kernel
        _lib.lib().geot_group_points_config(3 if what == "ti_old" else 0)
        F.three_interpolate(feat[:, :, :M].contiguous(), i3, w)
    elif what == "knn": F.knnquery(16, kx, kx, koff, koff, m_per=N)
    elif what == "knntie": F.knnquery(33, tx, tx, None, None, n_per=16000, m_per=16000)
    elif what == "sigt":
        modt.zero_grad(set_to_none=True); modt(pt, cmt).backward(got)
    elif what == "apply":
        lgt.grad = None; insTt.grad = None; T.apply_transition(lgt, emat, insTt, 0.9).backward(go2t)
    elif what == "manifold":
        insTt.grad = None; losst(post, labt, insTt, idx=idxt).backward()
torch.cuda.synchronize()
