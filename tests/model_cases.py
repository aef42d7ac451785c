"""BASELINE.json configs 2, 3 and 4 as runnable cases built from the reference's OWN Python (tests/refenv.py),
run once per backend: 'geot' = geot_b200.backends (what compat.install() registers), 'ref' = the reference's own
kernels (oracle/_ref/*.so).  Shared by tests/test_models_gpu.py (parity) and bench.py --config cfgN (timing).
Test infrastructure: nothing in geot_b200/ imports this.

  cfg2  Point-Transformer op pyramid, offset format, B=8 x 16384, k=16     SURVEY.md Appendix C
        (openpoints/cpp/pointops/functions/pointops.py:10-29,32-50,73-103,151-172,262-299)
  cfg3  PointNet++ MSG segmentation, B=16 x 16384                          SURVEY.md Appendix B
        (openpoints/models/backbone/pointnetv2.py:69-100,322-339,375-380; layers/local_aggregation.py:205-242)
  cfg4  GeoT semi-supervised step: WholePartSeg(PointTransformer_seg_T) on 2 labeled + 2 strong + 2 weak clouds of
        16000 points + the transition-matrix block (examples/segmentation/train.py:505-596)
"""
from __future__ import annotations

import os
import textwrap
import types

import numpy as np
import torch
import torch.nn.functional as F

from tests import refenv
from geot_b200 import synth


def _cuda(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def param_grads(model):
    return {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}


def rel_err(a, b):
    """max |a-b| / max |b| (what 'within 1e-5 relative' means for a whole tensor)."""
    d = (a.double() - b.double()).abs().max().item()
    s = b.double().abs().max().item()
    return d / s if s > 0 else d


# ------------------------------------------------------------------------------------------------- cfg3
def cfg3_model(num_classes=33):
    refenv.install(refenv.backend() or "geot")
    from openpoints.models import build_model_from_cfg
    from openpoints.utils import EasyConfig
    enc = dict(NAME='PointNet2Encoder', in_channels=6, width=None, strides=[4, 4, 4, 4], layers=3, use_res=False, sampler='fps',
               mlps=[[[16, 16, 32], [32, 32, 64], [32, 48, 64]], [[64, 64, 128], [64, 96, 128]],
                     [[128, 196, 256], [128, 196, 256]], [[256, 256, 512], [256, 384, 512]]],
               radius=[[0.05, 0.1, 0.2], [0.1, 0.2], [0.2, 0.4], [0.4, 0.8]], num_samples=[[16, 32, 64], [16, 32], [16, 32], [16, 32]],
               aggr_args=dict(NAME='convpool', feature_type='dp_fj', anisotropic=False, reduction='max'),
               group_args=dict(NAME='ballquery', use_xyz=True, normalize_dp=False),
               conv_args=dict(order='conv-norm-act'), act_args=dict(act='relu'), norm_args=dict(norm='bn'))
    seg = dict(NAME='BaseSeg', encoder_args=enc,
               decoder_args=dict(NAME='PointNet2Decoder', fp_mlps=[[128, 128, 128], [256, 128], [256, 256], [256, 256]]),
               cls_args=dict(NAME='SegHead', num_classes=num_classes, in_channels=None))
    cfg = EasyConfig()
    cfg.update(seg)
    torch.manual_seed(1609)
    return build_model_from_cfg(cfg).cuda()


def cfg3_inputs(B=16, N=16384):
    xyz = synth.tooth_batch(B, N)
    g = np.random.default_rng(3)
    normals = g.standard_normal((B, 3, N)).astype(np.float32)
    pos = _cuda(xyz)
    x = torch.cat([pos.transpose(1, 2), _cuda(normals)], 1).contiguous()       # (B, 6, N): xyz + 3 extra channels
    return {"pos": pos, "x": x}


def run_cfg3(backend, model, data, gout=None, backward=True, seed=7):
    """One forward (+ backward against a fixed cotangent) of the SAME model object on `backend`."""
    refenv.use(backend)
    model.train()
    model.zero_grad(set_to_none=True)
    torch.manual_seed(seed)                       # identical dropout masks on both backends
    with refenv.recording() as rec:
        logits = model({"pos": data["pos"], "x": data["x"]})
        if backward:
            if gout is None:
                gout = torch.ones_like(logits)
            (logits * gout).sum().backward()
    return {"out": logits.detach().clone(), "grads": param_grads(model) if backward else {}, "index_calls": rec}


# ------------------------------------------------------------------------------------------------- cfg4
def _train_block(first, last):
    """Statements [first, last] of the reference's examples/segmentation/train.py, dedented, unedited."""
    path = os.path.join(refenv.ref_root(), "examples", "segmentation", "train.py")
    lines = open(path).read().split("\n")
    return textwrap.dedent("\n".join(lines[first - 1:last]))


def cfg4_models(num_classes=17):
    refenv.install(refenv.backend() or "geot")
    from openpoints.models import build_model_from_cfg
    from openpoints.utils import EasyConfig
    cfg = EasyConfig()
    cfg.load(os.path.join(refenv.ref_root(), "cfgs", "tooth_semi", "transformer_finetune_fixmatch_ntm.yaml"), recursive=True)
    cfg.num_classes = num_classes
    cfg.model.segmentor_args.nclasses = num_classes
    cfg.t_predictor.T_args.nclasses = num_classes
    torch.manual_seed(1609)
    model = build_model_from_cfg(cfg.model).cuda()
    T_predictor = build_model_from_cfg(cfg.t_predictor).cuda()
    return cfg, model, T_predictor


def _rot(g, scale):
    a = g.uniform(-scale, scale, 3)
    cx, cy, cz = np.cos(a)
    sx, sy, sz = np.sin(a)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return (Rz @ Ry @ Rx).astype(np.float32)


def cfg4_inputs(N=16000, bl=2, bu=2, num_classes=17):
    """2 labeled + 2 unlabeled clouds; the unlabeled ones in a weak and a strong view (scale + rotation + shift,
    cfgs/tooth_semi/transformer_finetune_fixmatch_ntm.yaml datatransforms.train_s)."""
    g = np.random.default_rng(1609)
    pos_l, lab_l = synth.tooth_batch(bl, N, first_cloud=0, return_labels=True)
    pos_w, lab_w = synth.tooth_batch(bu, N, first_cloud=bl, return_labels=True)
    pos_s = np.stack([(p * g.uniform(0.8, 1.2)) @ _rot(g, 1.0).T + g.uniform(-0.2, 0.2, 3).astype(np.float32) for p in pos_w]).astype(np.float32)
    to = lambda a: _cuda(a)
    data = {"pos": to(pos_l), "x": to(pos_l).transpose(1, 2).contiguous(), "cls": torch.zeros((bl, 1), dtype=torch.long, device="cuda"),
            "y": _cuda(lab_l % num_classes, torch.long)}
    data_u = {"pos_w": to(pos_w), "x_w": to(pos_w).transpose(1, 2).contiguous(), "cls_w": torch.zeros((bu, 1), dtype=torch.long, device="cuda"),
              "pos_s": to(pos_s), "x_s": to(pos_s).transpose(1, 2).contiguous(), "cls_s": torch.zeros((bu, 1), dtype=torch.long, device="cuda"),
              "raw_pos": to(pos_w), "y": _cuda(lab_w % num_classes, torch.long).unsqueeze(-1)}
    return data, data_u


def run_cfg4_model(backend, model, data, data_u, gout=None, backward=True, seed=11):
    """The student forward of train.py:492 (`model(data, u0=data_u, fixmatch=True)`: 6 clouds through Group,
    the transformer, pointops.fps x3, PointnetFPModule x3, DGCNN_Propagation x2) + backward."""
    refenv.use(backend)
    model.train()
    model.zero_grad(set_to_none=True)
    torch.manual_seed(seed)
    C = model.segmentor.nclasses
    du = dict(data_u)
    du["T"] = torch.eye(C, device="cuda")
    with refenv.recording() as rec:
        pred_all, delta_T, sigma = model(data, u0=du, fixmatch=True)
        if backward:
            if gout is None:
                gout = torch.ones_like(pred_all)
            (pred_all * gout).sum().backward()
    return {"out": pred_all.detach().clone(), "grads": param_grads(model) if backward else {}, "index_calls": rec,
            "sigma": sigma.detach().clone()}


def cfg4_transition_inputs(C=17, B=2, N=16000, seed=5):
    """Synthetic student/teacher logits with tooth-like label structure for the transition block."""
    g = np.random.default_rng(seed)
    pos, lab = synth.tooth_batch(B, N, first_cloud=2, return_labels=True)
    lab = lab % C
    onehot = np.eye(C, dtype=np.float32)[lab]                                    # (B, N, C)
    teacher = (onehot * 4.0 + g.standard_normal((B, N, C)).astype(np.float32)).transpose(0, 2, 1)
    student = (onehot * 3.0 + 1.5 * g.standard_normal((B, N, C)).astype(np.float32)).transpose(0, 2, 1)
    cm = F.normalize(_cuda(g.random((C, C)).astype(np.float32)) + 0.01, p=1, dim=1)
    ema_t = F.normalize(_cuda(g.random((C, C)).astype(np.float32)) + torch.eye(C, device="cuda") * 5, p=1, dim=1)
    sigma = torch.full((C,), 0.4, device="cuda") + _cuda(g.random(C).astype(np.float32)) * 0.2
    return {"pred_u_strong": _cuda(student), "pred_u": F.softmax(_cuda(teacher), dim=1), "raw_pos": _cuda(pos),
            "cm": cm, "ema_t": ema_t, "sigma": sigma}


def label_proj(C):
    from geot_b200 import transition as gt
    return gt.LABEL_PROJ_17 if C == 17 else (gt.LABEL_PROJ_33 if C == 33 else list(range(C)))


def run_transition_reference(cfg, T_predictor, inp, threshold=0.0, double=False):
    """train.py:505-557,571,592-593 exec'd verbatim on the GPU (reference torch code; C != 17 replaces LABEL_PROJ
    train.py:48 by the harness-defined projection).  double=True runs the very same statements
    in float64 (inputs, predictor weights and torch's default dtype) - the rounding-free answer both fp32
    implementations are measured against."""
    if double:
        import copy
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.float64)
        try:
            inp64 = {k: (v.double() if v.is_floating_point() else v) for k, v in inp.items()}
            return run_transition_reference(cfg, copy.deepcopy(T_predictor).double(), inp64, threshold, False)
        finally:
            torch.set_default_dtype(prev)
    refenv.install(refenv.backend() or "geot")
    from openpoints.loss import build_criterion_from_cfg
    from openpoints.utils import EasyConfig
    from utils.insT_loss import threeD_space_loss
    C = cfg.num_classes
    pred_u_strong = inp["pred_u_strong"].detach().clone().requires_grad_(True)
    T_predictor.zero_grad(set_to_none=True)
    env = {"torch": torch, "F": F, "cfg": cfg, "pred_u": inp["pred_u"], "sigma": inp["sigma"], "ema_t": inp["ema_t"].clone(),
           "pred_u_strong": pred_u_strong, "T_predictor": T_predictor, "cm": inp["cm"],
           "batch_size_u": pred_u_strong.shape[0], "num_point": pred_u_strong.shape[2]}
    exec(_train_block(835, 836), env)                                           # gaussian()
    exec(_train_block(48, 48), env)                                             # LABEL_PROJ
    if C != 17:
        env["LABEL_PROJ"] = label_proj(C)
    exec(_train_block(505, 557), env)
    logits_u_aug, label_u_aug = torch.max(inp["pred_u"], dim=1)                 # train.py:478
    threed = threeD_space_loss(cfg.threed_k, cfg.threed_sigma, C)               # train.py:270
    l3 = threed(inp["raw_pos"], label_u_aug, env["insT"]) * cfg.threed_loss_weight     # train.py:571
    cu = EasyConfig()
    cu.update({"NAME": "Poly1FocalLoss_U_corr"})
    criterion_u = build_criterion_from_cfg(cu).cuda()
    lu = criterion_u(env["pred_u_strong_corr"], label_u_aug.detach(), logits_u_aug.detach(), thresh=threshold, mask=None)   # :592
    (lu + l3).backward()
    return {"class_T": env["class_T"], "ema_t_corr": env["ema_t_corr"], "ema_t": env["ema_t"], "insT": env["insT"].detach(),
            "corr": env["pred_u_strong_corr"].detach(), "loss_3d": l3.detach(), "loss_u": lu.detach(),
            "g_logits": pred_u_strong.grad.clone(),
            "g_W": torch.stack([T_predictor.T_predictor.fc[k].weight.grad for k in range(C)]).clone()}


def run_transition_geot(cfg, T_predictor_ref, inp, threshold=0.0):
    """The same block on geot_b200.transition / geot_b200.losses (same state_dict as the reference predictor)."""
    from geot_b200 import losses as gl
    from geot_b200 import transition as gt
    C = cfg.num_classes
    tp = gt.Ins_T_mean(T_args={"NAME": "sig_t_mean", "nclasses": C}).cuda()
    tp.load_state_dict(T_predictor_ref.state_dict())
    pred_u_strong = inp["pred_u_strong"].detach().clone().requires_grad_(True)
    class_T, _, _ = gt.class_T_pick(inp["pred_u"])
    ema_t_corr, ema_t = gt.class_T_step(class_T, inp["ema_t"], inp["sigma"], label_proj(C), cfg.geo_lambma, cfg.ema_t_decay)
    insT = tp(F.softmax(pred_u_strong, dim=1).detach(), inp["cm"])
    corr = gt.apply_transition(pred_u_strong, ema_t_corr, insT, cfg.lambma)
    logits_u_aug, label_u_aug = torch.max(inp["pred_u"], dim=1)
    l3 = gt.threeD_space_loss(cfg.threed_k, cfg.threed_sigma, C)(inp["raw_pos"], label_u_aug, insT) * cfg.threed_loss_weight
    lu = gl.Poly1FocalLoss_U_corr()(corr, label_u_aug.detach(), logits_u_aug.detach(), thresh=threshold, mask=None)
    (lu + l3).backward()
    return {"class_T": class_T, "ema_t_corr": ema_t_corr, "ema_t": ema_t, "insT": insT.detach(), "corr": corr.detach(),
            "loss_3d": l3.detach(), "loss_u": lu.detach(), "g_logits": pred_u_strong.grad.clone(),
            "g_W": torch.stack([tp.T_predictor.fc[k].weight.grad for k in range(C)]).clone()}


# ------------------------------------------------------------------------------------------------- cfg2
def cfg2_inputs(B=8, N=16384, C0=32):
    xyz = synth.tooth_batch(B, N).reshape(B * N, 3)
    g = np.random.default_rng(2)
    feat = g.standard_normal((B * N, C0)).astype(np.float32)
    offset = torch.arange(1, B + 1, dtype=torch.int32, device="cuda") * N
    return {"xyz": _cuda(xyz), "feat": _cuda(feat), "offset": offset, "B": B, "N": N}


def cfg2_weights(levels=(32, 64, 128, 256, 512), seed=22):
    """Stand-ins for the transition-down / transition-up linear layers (plain matmuls; library code)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    down = [torch.randn(3 + levels[i], levels[i + 1], device="cuda", generator=g) / np.sqrt(3 + levels[i]) for i in range(len(levels) - 1)]
    up = [torch.randn(levels[i + 1], levels[i], device="cuda", generator=g) / np.sqrt(levels[i + 1]) for i in range(len(levels) - 1)]
    return [w.requires_grad_(True) for w in down], [w.requires_grad_(True) for w in up]


def run_cfg2(backend, inp, weights, k=16, strides=(4, 4, 4, 4), backward=True):
    """per encoder level furthestsampling -> knnquery(16) -> queryandgroup (use_xyz) -> linear
    -> max over k; self-kNN + grouping within each level; decoder interpolation2 back up; loss = sum of outputs."""
    refenv.install(backend)
    import openpoints.cpp.pointops.functions.pointops as po
    down, up = weights
    for w in down + up:
        w.grad = None
    feat0 = inp["feat"].detach().clone().requires_grad_(True)
    xyzs, feats, offs = [inp["xyz"]], [feat0], [inp["offset"]]
    B, n = inp["B"], inp["N"]
    with refenv.recording() as rec:
        for lvl, s in enumerate(strides):
            xyz, feat, off = xyzs[-1], feats[-1], offs[-1]
            n = n // s
            new_off = torch.arange(1, B + 1, dtype=torch.int32, device="cuda") * n
            idx = po.furthestsampling(xyz, off, new_off)                                  # (B*n,)
            new_xyz = xyz[idx.long(), :].contiguous()
            # within-level self kNN + grouping (vector-attention neighbourhood) at the source resolution
            sidx, _ = po.knnquery(k, xyz, xyz, off, off)
            local = po.grouping(feat, sidx).max(1)[0]                                     # (n_src, C)
            featl = (feat + local).contiguous()
            g = po.queryandgroup(k, xyz, new_xyz, featl, None, off, new_off, use_xyz=True)  # (m, k, 3+C)
            new_feat = torch.relu(g @ down[lvl]).max(1)[0].contiguous()                   # (m, C')
            xyzs.append(new_xyz); feats.append(new_feat); offs.append(new_off)
        f = feats[-1]
        for lvl in range(len(strides) - 1, -1, -1):
            f = po.interpolation2(xyzs[lvl + 1], xyzs[lvl], (f @ up[lvl]).contiguous(), offs[lvl + 1], offs[lvl]) + feats[lvl]
        out = f
        if backward:
            # a dense random cotangent: `out.sum().backward()` would hand the op wrappers an EXPANDED (stride-0)
            # gradient, which the reference's backward kernels read as if it were contiguous (no check there;
            # geot_b200 raises) - real losses produce dense gradients
            gout = torch.randn(out.shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(99))
            (out * gout).sum().backward()
    res = {"out": out.detach().clone(), "index_calls": rec}
    if backward:
        res["grads"] = {"feat0": feat0.grad.clone(), **{f"down{i}": w.grad.clone() for i, w in enumerate(down)},
                        **{f"up{i}": w.grad.clone() for i, w in enumerate(up)}}
    return res
