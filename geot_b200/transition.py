"""Mirror of the reference's instance-dependent transition-matrix path on the sm_100a kernels.

Same class names, constructor arguments, parameter names (state_dict keys) and call signatures as the
reference, so its training loop can swap these in:
  sig_t_mean, sig_t        openpoints/models/backbone/transformer.py:1099-1131, :1071-1097
  Ins_T_mean               openpoints/models/segmentation/base_seg.py:254-263
  threeD_space_loss,
  feature_space_loss,
  Idenyity_loss            utils/insT_loss.py:61-110, :9-58, :113-132
  apply_transition         examples/segmentation/train.py:549-552   (inline code in the reference)
  class_T_pick / class_T_step / ClassTransitionState   train.py:505-545,556-557 (inline)
  mean_feature_update      train.py:884-895 (cal_mean_feature inner loop)
No CPU path: CPU tensors raise (geot_b200._lib.GeotError).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
from torch.autograd import Function

from . import functional as F_
from . import layers

LABEL_PROJ_17 = [0, 8, 7, 6, 5, 4, 3, 2, 1, 9, 10, 11, 12, 13, 14, 15, 16]  # train.py:48
# 33 classes (BASELINE config 4; harness-defined): lower jaw as the reference, upper jaw mirrored
LABEL_PROJ_33 = LABEL_PROJ_17 + [24, 23, 22, 21, 20, 19, 18, 17, 25, 26, 27, 28, 29, 30, 31, 32]


class _SigT(Function):
    @staticmethod
    def forward(ctx, x, W, cm):
        x = x.contiguous().float()
        Wc = W.contiguous()
        ctx.save_for_backward(x, Wc, cm)
        return F_.sigt_forward(x, Wc, cm)

    @staticmethod
    def backward(ctx, grad_insT):
        x, W, cm = ctx.saved_tensors
        grad_W, g_raw = F_.sigt_backward(x, W, cm, grad_insT.contiguous())
        grad_x = None
        if ctx.needs_input_grad[0]:
            # not on the reference's path (train.py:548 feeds a detached softmax): a plain library GEMM
            # over the kernel's g_raw is enough.  grad_p = g_raw (BN, C*C) @ A (C*C, C).
            B, C, N = x.shape
            A = W.reshape(C * C, -1)[:, :C]
            grad_x = (g_raw @ A).view(B, N, C).permute(0, 2, 1).contiguous()
        return grad_x, grad_W, None


class sig_t_mean(nn.Module):
    """transformer.py:1099-1131.  Parameters: fc.{kk}.weight (C, 2C), kk < C — identical state_dict."""

    def __init__(self, nclasses):
        super().__init__()
        self.nclasses = nclasses
        self.fc = nn.ModuleList()
        for _ in range(self.nclasses):
            self.fc.append(nn.Linear(nclasses * 2, nclasses, bias=False))

    def forward(self, x, cm):
        W = torch.stack([l.weight for l in self.fc])  # (C, C, 2C); autograd splits grad_W back per Linear
        return _SigT.apply(x, W, cm.contiguous().float())


class sig_t(nn.Module):
    """transformer.py:1071-1097.  Parameter: fc.weight (C*C, C)."""

    def __init__(self, nclasses):
        super().__init__()
        self.nclasses = nclasses
        self.fc = nn.Linear(nclasses, nclasses * nclasses, bias=False)

    def forward(self, x):
        return _SigT.apply(x, self.fc.weight, None)


class Ins_T_mean(nn.Module):
    """base_seg.py:254-263; T_args = {'NAME': 'sig_t_mean', 'nclasses': C} (cfg.t_predictor.T_args)."""

    def __init__(self, T_args=None, **kwargs):
        super().__init__()
        args = dict(T_args or {})
        name = args.pop("NAME", "sig_t_mean")
        assert name == "sig_t_mean", name
        self.T_predictor = sig_t_mean(**args)

    def forward(self, clean, cm):
        return self.T_predictor(clean, cm)


class _ApplyTransition(Function):
    @staticmethod
    def forward(ctx, logits, ema_t_corr, insT, lam):
        logits = logits.contiguous().float()
        ema = ema_t_corr.contiguous().float()
        insT = insT.contiguous()
        ctx.save_for_backward(logits, ema, insT)
        ctx.lam = float(lam)
        return F_.transition_apply_forward(logits, ema, insT, lam)

    @staticmethod
    def backward(ctx, grad_out):
        logits, ema, insT = ctx.saved_tensors
        gl, gi = F_.transition_apply_backward(logits, ema, insT, ctx.lam, grad_out.contiguous(),
                                              need_logits=ctx.needs_input_grad[0], need_insT=ctx.needs_input_grad[2])
        return gl, None, gi, None


def apply_transition(pred_u_strong, ema_t_corr, insT, lam):
    """train.py:549-552 in one pass: (B,C,N) logits . L1norm_rows(lam*ema_t_corr + (1-lam)*insT) -> (B,C,N).
    Differentiable w.r.t. the logits and insT (ema_t_corr is a statistic, no gradient — as in the reference,
    where it is built from detached tensors)."""
    return _ApplyTransition.apply(pred_u_strong, ema_t_corr, insT, lam)


class _Manifold(Function):
    @staticmethod
    def forward(ctx, insT, points, labels, idx, sigma, mode):
        insT = insT.contiguous()
        # points are visited cell by cell of a Morton grid: a neighbourhood's (C,C) matrices stay in L2 (the (BN,C,C)
        # tensor does not fit it); a scheduling order only, results are independent of it
        order = F_.spatial_order(points)
        loss, coef = F_.manifold_loss_forward(points, labels, insT, idx, sigma, mode, order=order)
        ctx.save_for_backward(insT, idx, coef, order)
        return loss.view(())

    @staticmethod
    def backward(ctx, grad_loss):
        insT, idx, coef, order = ctx.saved_tensors
        B, N, k = idx.shape
        plan = F_.ScatterPlan(idx, B, N)
        g = F_.manifold_loss_backward(insT, idx, coef, plan, grad_loss.contiguous().float().view(1), order=order)
        return g, None, None, None, None, None


def _neighbours(k, feats):
    """k+1 nearest, first column (self) dropped — insT_loss.py:19-21 / :69-70."""
    _, idx = layers.knn_point(k + 1, feats, feats)
    return idx[:, :, 1:].int().contiguous()


class threeD_space_loss(nn.Module):
    """utils/insT_loss.py:61-110.  forward(positions (B,N,3), labels (B,N), ins_T (BN,C,C)) -> scalar."""

    def __init__(self, k=7, sigma=1.0, num_classes=17):
        super().__init__()
        self.k, self.sigma, self.num_classes = k, sigma, num_classes

    def forward(self, positions, labels, ins_T, idx=None):
        positions = positions.contiguous().float()
        if idx is None:
            idx = _neighbours(self.k, positions)
        return _Manifold.apply(ins_T, positions, labels.contiguous().long(), idx.contiguous(), self.sigma, F_.MANIFOLD_3D)


class feature_space_loss(nn.Module):
    """utils/insT_loss.py:9-58.  forward(logits (B,C,N), labels (B,N), ins_T (BN,C,C)) -> scalar."""

    def __init__(self, k=7, sigma=1.0, num_classes=17):
        super().__init__()
        self.k, self.sigma, self.num_classes = k, sigma, num_classes

    def forward(self, logits, labels, ins_T, idx=None):
        feats = logits.permute(0, 2, 1).contiguous().float()
        if idx is None:
            idx = _neighbours(self.k, feats)  # C-dim kNN: the reference's own torch expression (knn.py:18-20)
        return _Manifold.apply(ins_T, feats, labels.contiguous().long(), idx.contiguous(), self.sigma, F_.MANIFOLD_FEATURE)


class Idenyity_loss(nn.Module):
    """utils/insT_loss.py:113-132 (disabled in the shipped config; a (BN,C*C) elementwise torch expression)."""

    def forward(self, insT, Identity):
        num = insT.size(0)
        Id = Identity.reshape(1, -1)
        diff = (insT.view(num, -1) - Id).pow(2)
        return (torch.sum(diff * Id, dim=1) / torch.sum(Id)).mean()


# --------------------------------------------------------------------------------------------------
# class-level statistics
# --------------------------------------------------------------------------------------------------
@torch.no_grad()
def class_T_pick(pred_u):
    """train.py:509-526 (filter_outlier off): -> class_T (C,C), best_val (C), best_idx (C) int64."""
    return F_.class_pick(pred_u.contiguous().float())


@torch.no_grad()
def class_T_step(class_T, ema_t, sigma, label_proj, geo_lambma, ema_t_decay):
    """train.py:528-545,556-557 -> (ema_t_corr, new ema_t)."""
    if not torch.is_tensor(label_proj):
        label_proj = torch.tensor(label_proj, dtype=torch.int32, device=class_T.device)
    return F_.class_t_step(class_T.contiguous(), ema_t.contiguous(), sigma.detach().contiguous().float(),
                           label_proj.int().contiguous(), geo_lambma, ema_t_decay)


@torch.no_grad()
def mean_feature_update(cm, c_num, probs, target):
    """train.py:884-895 in place; probs (BN,C) softmax rows, target (BN)."""
    return F_.mean_feature_update(probs.contiguous().float(), target.contiguous().long(), cm, c_num)


class ClassTransitionState:
    """The per-process statistics of train.py:274,505-557 (`ema_t`) with the multi-GPU semantics of
    SURVEY.md §8e: each rank picks its per-class anchor on its shard, ranks all-gather (value, row) and keep
    the arg-max (ties: lowest rank), then every rank applies the same deterministic update."""

    def __init__(self, num_classes, label_proj=None, geo_lambma=0.999, ema_t_decay=0.999, device="cuda"):
        C = num_classes
        self.C = C
        if label_proj is None:
            label_proj = LABEL_PROJ_17 if C == 17 else (LABEL_PROJ_33 if C == 33 else list(range(C)))
        self.label_proj = torch.tensor(label_proj, dtype=torch.int32, device=device)
        self.geo_lambma, self.ema_t_decay = geo_lambma, ema_t_decay
        self.ema_t = torch.eye(C, device=device)  # train.py:274

    @torch.no_grad()
    def step(self, pred_u, sigma, group=None):
        """pred_u (B,C,N) teacher softmax of THIS rank's shard; sigma (C).  -> ema_t_corr (C,C)."""
        class_T, best_v, _ = class_T_pick(pred_u)
        from . import dist as gdist
        class_T = gdist.combine_class_T(class_T, best_v, group)
        corr, self.ema_t = class_T_step(class_T, self.ema_t, sigma, self.label_proj, self.geo_lambma, self.ema_t_decay)
        return corr
