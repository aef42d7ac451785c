"""geot_b200.losses — the criteria that consume the transition-corrected logits.

`Poly1FocalLoss_U_corr` mirrors openpoints/loss/build.py:799-892 (constructor arguments, `forward(logits, labels,
logits_pred, thresh=0.95, mask=None)`); the 8-10 elementwise (B,C,N) passes of the reference and their saved
activations are one CUDA pass forward and one backward (geot_b200/csrc/unsup_loss.cu).  CUDA only."""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.autograd import Function

from . import functional as F_


class _Poly1(Function):
    @staticmethod
    def forward(ctx, logits, labels, conf, mask, thresh, epsilon, alpha, gamma):
        logits = logits.contiguous().float()
        labels = labels.contiguous().long()
        conf = None if conf is None else conf.contiguous().float()
        mask = None if mask is None else mask.contiguous().float()
        loss, denom = F_.poly1_focal_forward(logits, labels, conf, mask, thresh, epsilon, alpha, gamma)
        ctx.save_for_backward(logits, labels, conf, mask, denom)
        ctx.hp = (float(thresh), float(epsilon), float(alpha), float(gamma))
        return loss.view(())

    @staticmethod
    def backward(ctx, grad_loss):
        logits, labels, conf, mask, denom = ctx.saved_tensors
        thresh, epsilon, alpha, gamma = ctx.hp
        g = F_.poly1_focal_backward(logits, labels, conf, mask, thresh, epsilon, alpha, gamma, denom,
                                    grad_loss.contiguous().float().view(1))
        return g, None, None, None, None, None, None, None


class Poly1FocalLoss_U_corr(nn.Module):
    """openpoints/loss/build.py:799-892.  logits (B,C,N) (the corrected student logits, train.py:592), labels (B,N)
    pseudo labels, logits_pred (B,N) teacher confidence; points with confidence >= thresh (or mask != 0) are kept.
    `weight` / `pos_weight` / `label_is_onehot` / `reduction` exist for signature parity; the shipped configs leave
    them at their defaults and anything else raises."""

    def __init__(self, epsilon: float = 1.0, alpha: float = 0.25, gamma: float = 2.0, reduction: str = "mean",
                 weight=None, pos_weight=None, label_is_onehot: bool = False, **kwargs):
        super().__init__()
        if weight is not None or pos_weight is not None or label_is_onehot:
            raise NotImplementedError("Poly1FocalLoss_U_corr: weight / pos_weight / one-hot labels are not used by the "
                                      "reference configs and are not implemented by the CUDA path")
        self.epsilon, self.alpha, self.gamma, self.reduction = epsilon, alpha, gamma, reduction

    def forward(self, logits, labels, logits_pred, thresh=0.95, mask=None):
        if logits.dim() != 3:
            raise ValueError("Poly1FocalLoss_U_corr expects (B, C, N) logits")
        return _Poly1.apply(logits, labels, None if mask is not None else logits_pred, mask, thresh,
                            self.epsilon, self.alpha, self.gamma)


class _Poly1T(Function):
    @staticmethod
    def forward(ctx, logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma):
        logits = logits.contiguous().float()
        labels = labels.contiguous().long()
        conf = conf.contiguous().float()
        mask = None if mask is None else mask.contiguous().float()
        pred_t = pred_t.contiguous().float()
        loss, denom = F_.poly1_focal_t_forward(logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma)
        ctx.save_for_backward(logits, labels, conf, mask, pred_t, denom)
        ctx.hp = (float(thresh), float(epsilon), float(alpha), float(gamma))
        return loss.view(())

    @staticmethod
    def backward(ctx, grad_loss):
        logits, labels, conf, mask, pred_t, denom = ctx.saved_tensors
        thresh, epsilon, alpha, gamma = ctx.hp
        g, gt = F_.poly1_focal_t_backward(logits, labels, conf, mask, pred_t, thresh, epsilon, alpha, gamma, denom,
                                          grad_loss.contiguous().float().view(1))
        return g, None, None, None, gt, None, None, None, None


class Poly1FocalLoss_U_T(nn.Module):
    """openpoints/loss/build.py:564-688 (selected at train.py:594-596).  logits (B,C,N) = the UNcorrected student
    logits, labels (B,N), logits_pred (B,N) teacher confidence, T (C,C) class-level matrix (accepted and unused, as in the
    reference's forward), pred_u_t (B,C,N) = the transition-corrected prediction: every point's poly-1 focal terms are
    scaled by beta = logits_pred / pred_u_t[b, label, n].  Gradients reach `logits` and `pred_u_t` (the reference keeps
    beta in the graph; logits_pred arrives detached, train.py:595, and gets no gradient here)."""

    def __init__(self, epsilon: float = 1.0, alpha: float = 0.25, gamma: float = 2.0, reduction: str = "mean",
                 weight=None, pos_weight=None, label_is_onehot: bool = False, **kwargs):
        super().__init__()
        if weight is not None or pos_weight is not None or label_is_onehot:
            raise NotImplementedError("Poly1FocalLoss_U_T: weight / pos_weight / one-hot labels are not used by the "
                                      "reference configs and are not implemented by the CUDA path")
        self.epsilon, self.alpha, self.gamma, self.reduction = epsilon, alpha, gamma, reduction

    def forward(self, logits, labels, logits_pred, T, pred_u_t, thresh=0.95, mask=None):
        if logits.dim() != 3 or pred_u_t.shape != logits.shape:
            raise ValueError("Poly1FocalLoss_U_T expects (B, C, N) logits and pred_u_t of the same shape")
        return _Poly1T.apply(logits, labels, logits_pred, mask, pred_u_t, thresh, self.epsilon, self.alpha, self.gamma)


class Poly1FocalLoss_U_T_v1(nn.Module):
    """openpoints/loss/build.py:690-795: the class-level variant.  Instead of receiving the corrected prediction it forms
    it itself, `pred_u_t = (T + delta_T)^T . pred_u` (the matmul of build.py:766-770), takes beta = logits_pred /
    pred_u_t[b, label, n] and returns `(loss, delta_T)`.  The (C x C) . (C x BN) product is a dense contraction - a plain
    library GEMM, in the reference's own statement order so that pred_u_t has the same bits; everything after it (the
    8-10 elementwise (B,C,N) passes and the masked mean) is the fused kernel pair of Poly1FocalLoss_U_T.  Gradients reach
    `logits`, `pred_u` and - through the GEMM - `delta_T` (the learnable class-level revision) and `T`."""

    def __init__(self, epsilon: float = 1.0, alpha: float = 0.25, gamma: float = 2.0, reduction: str = "mean",
                 weight=None, pos_weight=None, label_is_onehot: bool = False, **kwargs):
        super().__init__()
        if weight is not None or pos_weight is not None or label_is_onehot:
            raise NotImplementedError("Poly1FocalLoss_U_T_v1: weight / pos_weight / one-hot labels are not used by the "
                                      "reference configs and are not implemented by the CUDA path")
        self.epsilon, self.alpha, self.gamma, self.reduction = epsilon, alpha, gamma, reduction

    def forward(self, logits, labels, logits_pred, T, pred_u, delta_T, thresh=0.95, mask=None):
        if logits.dim() != 3 or pred_u.shape != logits.shape:
            raise ValueError("Poly1FocalLoss_U_T_v1 expects (B, C, N) logits and pred_u of the same shape")
        batch_size, num_point = labels.shape
        ema_t_corr = T + delta_T                                                           # build.py:766
        pred_u_t = pred_u.transpose(1, 0).contiguous().view(-1, batch_size * num_point)    # :767
        pred_u_t = torch.matmul(ema_t_corr.t(), pred_u_t)                                  # :768
        pred_u_t = pred_u_t.view(-1, batch_size, num_point).transpose(1, 0).contiguous()   # :769  (B, C, N)
        loss = _Poly1T.apply(logits, labels, logits_pred, mask, pred_u_t, thresh, self.epsilon, self.alpha, self.gamma)
        return loss, delta_T
