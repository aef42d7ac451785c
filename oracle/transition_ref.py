"""oracle/transition_ref.py — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

CPU (torch, fp32/fp64) restatement of the reference's instance-dependent transition-matrix path.
Every function follows the reference expression it cites; device is whatever the inputs live on
(the reference hard-codes `.cuda()`).  Pinned against the reference's OWN code run on CPU by
tests/golden/make_transition_golden.py (fixtures in tests/golden/transition_*.npz).

Paths are relative to the reference root.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def sig_t_mean_forward(x, cm, W):
    """openpoints/models/backbone/transformer.py:1120-1131 (sig_t_mean.forward).

    x (B,C,N) class probabilities, cm (C,C) class mean features, W (C, C, 2C) = stacked fc[kk].weight
    -> ins_T (B*N, C, C): row kk = L1norm(clamp(W[kk] @ concat(p_n, cm[kk]), 1e-5, 1-1e-5))."""
    C = x.shape[1]
    out = x.permute(0, 2, 1).contiguous().view(-1, C)  # BN C
    ins_T = torch.empty((out.shape[0], C, C), dtype=out.dtype, device=out.device)
    for kk in range(C):
        new_in = torch.cat((out, cm[kk].unsqueeze(0).repeat(out.shape[0], 1)), dim=1)
        ins_T[:, kk, :] = new_in @ W[kk].t()
    ins_T = torch.clamp(ins_T, min=1e-5, max=1 - 1e-5)
    return F.normalize(ins_T, p=1, dim=2)


def sig_t_forward(x, W):
    """transformer.py:1071-1097 (sig_t.forward): one Linear(C -> C*C, bias=False), clamp, L1 rows.
    W (C*C, C)."""
    C = x.shape[1]
    out = x.permute(0, 2, 1).contiguous().view(-1, C)
    out = (out @ W.t()).view(-1, C, C)
    out = torch.clamp(out, min=1e-5, max=1 - 1e-5)
    return F.normalize(out, p=1, dim=2)


def transition_apply(logits, ema_t_corr, ins_T, lam):
    """examples/segmentation/train.py:549-552.

    logits (B,C,N) raw student logits of the strong view, ema_t_corr (C,C), ins_T (BN,C,C)
    -> corrected logits (B,C,N) = permute(bmm(logit_n (1,C), L1norm_rows(lam*ema + (1-lam)*ins_T_n)))."""
    B, C, N = logits.shape
    newT = lam * ema_t_corr + (1 - lam) * ins_T
    newT = F.normalize(newT, p=1, dim=2)
    out = torch.bmm(logits.permute(0, 2, 1).contiguous().view(-1, C).unsqueeze(1), newT).squeeze(1)
    return out.view(B, N, C).permute(0, 2, 1).contiguous()


def knn_point(k, query, support=None):
    """openpoints/models/layers/knn.py:7-20 (cdist + topk, ascending)."""
    if support is None:
        support = query
    dist = torch.cdist(query, support)
    kd = dist.topk(k=k, dim=-1, largest=False, sorted=True)
    return kd.values, kd.indices


def _manifold(points, labels, ins_T, idx, sigma, signed, normalised):
    """Shared body of utils/insT_loss.py:19-58 and :69-110 with the k index_select/cat rounds written
    as one gather.  points (B,N,D); idx (B,N,k) LOCAL neighbour indices (self already dropped)."""
    B, N, k = idx.shape
    glob = (idx.long() + (torch.arange(B, device=idx.device) * N).view(B, 1, 1)).view(B * N, k)
    pts = points.reshape(B * N, -1)
    lab = labels.reshape(-1)
    nb_pts = pts[glob]                      # BN k D
    nb_lab = lab[glob]                      # BN k
    nb_T = ins_T[glob].reshape(B * N, k, -1)
    same = (lab.unsqueeze(1) == nb_lab)
    if signed:   # feature_space_loss: -1 / +1  (insT_loss.py:42-45)
        dmap = torch.where(same, torch.ones((), dtype=pts.dtype, device=pts.device), -torch.ones((), dtype=pts.dtype, device=pts.device))
    else:        # threeD_space_loss: 0 / 1     (insT_loss.py:93-96)
        dmap = same.to(pts.dtype)
    eij = torch.exp((-torch.sum((pts.unsqueeze(1) - nb_pts) ** 2, dim=2)) / (2 * (sigma ** 2)))
    dmap = (dmap * eij).detach()
    T_dist = torch.sum((ins_T.reshape(B * N, 1, -1) - nb_T) ** 2, dim=2)
    if normalised:  # insT_loss.py:107-108
        return (torch.sum(dmap * T_dist, dim=1) / (torch.sum(dmap, dim=1) + 0.001)).mean()
    return torch.mean(dmap * T_dist)  # insT_loss.py:56


def threeD_space_loss(positions, labels, ins_T, k, sigma=1.0, idx=None):
    """utils/insT_loss.py:61-110.  idx (B,N,k): neighbours to use (default: the reference's own
    knn_point(k+1)[..., 1:], whose tie order is unspecified — pass idx to pin it)."""
    if idx is None:
        idx = knn_point(k + 1, positions, positions)[1][:, :, 1:]
    return _manifold(positions, labels, ins_T, idx, sigma, signed=False, normalised=True)


def feature_space_loss(logits, labels, ins_T, k, sigma=1.0, idx=None):
    """utils/insT_loss.py:9-58.  logits (B,C,N)."""
    feats = logits.permute(0, 2, 1).contiguous()
    if idx is None:
        idx = knn_point(k + 1, feats, feats)[1][:, :, 1:]
    return _manifold(feats, labels, ins_T, idx, sigma, signed=True, normalised=False)


def identity_loss(insT, Identity):
    """utils/insT_loss.py:113-132 (Idenyity_loss)."""
    num = insT.size(0)
    Id = Identity.repeat(num, 1, 1).view(num, -1)
    diff = (insT.view(num, -1) - Id).pow(2)
    return (torch.sum(diff * Id, dim=1) / torch.sum(Id, dim=1)).mean()


def gaussian(x, mu, s):
    """train.py:835-836."""
    return (1 / (s * torch.sqrt(torch.tensor(2 * torch.pi)))) * torch.exp(-((x - mu) ** 2) / (2 * s ** 2))


# train.py:48 (17 classes).  For C=33 the harness-defined extension of SURVEY.md Appendix D.
LABEL_PROJ_17 = [0, 8, 7, 6, 5, 4, 3, 2, 1, 9, 10, 11, 12, 13, 14, 15, 16]
LABEL_PROJ_33 = LABEL_PROJ_17 + [24, 23, 22, 21, 20, 19, 18, 17, 25, 26, 27, 28, 29, 30, 31, 32]


def class_T_pick(pred_u):
    """train.py:509-526 without filter_outlier: row cc of class_T = the softmax vector of the single point
    (over the whole unlabeled batch) with the largest probability for class cc; torch.argmax keeps the
    FIRST maximum in (b, n) order.  pred_u (B,C,N) softmax -> class_T (C,C), best value (C), flat index (C)."""
    B, C, N = pred_u.shape
    class_T = torch.empty((C, C), dtype=pred_u.dtype, device=pred_u.device)
    best_v = torch.empty((C,), dtype=pred_u.dtype, device=pred_u.device)
    best_i = torch.empty((C,), dtype=torch.int64, device=pred_u.device)
    for cc in range(C):
        robust = pred_u[:, cc, :].contiguous().view(B * N)
        ib = torch.argmax(robust)
        class_T[cc] = pred_u[ib // N, :, ib % N]
        best_v[cc] = robust[ib]
        best_i[cc] = ib
    return class_T, best_v, best_i


def class_T_step(class_T, ema_t, sigma, label_proj, geo_lambma, ema_t_decay):
    """train.py:507,528-545,556-557: Gaussian geometric prior, blend, the reference's keepdim-less
    'normalisation' (X / X.sum(1) divides COLUMN j by ROW-sum j — copied verbatim, Appendix D), EMA.
    -> (ema_t_corr used for this step's correction, new ema_t)."""
    C = class_T.shape[0]
    prior_T = torch.zeros((C, C), dtype=class_T.dtype, device=class_T.device)
    for cc in range(1, C):
        for ck in range(C):
            prior_T[cc, ck] = gaussian(label_proj[ck], label_proj[cc], sigma[cc])
    prior_T[:, 0] = 0
    prior_T[0, 0] = 1
    prior_T = prior_T / torch.sum(prior_T, 1)
    new_T = geo_lambma * class_T + (1 - geo_lambma) * prior_T
    new_T[0] = class_T[0]
    new_T = new_T / torch.sum(new_T, 1)
    ema_t_corr = ema_t * ema_t_decay + new_T * (1 - ema_t_decay)
    ema_t_corr = ema_t_corr / torch.sum(ema_t_corr, 1)
    ema_new = ema_t * ema_t_decay + class_T * (1 - ema_t_decay)
    ema_new = ema_new / torch.sum(ema_new, 1)
    return ema_t_corr, ema_new


def mean_feature_update(cm, c_num, probs, target):
    """train.py:884-895 (cal_mean_feature inner loop), bug-for-bug: `probs[target]` indexes ROWS BY CLASS ID.
    probs (BN,C) softmax, target (BN) int64; cm (C,C), c_num (C) updated in place and returned."""
    C = cm.shape[0]
    for kk in range(C):
        cur_num = torch.sum(target == kk)
        if cur_num == 0:
            continue
        mean_feats = probs[target].mean(0)
        cm[kk] = (cm[kk] * c_num[kk] + mean_feats * cur_num) / (c_num[kk] + cur_num)
        c_num[kk] += cur_num
    return cm, c_num


def poly1_focal_u_corr(logits, labels, logits_pred, thresh=0.95, mask=None, epsilon=1.0, alpha=0.25, gamma=2.0):
    """openpoints/loss/build.py:830-892 (Poly1FocalLoss_U_corr.forward; weight = pos_weight = None as in the
    shipped configs).  logits (B,C,N) corrected student logits, labels (B,N) pseudo labels, logits_pred (B,N)
    teacher confidence -> scalar: sigmoid-focal + poly1 term, masked mean over points with confidence >= thresh
    (or an explicit (B,N) mask), denominator C * #kept + 0.001.
    Pinned against the reference class by tests/golden/make_loss_golden.py."""
    C = logits.shape[1]
    p = torch.sigmoid(logits)
    onehot = F.one_hot(labels.unsqueeze(1), C).transpose(1, -1).squeeze(-1).to(dtype=logits.dtype)
    ce = F.binary_cross_entropy_with_logits(logits, onehot, reduction="none")
    pt = onehot * p + (1 - onehot) * (1 - p)
    FL = ce * ((1 - pt) ** gamma)
    if alpha >= 0:
        FL = (alpha * onehot + (1 - alpha) * (1 - onehot)) * FL
    poly1 = FL + epsilon * torch.pow(1 - pt, gamma + 1)
    keep = mask if mask is not None else logits_pred.ge(torch.tensor(thresh, device=logits_pred.device))
    keep = keep.unsqueeze(1).repeat(1, C, 1)
    return torch.sum(poly1 * keep) / (keep.sum() + 0.001)


def poly1_focal_u_t(logits, labels, logits_pred, pred_u_t, thresh=0.95, mask=None, epsilon=1.0, alpha=0.25, gamma=2.0):
    """openpoints/loss/build.py:595-688 (Poly1FocalLoss_U_T.forward; weight = pos_weight = None; T is accepted and unused
    there).  As poly1_focal_u_corr on the UNcorrected logits with every point scaled by
    beta = logits_pred / pred_u_t[b, label, n] (:661-670); beta stays in the autograd graph.
    Pinned against the reference class by tests/golden/make_loss_golden.py (loss_t_golden.npz)."""
    C = logits.shape[1]
    p = torch.sigmoid(logits)
    onehot = F.one_hot(labels.unsqueeze(1), C).transpose(1, -1).squeeze(-1).to(dtype=logits.dtype)
    ce = F.binary_cross_entropy_with_logits(logits, onehot, reduction="none")
    pt = onehot * p + (1 - onehot) * (1 - p)
    FL = ce * ((1 - pt) ** gamma)
    if alpha >= 0:
        FL = (alpha * onehot + (1 - alpha) * (1 - onehot)) * FL
    poly1 = FL + epsilon * torch.pow(1 - pt, gamma + 1)
    after = torch.gather(pred_u_t, 1, labels.unsqueeze(1)).squeeze(1)  # pred_u_t[b, label[b, n], n]
    poly1 = poly1 * (logits_pred / after).unsqueeze(1)
    keep = mask if mask is not None else logits_pred.ge(torch.tensor(thresh, device=logits_pred.device))
    keep = keep.unsqueeze(1).repeat(1, C, 1)
    return torch.sum(poly1 * keep) / (keep.sum() + 0.001)


def poly1_focal_u_t_v1(logits, labels, logits_pred, T, pred_u, delta_T, thresh=0.95, mask=None, epsilon=1.0, alpha=0.25, gamma=2.0):
    """openpoints/loss/build.py:690-795 (Poly1FocalLoss_U_T_v1.forward): the corrected prediction is formed inside the
    loss from the class-level matrix, pred_u_t[b,:,n] = (T + delta_T)^T . pred_u[b,:,n] (build.py:766-770), then
    beta = logits_pred / pred_u_t[b, label, n] scales the poly-1 focal terms as in Poly1FocalLoss_U_T."""
    pred_u_t = torch.einsum("ij,bin->bjn", T + delta_T, pred_u)
    return poly1_focal_u_t(logits, labels, logits_pred, pred_u_t, thresh=thresh, mask=mask, epsilon=epsilon, alpha=alpha, gamma=gamma)
