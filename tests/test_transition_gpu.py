"""Transition-matrix path on the GPU (through the C ABI) against (1) the golden fixtures produced by the
reference's own Python and (2) oracle/transition_ref.py on seeded inputs at larger sizes.
Tolerance: 1e-5 relative (north_star, fp32), written per assert."""
import os

import numpy as np
import pytest
import torch

from geot_b200 import functional as F_
from geot_b200 import synth, transition as T
from oracle import transition_ref as R

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "transition_golden.npz"))


def t(name, dev="cuda"):
    return torch.from_numpy(G[name]).to(dev)


def close(a, b, rtol=1e-5, atol=1e-7):
    torch.testing.assert_close(a.cpu(), b.cpu(), rtol=rtol, atol=atol)


def rel(a, b):
    """relative error in the max norm — the north_star's '1e-5 relative' for gradient tensors."""
    return float((a.double().cpu() - b.double().cpu()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
def test_golden_sig_t_mean_and_apply(cuda, tag):
    C = G[f"{tag}_cm"].shape[0]
    mod = T.sig_t_mean(C).cuda()
    with torch.no_grad():
        for kk in range(C):
            mod.fc[kk].weight.copy_(t(f"{tag}_W")[kk])
    x = t(f"{tag}_x").requires_grad_(True)
    insT = mod(x, t(f"{tag}_cm"))
    close(insT, t(f"{tag}_insT"), rtol=1e-5, atol=1e-7)  # entries just above the 1e-5 clamp come from cancelling O(0.1) terms
    insT.backward(t(f"{tag}_go"))
    gW = torch.stack([mod.fc[kk].weight.grad for kk in range(C)])
    assert rel(gW, t(f"{tag}_gW")) < 1e-5
    assert rel(x.grad, t(f"{tag}_gx")) < 1e-5
    m2 = T.sig_t(C).cuda()
    with torch.no_grad():
        m2.fc.weight.copy_(t(f"{tag}_sigt_W"))
    close(m2(t(f"{tag}_x")), t(f"{tag}_sigt_T"), rtol=1e-5, atol=1e-7)
    # train.py:549-552
    logits = t(f"{tag}_logits").requires_grad_(True)
    ins = t(f"{tag}_insT").requires_grad_(True)
    out = T.apply_transition(logits, t(f"{tag}_ema_t_corr"), ins, 0.9)
    close(out, t(f"{tag}_corr"), rtol=1e-5, atol=1e-6)
    out.backward(t(f"{tag}_go2"))
    assert rel(logits.grad, t(f"{tag}_g_logits")) < 1e-5
    assert rel(ins.grad, t(f"{tag}_g_insT_apply")) < 1e-5


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
def test_golden_manifold_losses(cuda, tag):
    k = int(G[f"{tag}_k"])
    pos, labels = t(f"{tag}_pos"), t(f"{tag}_labels")
    ins = t(f"{tag}_insT").requires_grad_(True)
    l3 = T.threeD_space_loss(k=k, sigma=0.5)(pos, labels, ins, idx=t(f"{tag}_idx3").int().contiguous())
    close(l3, t(f"{tag}_loss3"), rtol=1e-5, atol=1e-9)
    l3.backward()
    assert rel(ins.grad, t(f"{tag}_g_insT_loss3")) < 1e-5
    # own kNN (sm_100a kernel) must find the same neighbours on this tie-free input
    ins.grad = None
    l3b = T.threeD_space_loss(k=k, sigma=0.5)(pos, labels, ins)
    close(l3b, t(f"{tag}_loss3"), rtol=1e-5, atol=1e-9)
    ins.grad = None
    lf = T.feature_space_loss(k=k, sigma=2.0)(t(f"{tag}_logits"), labels, ins, idx=t(f"{tag}_idxf").int().contiguous())
    close(lf, t(f"{tag}_lossf"), rtol=1e-5, atol=1e-9)
    lf.backward()
    assert rel(ins.grad, t(f"{tag}_g_insT_lossf")) < 1e-5
    C = ins.shape[1]
    close(T.Idenyity_loss()(ins.detach(), torch.eye(C, device="cuda").unsqueeze(0)), t(f"{tag}_loss_id"), rtol=1e-5, atol=1e-9)


def test_golden_class_statistics(cuda):
    tag = "C17"
    class_T, best_v, best_i = T.class_T_pick(t(f"{tag}_pred_u"))
    assert torch.equal(class_T.cpu(), t(f"{tag}_class_T", "cpu"))  # a row copy: bit-exact
    corr, ema = T.class_T_step(class_T, t(f"{tag}_ema_t_in"), t(f"{tag}_sigma"), T.LABEL_PROJ_17, 0.999, 0.999)
    close(corr, t(f"{tag}_ema_t_corr_out"), rtol=1e-5, atol=1e-8)
    close(ema, t(f"{tag}_ema_t_out"), rtol=1e-5, atol=1e-8)
    C = 17
    cm, cn = torch.zeros((C, C), device="cuda"), torch.zeros((C,), device="cuda")
    for _ in range(2):
        T.mean_feature_update(cm, cn, t(f"{tag}_mf_probs"), t(f"{tag}_mf_target"))
    close(cm, t(f"{tag}_mf_cm"), rtol=1e-5, atol=1e-8)
    assert torch.equal(cn.cpu(), t(f"{tag}_mf_cnum", "cpu"))


@pytest.mark.parametrize("C,B,N", [(17, 2, 16000), (33, 2, 4096), (33, 3, 1000), (33, 2, 1003), (17, 3, 7), (8, 2, 333), (64, 1, 130)])
def test_oracle_sigt_apply_large(cuda, C, B, N):
    g = torch.Generator().manual_seed(C * 1000 + N)
    x = torch.softmax(torch.randn(B, C, N, generator=g) * 3, 1)
    W = (torch.rand(C, C, 2 * C, generator=g) - 0.5) * (2 / (2 * C) ** 0.5)
    W[:, :, :C] += torch.eye(C) * 0.5  # some rows saturate the clamp on both sides
    cm = torch.softmax(torch.randn(C, C, generator=g), 1)
    go = torch.randn(B * N, C, C, generator=g)
    Wr = W.clone().requires_grad_(True)
    ref = R.sig_t_mean_forward(x, cm, Wr)
    gWr, = torch.autograd.grad(ref, Wr, go)
    Wc = W.cuda().requires_grad_(True)
    out = T._SigT.apply(x.cuda(), Wc, cm.cuda())
    close(out, ref, rtol=1e-5, atol=1e-7)
    out.backward(go.cuda())
    assert rel(Wc.grad, gWr) < 1e-5
    # determinism of the two-stage reduction
    Wc.grad = None
    T._SigT.apply(x.cuda(), Wc, cm.cuda()).backward(go.cuda())
    g1 = Wc.grad.clone(); Wc.grad = None
    T._SigT.apply(x.cuda(), Wc, cm.cuda()).backward(go.cuda())
    assert torch.equal(g1, Wc.grad)
    # apply
    logits = (torch.randn(B, C, N, generator=g) * 2).requires_grad_(True)
    ema = torch.softmax(torch.randn(C, C, generator=g) + torch.eye(C) * 4, 1)
    insT = ref.detach().clone().requires_grad_(True)
    r2 = R.transition_apply(logits, ema, insT, 0.9)
    go2 = torch.randn(B, C, N, generator=g)
    gl, gi = torch.autograd.grad(r2, [logits, insT], go2)
    lc = logits.detach().cuda().requires_grad_(True)
    ic = insT.detach().cuda().requires_grad_(True)
    o2 = T.apply_transition(lc, ema.cuda(), ic, 0.9)
    close(o2, r2, rtol=1e-5, atol=1e-6)
    o2.backward(go2.cuda())
    assert rel(lc.grad, gl) < 1e-5 and rel(ic.grad, gi) < 1e-5


@pytest.mark.parametrize("C,B,N,k", [(17, 2, 3000, 32), (33, 2, 1500, 32), (5, 3, 700, 40), (17, 1, 500, 7)])
def test_oracle_manifold_large(cuda, C, B, N, k):
    g = torch.Generator().manual_seed(C + N + k)
    pos = torch.from_numpy(synth.tooth_batch(B, N))
    labels = torch.randint(0, 4, (B, N), generator=g)
    insT = torch.softmax(torch.randn(B * N, C, C, generator=g), 2).requires_grad_(True)
    idx = R.knn_point(k + 1, pos, pos)[1][:, :, 1:]
    l = R.threeD_space_loss(pos, labels, insT, k, sigma=0.05, idx=idx)
    gr, = torch.autograd.grad(l, insT)
    ic = insT.detach().cuda().requires_grad_(True)
    lc = T.threeD_space_loss(k=k, sigma=0.05)(pos.cuda(), labels.cuda(), ic)  # own kNN
    close(lc, l, rtol=1e-5, atol=1e-10)
    lc.backward()
    assert rel(ic.grad, gr) < 1e-5
    g1 = ic.grad.clone(); ic.grad = None
    T.threeD_space_loss(k=k, sigma=0.05)(pos.cuda(), labels.cuda(), ic).backward()
    assert torch.equal(g1, ic.grad)  # deterministic
    logits = torch.randn(B, C, N, generator=g)
    idxf = R.knn_point(k + 1, logits.permute(0, 2, 1).contiguous())[1][:, :, 1:]
    lf = R.feature_space_loss(logits, labels, insT, k, sigma=1.5, idx=idxf)
    gf, = torch.autograd.grad(lf, insT)
    ic.grad = None
    lfc = T.feature_space_loss(k=k, sigma=1.5)(logits.cuda(), labels.cuda(), ic, idx=idxf.int().cuda().contiguous())
    close(lfc, lf, rtol=2e-5, atol=1e-9)  # signed sum: cancellation
    lfc.backward()
    assert rel(ic.grad, gf) < 1e-5


def test_oracle_class_pick_large(cuda):
    g = torch.Generator().manual_seed(7)
    B, C, N = 4, 33, 16000
    pred = torch.softmax(torch.randn(B, C, N, generator=g) * 2, 1)
    pred[1, 5, 77] = pred[3, 5, 9] = 2.0  # exact tie across clouds: the first in (b,n) order wins
    ct, bv, bi = R.class_T_pick(pred)
    ctc, bvc, bic = T.class_T_pick(pred.cuda())
    assert torch.equal(ctc.cpu(), ct) and torch.equal(bvc.cpu(), bv) and torch.equal(bic.cpu(), bi)
    assert int(bi[5]) == 1 * N + 77
    sigma = torch.full((C,), 0.4)
    ema = torch.softmax(torch.randn(C, C, generator=g) + torch.eye(C) * 5, 1)
    corr, ema2 = R.class_T_step(ct, ema, sigma, R.LABEL_PROJ_33, 0.999, 0.999)
    corrc, ema2c = T.class_T_step(ctc, ema.cuda(), sigma.cuda(), T.LABEL_PROJ_33, 0.999, 0.999)
    close(corrc, corr, rtol=1e-5, atol=1e-8)
    close(ema2c, ema2, rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize("C,B,N", [(33, 2, 16000), (17, 5, 1001), (5, 1, 15), (33, 1, 37)])
def test_apply_tile_kernel_vs_warp_kernel(cuda, C, B, N):
    """The lane-per-point tile kernel (bulk async tiles of 16 points, default for C in {5,17,33}) against the
    warp-per-point kernel (geot_group_points_config(1) selects it): forward, grad_logits, grad_insT, including a
    ragged last tile, negative entries (sign term of the L1-norm gradient) and only-logits / only-insT gradients."""
    from geot_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(C + N)
    logits = torch.randn(B, C, N, device="cuda", generator=g)
    ema = torch.softmax(torch.randn(C, C, device="cuda", generator=g), 1)
    insT = torch.randn(B * N, C, C, device="cuda", generator=g)  # signed on purpose
    go = torch.randn(B, C, N, device="cuda", generator=g)

    def run(need_l, need_i):
        l = logits.clone().requires_grad_(need_l); i = insT.clone().requires_grad_(need_i)
        out = T.apply_transition(l, ema, i, 0.9)
        out.backward(go)
        return out.detach(), l.grad, i.grad

    res = {}
    for direct in (0, 1):
        _lib.lib().geot_group_points_config(direct)
        try:
            res[direct] = [run(True, True), run(True, False), run(False, True)]
        finally:
            _lib.lib().geot_group_points_config(0)
    for a, b in zip(res[0], res[1]):
        for x, y in zip(a, b):
            assert (x is None) == (y is None)
            if x is not None:
                assert rel(x, y) < 1e-5
    ref = R.transition_apply(logits, ema, insT, 0.9)
    assert rel(res[0][0][0], ref) < 1e-5


# ---------------------------------------------------------------------------------------------
# Poly1FocalLoss_U_corr (the criterion consuming the corrected logits)
# ---------------------------------------------------------------------------------------------
from tests.test_transition_oracle_cpu import LOSS_VARIANTS, loss_case  # noqa: E402  (golden fixtures of the reference class)


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_loss(cuda, tag, variant, hp):
    from geot_b200 import losses
    logits, labels, conf, mask, thresh, exp_loss, exp_grad = loss_case(tag, variant)
    x = logits.cuda().requires_grad_(True)
    crit = losses.Poly1FocalLoss_U_corr(**hp)
    loss = crit(x, labels.cuda(), conf.cuda(), thresh=thresh, mask=None if mask is None else mask.cuda())
    close(loss, exp_loss, rtol=1e-5, atol=1e-8)   # north_star tolerance: 1e-5 relative
    loss.backward()
    assert rel(x.grad, exp_grad) < 1e-5


def test_poly1_focal_loss_large_vs_oracle(cuda):
    """cfg4 size (B=2+2+2 views -> 6 x 16000 points, C=33): against the oracle's torch expression on the GPU,
    nothing kept (denominator 0.001), everything kept, and run-to-run identical bits."""
    from geot_b200 import losses
    g = torch.Generator(device="cuda").manual_seed(7)
    B, C, N = 6, 33, 16000
    logits = (torch.randn(B, C, N, device="cuda", generator=g) * 4).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
    conf = torch.rand(B, N, device="cuda", generator=g)
    crit = losses.Poly1FocalLoss_U_corr()
    for thresh in (0.95, 0.0, 2.0):
        logits.grad = None
        loss = crit(logits, labels, conf, thresh=thresh)
        loss.backward()
        g1 = logits.grad.clone(); logits.grad = None
        ref = R.poly1_focal_u_corr(logits, labels, conf, thresh=thresh)
        ref.backward()
        close(loss, ref, rtol=1e-5, atol=1e-8)
        if thresh < 2.0:
            assert rel(g1, logits.grad) < 1e-5
        else:
            assert float(g1.abs().max()) == 0.0 and float(loss) == 0.0
        logits.grad = None
        l2 = crit(logits, labels, conf, thresh=thresh); l2.backward()
        assert torch.equal(l2, loss) and torch.equal(logits.grad, g1)


# ---------------------------------------------------------------------------------------------
# Poly1FocalLoss_U_T (the beta-scaled variant, train.py:594-596)
# ---------------------------------------------------------------------------------------------
from tests.test_transition_oracle_cpu import loss_t_case  # noqa: E402


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_u_t(cuda, tag, variant, hp):
    from geot_b200 import losses
    logits, labels, conf, pred_t, mask, thresh, exp_loss, exp_grad, exp_grad_t = loss_t_case(tag, variant)
    x = logits.cuda().requires_grad_(True)
    pt = pred_t.cuda().requires_grad_(True)
    crit = losses.Poly1FocalLoss_U_T(**hp)
    loss = crit(x, labels.cuda(), conf.cuda(), torch.eye(x.shape[1], device="cuda"), pt, thresh=thresh,
                mask=None if mask is None else mask.cuda())
    close(loss, exp_loss, rtol=1e-5, atol=1e-8)   # north_star tolerance: 1e-5 relative
    loss.backward()
    assert rel(x.grad, exp_grad) < 1e-5
    assert rel(pt.grad, exp_grad_t) < 1e-5
    assert x.grad.shape == x.shape and pt.grad.shape == pt.shape


def test_poly1_focal_u_t_large_vs_oracle(cuda):
    """cfg4 size (6 x 16000 points, C=33) against the oracle's torch expression on the GPU; with pred_u_t = conf in the
    label row beta is exactly 1 and the loss equals Poly1FocalLoss_U_corr's on the same logits (a size-independent property)."""
    from geot_b200 import losses
    g = torch.Generator(device="cuda").manual_seed(9)
    B, C, N = 6, 33, 16000
    logits = (torch.randn(B, C, N, device="cuda", generator=g) * 4).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
    conf = torch.rand(B, N, device="cuda", generator=g) * 0.9 + 0.1
    pred_t = torch.softmax(torch.randn(B, C, N, device="cuda", generator=g) * 2, dim=1).requires_grad_(True)
    crit = losses.Poly1FocalLoss_U_T()
    T = torch.eye(C, device="cuda")
    for thresh in (0.95, 0.0, 2.0):
        logits.grad = None; pred_t.grad = None
        loss = crit(logits, labels, conf, T, pred_t, thresh=thresh)
        loss.backward()
        g1, g1t = logits.grad.clone(), pred_t.grad.clone()
        logits.grad = None; pred_t.grad = None
        ref = R.poly1_focal_u_t(logits, labels, conf, pred_t, thresh=thresh)
        ref.backward()
        close(loss, ref, rtol=1e-5, atol=1e-8)
        if thresh < 2.0:
            assert rel(g1, logits.grad) < 1e-5 and rel(g1t, pred_t.grad) < 1e-5
        else:
            assert float(g1.abs().max()) == 0.0 and float(g1t.abs().max()) == 0.0 and float(loss) == 0.0
        logits.grad = None; pred_t.grad = None
        l2 = crit(logits, labels, conf, T, pred_t, thresh=thresh); l2.backward()
        assert torch.equal(l2, loss) and torch.equal(logits.grad, g1) and torch.equal(pred_t.grad, g1t)
    ones = torch.zeros(B, C, N, device="cuda").scatter_(1, labels.unsqueeze(1), conf.unsqueeze(1)).contiguous()
    l_beta1 = crit(logits.detach(), labels, conf, T, ones + (ones == 0).float(), thresh=0.5)
    l_corr = losses.Poly1FocalLoss_U_corr()(logits.detach(), labels, conf, thresh=0.5)
    assert torch.equal(l_beta1, l_corr)


def test_sigt_clamp_edge_flips_are_counted_and_bounded(cuda):
    """sig_t_mean fwd+bwd at B=16 x 16384, C=33 (profiles/REPORT_r01g.md reported a 2.7e-3 max relative error of the
    weight gradient there and attributed it to clamp-edge flips).  This test makes that explicit: against the float64
    evaluation of the same expression, (1) every pre-clamp gradient entry whose pre-activation is NOT within fp32
    rounding of a clamp edge agrees to 1e-5; (2) the entries that ARE near an edge either agree or are exactly zero
    (the clamp decided the other way in fp32 - torch's own fp32 does the same) and there are only a handful of them;
    (3) the weight gradient is within 1e-5 of the float64 one plus the sum of the magnitudes of exactly those flipped
    terms."""
    C, B, N = 33, 16, 16384
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.softmax(torch.randn((B, C, N), device="cuda", generator=g) * 3, 1)
    W = (torch.rand((C, C, 2 * C), device="cuda", generator=g) - 0.5) * (2 / (2 * C) ** 0.5)
    W[:, :, :C] += torch.eye(C, device="cuda") * 0.5
    cm = torch.softmax(torch.randn((C, C), device="cuda", generator=g), 1)
    go = torch.randn((B * N, C, C), device="cuda", generator=g)
    insT = F_.sigt_forward(x, W, cm)
    gW, g_raw = F_.sigt_backward(x, W, cm, go)
    g_raw = g_raw.view(B * N, C, C)
    # float64 evaluation, chunked over points
    p = x.permute(0, 2, 1).reshape(-1, C).double()
    W64, cm64 = W.double(), cm.double()
    bias = torch.einsum("kc,kjc->kj", cm64, W64[:, :, C:])                      # (C, C): W[kk][:, C:] @ cm[kk]
    gW64 = torch.zeros_like(W64)
    flipped_terms = torch.zeros_like(W64)
    n_near = n_flip = 0
    worst = 0.0
    for s in range(0, B * N, 32768):
        ps = p[s:s + 32768]
        z = (torch.einsum("nc,kjc->nkj", ps, W64[:, :, :C]) + bias[None]).requires_grad_(True)
        out = torch.nn.functional.normalize(torch.clamp(z, 1e-5, 1 - 1e-5), p=1, dim=2)
        gz, = torch.autograd.grad(out, z, go[s:s + 32768].double())
        # the gradient the entry WOULD get if the clamp let it through (for entries fp64 blocks but fp32 may pass)
        zc = torch.clamp(z.detach(), 1e-5, 1 - 1e-5)
        sden = zc.sum(2, keepdim=True)
        gos = go[s:s + 32768].double()
        g_pass = (gos - (gos * zc).sum(2, keepdim=True) / sden) / sden
        # fp32 evaluates z (a 2C-term dot product) with an absolute error of up to ~2C * 2^-24 * sum|terms|: an entry is
        # "near an edge" when its float64 pre-activation is closer to the edge than that
        zabs = torch.einsum("nc,kjc->nkj", ps, W64[:, :, :C].abs()) + torch.einsum("kc,kjc->kj", cm64, W64[:, :, C:].abs())[None]
        delta = 16 * 2.0 ** -24 * zabs
        near = ((z.detach() - 1e-5).abs() <= delta) | ((z.detach() - (1 - 1e-5)).abs() <= delta)
        # forward: entries just above the 1e-5 clamp come from cancelling O(0.1) terms, so the bound is relative to
        # the magnitude of the terms of the dot product (normalised like the output), not to the tiny result
        sd = torch.clamp(z.detach(), 1e-5, 1 - 1e-5).sum(2, keepdim=True)
        worst = max(worst, float(((insT[s:s + 32768].double() - out.detach()).abs() / (zabs / sd + out.detach())).max()))
        ours = g_raw[s:s + 32768].double()
        # g_pass is a difference of two terms: the 1e-5 bound is relative to their magnitudes (tests/tolerance.py)
        scale = (gos.abs() + (gos.abs() * zc).sum(2, keepdim=True) / sden) / sden
        ok = (ours - gz).abs() <= 1e-5 * scale
        assert bool(ok[~near].all()), "pre-clamp gradient differs away from the clamp edges"
        flip = near & ~ok
        assert bool(((ours[flip] == 0) | ((ours[flip] - g_pass[flip]).abs() <= 1e-5 * scale[flip])).all())
        n_near += int(near.sum()); n_flip += int(flip.sum())
        inp = torch.cat([ps[:, None, :].expand(-1, C, -1), cm64[None].expand(ps.shape[0], -1, -1)], 2)   # (n, C, 2C)
        gW64 += torch.einsum("nkj,nkc->kjc", gz, inp)
        flipped_terms += torch.einsum("nkj,nkc->kjc", (g_pass.abs() * flip), inp)
    print(f"\n[sig_t clamp edges] {n_near} of {B * N * C * C} pre-activations within fp32 rounding of a clamp edge, {n_flip} decided "
          f"the other way in fp32; forward max err relative to sum|terms| {worst:.2e}")
    assert worst <= 1e-5
    assert n_near <= 1e-4 * B * N * C * C and n_flip <= n_near
    err = (gW.double() - gW64).abs()
    bound = 1e-5 * gW64.abs().max() + flipped_terms + 1e-9
    assert bool((err <= bound).all()), float((err / bound).max())


# ---------------------------------------------------------------------------------------------
# Poly1FocalLoss_U_T_v1 (class-level T^T . p inside the loss, build.py:690-795)
# ---------------------------------------------------------------------------------------------
from tests.test_transition_oracle_cpu import loss_tv1_case  # noqa: E402


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_u_t_v1(cuda, tag, variant, hp):
    from geot_b200 import losses
    c = loss_tv1_case(tag, variant)
    x = c["logits"].cuda().requires_grad_(True)
    pu = c["pred_u"].cuda().requires_grad_(True)
    dT = c["delta_T"].cuda().requires_grad_(True)
    loss, ret = losses.Poly1FocalLoss_U_T_v1(**hp)(x, c["labels"].cuda(), c["conf"].cuda(), c["T"].cuda(), pu, dT, thresh=c["thresh"],
                                                  mask=None if c["mask"] is None else c["mask"].cuda())
    assert ret is dT                                            # the reference returns (loss, delta_T)
    close(loss, c["loss"], rtol=1e-5, atol=1e-8)
    loss.backward()
    assert rel(x.grad, c["grad"]) < 1e-5 and rel(pu.grad, c["grad_u"]) < 1e-5 and rel(dT.grad, c["grad_dT"]) < 1e-5


def test_poly1_focal_u_t_v1_cfg4_size_vs_oracle(cuda):
    """6 x 16000 points, C=33 against the oracle's torch expression in float64 on the GPU."""
    from geot_b200 import losses
    g = torch.Generator(device="cuda").manual_seed(10)
    B, C, N = 6, 33, 16000
    logits = (torch.randn(B, C, N, device="cuda", generator=g) * 4).requires_grad_(True)
    labels = torch.randint(0, C, (B, N), device="cuda", generator=g)
    conf = torch.rand(B, N, device="cuda", generator=g) * 0.9 + 0.1
    pred_u = torch.softmax(torch.randn(B, C, N, device="cuda", generator=g) * 2, dim=1).requires_grad_(True)
    T = torch.nn.functional.normalize(torch.rand(C, C, device="cuda", generator=g) + torch.eye(C, device="cuda") * 4, p=1, dim=1)
    dT = (torch.randn(C, C, device="cuda", generator=g) * 0.01).requires_grad_(True)
    loss, _ = losses.Poly1FocalLoss_U_T_v1()(logits, labels, conf, T, pred_u, dT, thresh=0.5)
    loss.backward()
    x64, pu64, dT64 = (t.detach().double().requires_grad_(True) for t in (logits, pred_u, dT))
    ref = R.poly1_focal_u_t_v1(x64, labels, conf.double(), T.double(), pu64, dT64, thresh=0.5)
    ref.backward()
    assert rel(loss, ref) < 1e-5
    assert rel(logits.grad, x64.grad) < 1e-5 and rel(pred_u.grad, pu64.grad) < 1e-5 and rel(dT.grad, dT64.grad) < 1e-5


def test_manifold_processing_order_changes_no_bit(cuda):
    """The Morton-cell processing order of the manifold kernels (L2 locality) must not change a single bit of the loss,
    the coefficients or the gradient - with the order, without it, and with a reversed one."""
    g = torch.Generator(device="cuda").manual_seed(3)
    B, N, C, k = 2, 6000, 17, 16
    pos = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=11)).cuda()
    labels = torch.randint(0, 4, (B, N), device="cuda", generator=g)
    insT = torch.softmax(torch.randn((B * N, C, C), device="cuda", generator=g), 2).contiguous()
    idx = T._neighbours(k, pos)
    order = F_.spatial_order(pos)
    assert torch.equal(torch.sort(order.long(), 1).values, torch.arange(N, device="cuda").expand(B, N))   # a permutation
    plan = F_.ScatterPlan(idx, B, N)
    gl = torch.ones((1,), device="cuda")
    res = []
    for o in (None, order, order.flip(1).contiguous()):
        loss, coef = F_.manifold_loss_forward(pos, labels, insT, idx, 1.0, F_.MANIFOLD_3D, order=o)
        grad = F_.manifold_loss_backward(insT, idx, coef, plan, gl, order=o)
        res.append((loss, coef, grad))
    for r in res[1:]:
        assert all(torch.equal(a, b) for a, b in zip(res[0], r))


@pytest.mark.parametrize("case", ["cfg4", "small_k", "ragged_tail", "random_idx", "hub", "feature_mode", "k33"])
def test_manifold_tile_kernels_equal_warp_kernels(cuda, case):
    """The tile kernels (32 Morton-consecutive points per CTA, the distinct matrices they touch staged in shared memory
    slice by slice; mutual neighbours merged in the backward) against the warp-per-point kernels (mode 6): loss,
    coefficients and gradient bit for bit - at the
    cfg4 size (slices of 256 entries), with N % 32 != 0, with random neighbours (1056 distinct rows per tile: slices of
    32 entries), with one point listed by every other (its in-neighbours overflow the table: global reads), in the
    feature mode, with and without the processing order; k = 33 stays on the warp-per-point kernels."""
    from geot_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(11)
    B, N, C, k, mode = {"cfg4": (2, 16000, 33, 32, F_.MANIFOLD_3D), "small_k": (3, 4099, 17, 7, F_.MANIFOLD_3D),
                        "ragged_tail": (2, 1001, 5, 32, F_.MANIFOLD_3D), "random_idx": (2, 5000, 9, 32, F_.MANIFOLD_3D),
                        "hub": (1, 6000, 6, 16, F_.MANIFOLD_3D), "feature_mode": (2, 3000, 17, 12, F_.MANIFOLD_FEATURE),
                        "k33": (1, 2000, 5, 33, F_.MANIFOLD_3D)}[case]
    pos = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=21)).cuda()
    labels = torch.randint(0, 3, (B, N), device="cuda", generator=g)
    insT = torch.softmax(torch.randn((B * N, C, C), device="cuda", generator=g), 2).contiguous()
    if case == "random_idx":
        idx = torch.randint(0, N, (B, N, k), device="cuda", generator=g, dtype=torch.int32)
    else:
        idx = T._neighbours(k, pos)
    if case == "hub":
        idx[:, :, 3] = 77          # every point lists point 77: 6000 in-neighbours in one tile
        idx[:, 100:200, 5] = 4242  # and a second, smaller hub
    order = F_.spatial_order(pos)
    plan = F_.ScatterPlan(idx, B, N)
    gl = torch.full((1,), 0.7, device="cuda")
    res = {}
    try:
        for name, m in (("tile", 7), ("warp", 6), ("default", 0)):   # 7: the tile kernels whatever C; 0: chosen by C
            _lib.lib().geot_group_points_config(m)
            for o in (order, None):
                loss, coef = F_.manifold_loss_forward(pos, labels, insT, idx, 0.8, mode, order=o)
                grad = F_.manifold_loss_backward(insT, idx, coef, plan, gl, order=o)
                res[(name, o is None)] = (loss, coef, grad)
    finally:
        _lib.lib().geot_group_points_config(0)
    ref = res[("warp", True)]
    assert bool(torch.isfinite(ref[2]).all()) and float(ref[2].abs().max()) > 0
    for key, r in res.items():
        for what, a, b in zip(("loss", "coef", "grad"), ref, r):
            assert torch.equal(a, b), f"{key} {what}: {(a != b).sum().item()} of {a.numel()} values differ"


@pytest.mark.parametrize("B,N,D", [(1, 1, 3), (3, 5, 3), (2, 1000, 3), (2, 1025, 3), (2, 4096, 3), (2, 6000, 3), (2, 16000, 3), (2, 16384, 3), (1, 16385, 3),
                                   (2, 40000, 3), (2, 3000, 17), (1, 777, 1)])
def test_spatial_order_is_a_permutation_and_local(cuda, B, N, D):
    """geot_spatial_order: a permutation of every cloud (fine 30-bit Morton sort up to 16384 points, 16^3-cell counting
    sort above), and consecutive points of it are closer to each other than consecutive points of the input."""
    g = torch.Generator(device="cuda").manual_seed(N + D)
    if D == 3 and N >= 1000:
        pos = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=3)).cuda()
    else:
        pos = torch.randn((B, N, D), device="cuda", generator=g)
    order = F_.spatial_order(pos)
    assert order.dtype == torch.int32 and order.shape == (B, N)
    assert torch.equal(torch.sort(order.long(), 1).values, torch.arange(N, device="cuda").expand(B, N))
    if N >= 1000 and D == 3:
        walk = lambda p: (p[:, 1:] - p[:, :-1]).norm(dim=-1).mean().item()
        sorted_pos = torch.gather(pos, 1, order.long().unsqueeze(-1).expand(-1, -1, D))
        assert walk(sorted_pos) < 0.25 * walk(pos)
    pos[0, 0, 0] = float("nan")          # scheduling only: still a permutation
    order = F_.spatial_order(pos)
    assert torch.equal(torch.sort(order.long(), 1).values, torch.arange(N, device="cuda").expand(B, N))
