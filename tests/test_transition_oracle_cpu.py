"""oracle/transition_ref.py against fixtures produced by the REFERENCE'S OWN code (tests/golden/
make_transition_golden.py ran transformer.py / insT_loss.py / train.py statements on CPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import transition_ref as R

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "transition_golden.npz"))


def t(name):
    return torch.from_numpy(G[name])


CASES = ["C17", "C33", "C5"]


@pytest.mark.parametrize("tag", CASES)
def test_sig_t_mean_forward_and_grads(tag):
    W = t(f"{tag}_W").requires_grad_(True)
    x = t(f"{tag}_x").requires_grad_(True)
    insT = R.sig_t_mean_forward(x, t(f"{tag}_cm"), W)
    torch.testing.assert_close(insT, t(f"{tag}_insT"), rtol=1e-6, atol=1e-7)
    gW, gx = torch.autograd.grad(insT, [W, x], t(f"{tag}_go"))
    torch.testing.assert_close(gW, t(f"{tag}_gW"), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(gx, t(f"{tag}_gx"), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(R.sig_t_forward(t(f"{tag}_x"), t(f"{tag}_sigt_W")), t(f"{tag}_sigt_T"), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("tag", CASES)
def test_transition_apply(tag):
    logits = t(f"{tag}_logits").requires_grad_(True)
    insT = t(f"{tag}_insT").requires_grad_(True)
    out = R.transition_apply(logits, t(f"{tag}_ema_t_corr"), insT, 0.9)
    torch.testing.assert_close(out, t(f"{tag}_corr"), rtol=1e-6, atol=1e-7)
    gl, gi = torch.autograd.grad(out, [logits, insT], t(f"{tag}_go2"))
    torch.testing.assert_close(gl, t(f"{tag}_g_logits"), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(gi, t(f"{tag}_g_insT_apply"), rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("tag", CASES)
def test_manifold_losses(tag):
    k = int(G[f"{tag}_k"])
    insT = t(f"{tag}_insT").requires_grad_(True)
    pos, labels = t(f"{tag}_pos"), t(f"{tag}_labels")
    # default neighbours (the reference's own knn_point) and pinned neighbours must both reproduce it
    for idx in (None, t(f"{tag}_idx3")):
        l3 = R.threeD_space_loss(pos, labels, insT, k, sigma=0.5, idx=idx)
        torch.testing.assert_close(l3, t(f"{tag}_loss3"), rtol=1e-6, atol=1e-9)
        g3, = torch.autograd.grad(l3, insT)
        torch.testing.assert_close(g3, t(f"{tag}_g_insT_loss3"), rtol=1e-5, atol=1e-9)
    lf = R.feature_space_loss(t(f"{tag}_logits"), labels, insT, k, sigma=2.0, idx=t(f"{tag}_idxf"))
    torch.testing.assert_close(lf, t(f"{tag}_lossf"), rtol=1e-5, atol=1e-9)
    gf, = torch.autograd.grad(lf, insT)
    torch.testing.assert_close(gf, t(f"{tag}_g_insT_lossf"), rtol=1e-5, atol=1e-9)
    C = insT.shape[1]
    torch.testing.assert_close(R.identity_loss(insT.detach(), torch.eye(C).unsqueeze(0)), t(f"{tag}_loss_id"), rtol=1e-6, atol=1e-9)


def test_class_statistics():
    tag = "C17"
    assert R.LABEL_PROJ_17 == G[f"{tag}_label_proj"].tolist()
    assert R.LABEL_PROJ_33[:17] == R.LABEL_PROJ_17 and sorted(R.LABEL_PROJ_33) == list(range(33))
    class_T, best_v, best_i = R.class_T_pick(t(f"{tag}_pred_u"))
    assert torch.equal(class_T, t(f"{tag}_class_T"))
    corr, ema = R.class_T_step(class_T, t(f"{tag}_ema_t_in"), t(f"{tag}_sigma"), R.LABEL_PROJ_17, 0.999, 0.999)
    torch.testing.assert_close(corr, t(f"{tag}_ema_t_corr_out"), rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(ema, t(f"{tag}_ema_t_out"), rtol=1e-6, atol=1e-8)
    C = class_T.shape[0]
    cm, cn = torch.zeros((C, C)), torch.zeros((C,))
    for _ in range(2):
        R.mean_feature_update(cm, cn, t(f"{tag}_mf_probs"), t(f"{tag}_mf_target"))
    torch.testing.assert_close(cm, t(f"{tag}_mf_cm"), rtol=1e-6, atol=1e-8)
    assert torch.equal(cn, t(f"{tag}_mf_cnum"))


LG = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_golden.npz"))
LOSS_VARIANTS = [("thr", {}), ("low", {}), ("mask", {}), ("hp", dict(epsilon=0.5, alpha=-1.0, gamma=1.5))]


def loss_case(tag, variant):
    """Inputs and expected outputs of one golden case of tests/golden/make_loss_golden.py."""
    logits = torch.from_numpy(LG[f"{tag}_logits"]); labels = torch.from_numpy(LG[f"{tag}_labels"]); conf = torch.from_numpy(LG[f"{tag}_conf"])
    mask = torch.from_numpy(LG[f"{tag}_{variant}_mask"]) if f"{tag}_{variant}_mask" in LG else None
    thresh = float(LG[f"{tag}_{variant}_thresh"]) if f"{tag}_{variant}_thresh" in LG else 0.5
    return logits, labels, conf, mask, thresh, torch.from_numpy(LG[f"{tag}_{variant}_loss"]), torch.from_numpy(LG[f"{tag}_{variant}_grad"])


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_u_corr(tag, variant, hp):
    """oracle restatement of Poly1FocalLoss_U_corr against the reference class's own outputs (loss and gradient)."""
    logits, labels, conf, mask, thresh, exp_loss, exp_grad = loss_case(tag, variant)
    x = logits.clone().requires_grad_(True)
    loss = R.poly1_focal_u_corr(x, labels, conf, thresh=thresh, mask=mask, **hp)
    torch.testing.assert_close(loss, exp_loss, rtol=1e-6, atol=1e-9)
    loss.backward()
    torch.testing.assert_close(x.grad, exp_grad, rtol=1e-5, atol=1e-10)


LTG = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_t_golden.npz"))


def loss_t_case(tag, variant):
    """One golden case of Poly1FocalLoss_U_T (tests/golden/make_loss_golden.py, second half)."""
    f = lambda k: torch.from_numpy(LTG[k])
    mask = f(f"{tag}_{variant}_mask") if f"{tag}_{variant}_mask" in LTG else None
    return (f(f"{tag}_logits"), f(f"{tag}_labels"), f(f"{tag}_conf"), f(f"{tag}_pred_t"), mask, float(LTG[f"{tag}_{variant}_thresh"]),
            f(f"{tag}_{variant}_loss"), f(f"{tag}_{variant}_grad"), f(f"{tag}_{variant}_grad_t"))


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_u_t(tag, variant, hp):
    """oracle restatement of Poly1FocalLoss_U_T against the reference class's own outputs (loss and both gradients)."""
    logits, labels, conf, pred_t, mask, thresh, exp_loss, exp_grad, exp_grad_t = loss_t_case(tag, variant)
    x = logits.clone().requires_grad_(True)
    pt = pred_t.clone().requires_grad_(True)
    loss = R.poly1_focal_u_t(x, labels, conf, pt, thresh=thresh, mask=mask, **hp)
    torch.testing.assert_close(loss, exp_loss, rtol=1e-6, atol=1e-9)
    loss.backward()
    torch.testing.assert_close(x.grad, exp_grad, rtol=1e-5, atol=1e-10)
    torch.testing.assert_close(pt.grad, exp_grad_t, rtol=1e-5, atol=1e-10)


LV1 = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_tv1_golden.npz"))


def loss_tv1_case(tag, variant):
    """One golden case of Poly1FocalLoss_U_T_v1 (tests/golden/make_loss_golden.py, third part)."""
    f = lambda k: torch.from_numpy(LV1[k])
    mask = f(f"{tag}_{variant}_mask") if f"{tag}_{variant}_mask" in LV1 else None
    return dict(logits=f(f"{tag}_logits"), labels=f(f"{tag}_labels"), conf=f(f"{tag}_conf"), pred_u=f(f"{tag}_pred_u"), T=f(f"{tag}_T"),
                delta_T=f(f"{tag}_delta_T"), mask=mask, thresh=float(LV1[f"{tag}_{variant}_thresh"]), loss=f(f"{tag}_{variant}_loss"),
                grad=f(f"{tag}_{variant}_grad"), grad_u=f(f"{tag}_{variant}_grad_u"), grad_dT=f(f"{tag}_{variant}_grad_dT"))


@pytest.mark.parametrize("tag", ["C17", "C33", "C5"])
@pytest.mark.parametrize("variant,hp", LOSS_VARIANTS)
def test_golden_poly1_focal_u_t_v1(tag, variant, hp):
    """oracle restatement of Poly1FocalLoss_U_T_v1 against the reference class's own outputs (loss and three gradients)."""
    c = loss_tv1_case(tag, variant)
    x = c["logits"].clone().requires_grad_(True)
    pu = c["pred_u"].clone().requires_grad_(True)
    dT = c["delta_T"].clone().requires_grad_(True)
    loss = R.poly1_focal_u_t_v1(x, c["labels"], c["conf"], c["T"], pu, dT, thresh=c["thresh"], mask=c["mask"], **hp)
    torch.testing.assert_close(loss, c["loss"], rtol=1e-6, atol=1e-9)
    loss.backward()
    torch.testing.assert_close(x.grad, c["grad"], rtol=1e-5, atol=1e-10)
    torch.testing.assert_close(pu.grad, c["grad_u"], rtol=1e-5, atol=1e-9)
    torch.testing.assert_close(dT.grad, c["grad_dT"], rtol=2e-5, atol=1e-8)
