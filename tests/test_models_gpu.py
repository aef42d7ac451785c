"""BASELINE configs 2, 3, 4: the reference's OWN models / op wrappers, unchanged, run on the B200 once on
geot_b200 (through the pybind-shaped facades compat.install() registers) and once on the reference's own kernels
(oracle/_ref/*.so), same model object, same inputs, same seeds.

Bars (north_star): every index-producing call (FPS / ball query / kNN / three_nn) bit-equal between the two
backends; outputs and gradients within 1e-5 relative (max |a-b| <= 1e-5 * max |b| per tensor).  The reference's
backward kernels use float atomics, so the reference is not bit-reproducible against ITSELF: each test also runs
the reference backend twice and reports that run-to-run noise; where the noise floor alone exceeds 1e-5 (deep
networks amplify the last-bit differences of the atomics) the bound used is max(1e-5, 4 x noise), printed.
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from tests import refenv, refgpu  # noqa: E402


@pytest.fixture(scope="module")
def env(cuda):
    if not refenv.available():
        pytest.fail("no reference tree (oracle/_ref/src missing): run tools/stage_reference.sh in the authoring container")
    for name in ("pn2batch", "pointops_full", "pointops", "pn2ext"):
        assert refgpu.available(name), f"oracle/_ref/libref_{name}.so missing"
    refenv.install("geot")
    from tests import model_cases
    return model_cases


def _compare_index_calls(a, b):
    assert [n for n, _ in a] == [n for n, _ in b], "the two backends saw different call sequences"
    assert len(a) > 0
    for i, ((na, ta), (nb, tb)) in enumerate(zip(a, b)):
        assert ta.shape == tb.shape and ta.dtype == tb.dtype, (i, na)
        if not torch.equal(ta, tb):
            bad = (ta != tb).sum().item()
            raise AssertionError(f"index call #{i} {na}: {bad} of {ta.numel()} entries differ between geot_b200 and the reference kernels")


def _report_self_noise(mc, a, b, what):
    """geot_b200 against itself (second run): the segmented reductions are deterministic, so whatever differs here
    comes from the torch / cuDNN layers between the ops."""
    n = max(mc.rel_err(a[k], b[k]) for k in a)
    same = sum(bool(torch.equal(a[k], b[k])) for k in a)
    print(f"\n[{what}] geot_b200 run-to-run: {same}/{len(a)} tensors bit-identical, worst rel diff {n:.3e}")


def _compare_tensors(mc, ours, ref, ref2, what, tol=1e-5):
    print()
    worst = 0.0
    report = []
    for k in ref:
        noise = mc.rel_err(ref2[k], ref[k])
        err = mc.rel_err(ours[k], ref[k])
        bound = max(tol, 4.0 * noise)
        report.append((err / bound, k, err, noise))
        worst = max(worst, err)
    report.sort(reverse=True)
    print(f"[{what}] {len(ref)} tensors, worst rel err {worst:.3e}; top: " +
          "; ".join(f"{k} err={e:.2e} ref-noise={n:.2e}" for _, k, e, n in report[:4]))
    bad = [(k, e, n) for r, k, e, n in report if r > 1.0]
    assert not bad, f"{what}: beyond max(1e-5, 4 x reference run-to-run noise): {bad[:6]}"


def test_cfg3_pointnet2_msg_fwd_bwd(env):
    """BASELINE config 3: PointNet++ MSG segmentation fwd+bwd, B=16 x 16384."""
    mc = env
    model = mc.cfg3_model()
    data = mc.cfg3_inputs(16, 16384)
    with torch.no_grad():
        refenv.use("geot")
        shape = model.eval()({"pos": data["pos"][:1], "x": data["x"][:1]}).shape
    gout = torch.randn((16,) + tuple(shape[1:]), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    ours = mc.run_cfg3("geot", model, data, gout)
    ref = mc.run_cfg3("ref", model, data, gout)
    ref2 = mc.run_cfg3("ref", model, data, gout)
    # 4 FPS + 9 ball queries + 4 three_nn at the shipped shapes, all bit-equal
    names = [n for n, _ in ours["index_calls"]]
    assert names.count("furthest_point_sampling_wrapper") == 4 and names.count("ball_query_wrapper") == 9 and names.count("three_nn_wrapper") == 4, names
    _compare_index_calls(ours["index_calls"], ref["index_calls"])
    _compare_tensors(mc, {"logits": ours["out"]}, {"logits": ref["out"]}, {"logits": ref2["out"]}, "cfg3 logits")
    assert set(ours["grads"]) == set(ref["grads"]) and len(ref["grads"]) > 50
    _compare_tensors(mc, ours["grads"], ref["grads"], ref2["grads"], "cfg3 parameter gradients")
    _report_self_noise(mc, ours["grads"], mc.run_cfg3("geot", model, data, gout)["grads"], "cfg3 parameter gradients")


def test_cfg4_geot_student_fwd_bwd(env):
    """BASELINE config 4, model part: WholePartSeg(PointTransformer_seg_T) on 2 labeled + 2 strong + 2 weak clouds of
    16000 points - Group (FPS + knn_cuda), pointops.fps x3, PointnetFPModule x3 (three_nn + three_interpolate),
    DGCNN_Propagation x2 (transformer.py:282-303,343-379,1037-1049; pointnet2_modules.py:597-641)."""
    mc = env
    cfg, model, _ = mc.cfg4_models(17)
    data, data_u = mc.cfg4_inputs(16000)
    gout = torch.randn((6, 17, 16000), device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    ours = mc.run_cfg4_model("geot", model, data, data_u, gout)
    ref = mc.run_cfg4_model("ref", model, data, data_u, gout)
    ref2 = mc.run_cfg4_model("ref", model, data, data_u, gout)
    names = [n for n, _ in ours["index_calls"]]
    assert names.count("furthest_point_sampling") == 1 and names.count("furthestsampling_cuda") == 3 and names.count("three_nn") == 3, names
    _compare_index_calls(ours["index_calls"], ref["index_calls"])
    _compare_tensors(mc, {"logits": ours["out"]}, {"logits": ref["out"]}, {"logits": ref2["out"]}, "cfg4 logits")
    assert set(ours["grads"]) == set(ref["grads"]) and len(ref["grads"]) > 100
    _compare_tensors(mc, ours["grads"], ref["grads"], ref2["grads"], "cfg4 parameter gradients")
    _report_self_noise(mc, ours["grads"], mc.run_cfg4_model("geot", model, data, data_u, gout)["grads"], "cfg4 parameter gradients")


@pytest.mark.parametrize("C", [17, 33])
def test_cfg4_transition_block_vs_reference_statements(env, C):
    """BASELINE config 4, loss part: train.py:505-557 + :571 + :592 exec'd verbatim on the GPU (reference torch code)
    against the same block on geot_b200.transition / geot_b200.losses, B=2 x 16000, k=32."""
    mc = env
    cfg, _, T_predictor = mc.cfg4_models(C)
    inp = mc.cfg4_transition_inputs(C, 2, 16000)
    ref = mc.run_transition_reference(cfg, T_predictor, inp)
    ours = mc.run_transition_geot(cfg, T_predictor, inp)
    assert torch.equal(ours["class_T"], ref["class_T"])
    for k in ("ema_t_corr", "ema_t", "insT", "corr", "loss_3d", "loss_u", "g_logits"):
        e = mc.rel_err(ours[k], ref[k])
        print(f"[cfg4 transition C={C}] {k}: rel err vs reference fp32 {e:.3e}")
        assert e <= 1e-5, (k, e)
    # g_W sums BN = 32000 per-point terms per weight: two fp32 summation orders differ by ~sqrt(BN) ulp, so both are
    # measured against the SAME reference statements run in float64; ours must be within 1e-5 of that truth or at
    # least as close to it as the reference's own fp32 run.
    ref64 = mc.run_transition_reference(cfg, T_predictor, inp, double=True)
    for k in ("insT", "corr", "loss_3d", "loss_u", "g_logits", "g_W"):
        e_ours, e_ref = mc.rel_err(ours[k].double(), ref64[k]), mc.rel_err(ref[k].double(), ref64[k])
        print(f"[cfg4 transition C={C}] {k}: rel err vs float64 statements: ours {e_ours:.3e}, reference fp32 {e_ref:.3e}")
        assert e_ours <= max(1e-5, 1.5 * e_ref), (k, e_ours, e_ref)


def test_cfg2_point_transformer_pyramid_fwd_bwd(env):
    """BASELINE config 2: offset-format kNN(16) + grouping + interpolation pyramid fwd+bwd, B=8 x 16384
    (reference op wrappers openpoints/cpp/pointops/functions/pointops.py unchanged)."""
    mc = env
    inp = mc.cfg2_inputs(8, 16384)
    weights = mc.cfg2_weights()
    ours = mc.run_cfg2("geot", inp, weights)
    ref = mc.run_cfg2("ref", inp, weights)
    ref2 = mc.run_cfg2("ref", inp, weights)
    names = [n for n, _ in ours["index_calls"]]
    assert names.count("furthestsampling_cuda") == 4 and names.count("knnquery_cuda") == 12, names
    _compare_index_calls(ours["index_calls"], ref["index_calls"])
    _compare_tensors(mc, {"out": ours["out"]}, {"out": ref["out"]}, {"out": ref2["out"]}, "cfg2 output")
    _compare_tensors(mc, ours["grads"], ref["grads"], ref2["grads"], "cfg2 gradients")
    _report_self_noise(mc, ours["grads"], mc.run_cfg2("geot", inp, weights)["grads"], "cfg2 gradients")
