#!/usr/bin/env python
"""GPU box: the transition-matrix path of this library against the reference's own torch
expressions (oracle/transition_ref.py: the reference Python with `device=` made a parameter) on the SAME GPU and
inputs — forward + backward per op, at cfg4 (B_u=2 x N=16000, C=17 and C=33, k=32) and a cfg5-sized batch.

    python tests/transition_sweep.py [--out gpurun_out/transition_sweep.json]

Lives under tests/ because it executes oracle/ code (as the baseline being compared with, never as the product).
Algorithmic bytes per SURVEY.md §8(d): predictor writes 4*BN*C^2; apply reads 4*BN*C^2 fwd (+ writes it bwd);
manifold loss reads 4*BN*C^2 once in the ideal case."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from geot_b200 import functional as F_
from geot_b200 import synth, transition as T
from oracle import transition_ref as R

ap = argparse.ArgumentParser(); ap.add_argument("--out", default=None); args = ap.parse_args()
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0
rows = []


def timeit(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def report(op, shape, ours, ref, nbytes, err):
    to, tr = timeit(ours), timeit(ref)
    gbs = nbytes / to / 1e6
    rows.append({"op": op, "shape": shape, "ours_ms": to, "ref_ms": tr, "speedup": tr / to, "algo_bytes": nbytes,
                 "gbs": gbs, "hbm_frac": gbs / PEAK, "max_rel_err": err})
    print(f"{op:34s} {shape:28s} ours {to:8.4f} ms  ref(torch, same GPU) {tr:9.4f} ms  x{tr / to:7.2f}  "
          f"{gbs:7.0f} GB/s ({100 * gbs / PEAK:5.1f}% HBM)  max rel err {err:.1e}", flush=True)


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


for (B, N, C, k) in [(2, 16000, 17, 32), (2, 16000, 33, 32), (16, 16384, 33, 32)]:
    tag = f"B={B} N={N} C={C} k={k}"
    BN = B * N
    g = torch.Generator(device="cuda").manual_seed(C)
    logits = torch.randn(B, C, N, device="cuda", generator=g) * 3
    p = logits.softmax(1)
    cm = torch.rand(C, C, device="cuda", generator=g); cm = cm / cm.sum(1, keepdim=True)
    mod = T.sig_t_mean(C).cuda()
    W = torch.stack([l.weight for l in mod.fc]).detach()
    # ---- a16: T predictor, forward and backward w.r.t. the weights
    insT = mod(p, cm).detach()
    ref_T = R.sig_t_mean_forward(p, cm, W)
    report("sig_t_mean fwd", tag, lambda: mod(p, cm), lambda: R.sig_t_mean_forward(p, cm, W), 4 * BN * C * C + 4 * BN * C, rel(insT, ref_T))
    go = torch.randn(BN, C, C, device="cuda", generator=g)

    def ours_fb():
        mod.zero_grad(set_to_none=True)
        mod(p, cm).backward(go)

    Wr = W.clone().requires_grad_(True)

    def ref_fb():
        Wr.grad = None
        R.sig_t_mean_forward(p, cm, Wr).backward(go)

    ours_fb(); ref_fb()
    gW = torch.stack([l.weight.grad for l in mod.fc])
    report("sig_t_mean fwd+bwd", tag, ours_fb, ref_fb, 3 * 4 * BN * C * C, rel(gW, Wr.grad))
    # ---- a17: blend + normalise + apply, forward and backward (logits and insT)
    ema = torch.rand(C, C, device="cuda", generator=g); ema = ema / ema.sum(1, keepdim=True)
    lg = logits.clone().requires_grad_(True); it = insT.clone().requires_grad_(True)
    lg2 = logits.clone().requires_grad_(True); it2 = insT.clone().requires_grad_(True)
    go2 = torch.randn(B, C, N, device="cuda", generator=g)

    def ours_apply():
        lg.grad = None; it.grad = None
        T.apply_transition(lg, ema, it, 0.9).backward(go2)

    def ref_apply():
        lg2.grad = None; it2.grad = None
        R.transition_apply(lg2, ema, it2, 0.9).backward(go2)

    ours_apply(); ref_apply()
    with torch.no_grad():
        e_f = rel(T.apply_transition(logits, ema, insT, 0.9), R.transition_apply(logits, ema, insT, 0.9))
    report("apply_transition fwd", tag, lambda: T.apply_transition(logits, ema, insT, 0.9),
           lambda: R.transition_apply(logits, ema, insT, 0.9), 4 * BN * C * C + 8 * BN * C, e_f)
    report("apply_transition fwd+bwd", tag, ours_apply, ref_apply, 3 * 4 * BN * C * C, max(rel(lg.grad, lg2.grad), rel(it.grad, it2.grad)))
    # ---- a18: 3-D manifold loss, forward and backward (same neighbour indices on both sides)
    if BN * k * C * C * 4 * 3 < 60e9:  # the torch expression materialises (BN,k,C,C) several times
        pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
        labels = logits.argmax(1)
        loss_mod = T.threeD_space_loss(k=k, sigma=1.0, num_classes=C)
        idx, _ = F_.knnquery(k + 1, pos.view(-1, 3), pos.view(-1, 3), None, None, n_per=N, m_per=N)
        idx_local = (idx.view(B, N, k + 1)[:, :, 1:] - (torch.arange(B, device="cuda", dtype=torch.int32) * N).view(B, 1, 1)).contiguous()
        t1 = insT.clone().requires_grad_(True); t2 = insT.clone().requires_grad_(True)

        def ours_loss():
            t1.grad = None
            loss_mod(pos, labels, t1, idx=idx_local).backward()

        def ref_loss():
            t2.grad = None
            R.threeD_space_loss(pos, labels, t2, k, 1.0, idx=idx_local).backward()

        try:
            ours_loss(); ref_loss()
            report("threeD_space_loss fwd+bwd", tag, ours_loss, ref_loss, 3 * 4 * BN * C * C, rel(t1.grad, t2.grad))
        except Exception as e:  # keep the sweep going; the parity tests own correctness
            print("threeD_space_loss skipped:", repr(e)[:200])
    # ---- (f) rank 3: the criterion on the corrected logits, forward + backward
    from geot_b200 import losses
    crit = losses.Poly1FocalLoss_U_corr()
    conf = torch.rand(B, N, device="cuda", generator=g)
    labels_u = logits.argmax(1)
    x1 = logits.clone().requires_grad_(True); x2 = logits.clone().requires_grad_(True)

    def ours_crit():
        x1.grad = None
        crit(x1, labels_u, conf, thresh=0.5).backward()

    def ref_crit():
        x2.grad = None
        R.poly1_focal_u_corr(x2, labels_u, conf, thresh=0.5).backward()

    ours_crit(); ref_crit()
    report("Poly1FocalLoss_U_corr fwd+bwd", tag, ours_crit, ref_crit, 3 * 4 * BN * C + 12 * BN, rel(x1.grad, x2.grad))
    del logits, p, insT, go
    torch.cuda.empty_cache()

if args.out:
    json.dump({"hbm_peak_gbs": PEAK, "rows": rows}, open(args.out, "w"), indent=1)
