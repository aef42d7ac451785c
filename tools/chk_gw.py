import sys; sys.path.insert(0, "/root/repo")
import torch
from geot_b200 import transition as T
from oracle import transition_ref as R
def rel(a, b): return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))
for (B, N, C) in [(2, 16000, 17), (2, 16000, 33), (16, 16384, 33)]:
    g = torch.Generator(device="cuda").manual_seed(C)
    logits = torch.randn(B, C, N, device="cuda", generator=g) * 3
    p = logits.softmax(1)
    cm = torch.rand(C, C, device="cuda", generator=g); cm = cm / cm.sum(1, keepdim=True)
    mod = T.sig_t_mean(C).cuda()
    W = torch.stack([l.weight for l in mod.fc]).detach()
    go = torch.randn(B * N, C, C, device="cuda", generator=g)
    mod(p, cm).backward(go)
    gW = torch.stack([l.weight.grad for l in mod.fc])
    Wr = W.clone().requires_grad_(True)
    R.sig_t_mean_forward(p, cm, Wr).backward(go)
    Wd = W.double().clone().requires_grad_(True)
    R.sig_t_mean_forward(p.double(), cm.double(), Wd).backward(go.double())
    print(B, N, C, "ours vs f64", rel(gW, Wd.grad), " torch f32 vs f64", rel(Wr.grad, Wd.grad), " max|gW|", float(Wd.grad.abs().max()))
    # forward clamp sensitivity: count entries at the clamp boundary
