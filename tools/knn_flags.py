import os, sys
sys.path.insert(0, "/root/repo")
import torch
from geot_b200 import _lib, synth
B, N = 2, 16000
pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
x = pos.view(-1, 3).contiguous()
m = n = B * N
for k in (16, 17, 32, 33):
    nb = int(_lib.lib().geot_knnquery_workspace_bytes(B, n, m))
    ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    idx = torch.empty((m, k), dtype=torch.int32, device="cuda"); d2 = torch.empty((m, k), device="cuda")
    _lib.call("geot_knnquery", B, n, m, k, _lib.ptr(x), _lib.ptr(x), None, None, N, N, _lib.ptr(idx), _lib.ptr(d2), _lib.ptr(ws), nb, 0, _lib.stream_ptr())
    torch.cuda.synchronize()
    fl = ws[:m]
    print("k", k, "flagged", int(fl.sum()), "of", m, " adjacent equal d2 in output:", int((d2[:, 1:] == d2[:, :-1]).any(1).sum()))
    t = d2[fl.bool()][:3]
    print(t[:, :8] if len(t) else "")


def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


for k in (16, 24, 31, 32, 33, 40):
    nb = int(_lib.lib().geot_knnquery_workspace_bytes(B, n, m))
    ws = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    idx = torch.empty((m, k), dtype=torch.int32, device="cuda"); d2 = torch.empty((m, k), device="cuda")
    f = lambda: _lib.call("geot_knnquery", B, n, m, k, _lib.ptr(x), _lib.ptr(x), None, None, N, N, _lib.ptr(idx), _lib.ptr(d2), _lib.ptr(ws), nb, 0, _lib.stream_ptr())
    print("k", k, f"{timeit(f):.4f} ms  flagged {int(ws[:m].sum())}")
