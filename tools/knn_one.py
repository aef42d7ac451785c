import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from geot_b200 import _lib, synth, functional as F
b, n_per, k = 8, int(sys.argv[2]) if len(sys.argv) > 2 else 16384, int(sys.argv[1])
n = b * n_per
xyz = torch.from_numpy(synth.tooth_batch(b, n_per)).cuda().reshape(n, 3).contiguous()
off = torch.arange(1, b + 1, device="cuda", dtype=torch.int32) * n_per
for fl in (0, _lib.GEOT_KNN_NO_GRID):
    for _ in range(2):
        F.knnquery(k, xyz, xyz, off, off, m_per=n_per, flags=fl)
torch.cuda.synchronize()
