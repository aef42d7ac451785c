#!/usr/bin/env python
"""Dev tool (GPU box): self-kNN k=33 on the 2 x 16000 cfg4 clouds (one query of 32000 has an exact tie among its 34 best
and is replayed by knn_heap_warp_kernel) - for ncu captures of the replay."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from geot_b200 import synth
from geot_b200 import functional as F
B, N, k = 2, 16000, 33
pos = torch.from_numpy(synth.tooth_batch(B, N)).cuda()
x = pos.view(-1, 3).contiguous()
for _ in range(3):
    F.knnquery(k, x, x, None, None, n_per=N, m_per=N)
torch.cuda.synchronize()
