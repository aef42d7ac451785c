"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in geot_b200/dist.py: cloud sharding, the
class-level anchor combine (must equal the single-process pick over the concatenated batch), the class-mean
combine and the packed gradient all-reduce.  No CUDA compute here — that is the -m gpu suite."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from geot_b200 import dist as gd
        from oracle import transition_ref as R
        g = torch.Generator().manual_seed(11)
        B, C, N = 6, 17, 200
        pred = torch.softmax(torch.randn(B, C, N, generator=g) * 2, 1)
        pred[1, 3, 5] = pred[4, 3, 9] = 2.0       # exact tie across ranks -> the lower rank's point must win
        first, cnt = gd.shard_clouds(B)
        assert (first, cnt) == (rank * 3, 3)
        mine = pred[first:first + cnt]
        ct_local, bv_local, _ = R.class_T_pick(mine)  # (the GPU kernel's CPU restatement; host logic under test)
        ct = gd.combine_class_T(ct_local, bv_local)
        ct_ref, _, bi_ref = R.class_T_pick(pred)
        assert torch.equal(ct, ct_ref), "combined anchor rows differ from the single-process pick"
        assert int(bi_ref[3]) == 1 * N + 5
        # class means: two ranks with different counts
        cm = torch.full((C, C), float(rank + 1))
        cn = torch.arange(C, dtype=torch.float32) * (rank + 1)
        gd.allreduce_class_means(cm, cn)
        exp_cn = torch.arange(C, dtype=torch.float32) * 3
        assert torch.equal(cn, exp_cn)
        exp = (1 * 1 + 2 * 2) / 3.0
        assert torch.allclose(cm[1:], torch.full((C - 1, C), exp)) and torch.equal(cm[0], torch.full((C,), float(rank + 1)))
        # packed gradient all-reduce
        p1 = torch.nn.Parameter(torch.zeros(4, 3)); p2 = torch.nn.Parameter(torch.zeros(5))
        p1.grad = torch.full((4, 3), float(rank + 1)); p2.grad = None if rank == 0 else torch.ones(5)
        h = gd.GradAllReduce([p1, p2], average=True).launch()
        h.wait()
        assert torch.allclose(p1.grad, torch.full((4, 3), 1.5)) and torch.allclose(p2.grad, torch.full((5,), 0.5))
        # offset-format sharding
        offset = torch.tensor([100, 250, 300, 420, 500, 640], dtype=torch.int32)
        f, c, lo, hi, local = gd.shard_offsets(offset)
        assert (f, c) == (rank * 3, 3)
        assert (lo, hi) == ((0, 300) if rank == 0 else (300, 640))
        assert local.tolist() == ([100, 250, 300] if rank == 0 else [120, 200, 340])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world2_gloo(tmp_path):
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_shard_clouds_remainders():
    sys.path.insert(0, ROOT)
    from geot_b200 import dist as gd
    parts = [gd.shard_clouds(10, r, 4) for r in range(4)]
    assert parts == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert gd.shard_clouds(64, 5, 8) == (40, 8)
    # single process: everything is a no-op
    ct = torch.eye(3)
    assert gd.combine_class_T(ct, torch.ones(3)) is ct
