"""Needs 2 GPUs (skips otherwise; the 2-GPU gpurun leg runs it).
(1) Two devices in ONE process: every >48 KB dynamic-shared-memory kernel must also launch on cuda:1 after having run on
    cuda:0 (cudaFuncAttributeMaxDynamicSharedMemorySize is per device context - ADVICE r1).
(2) The statistics exchange of geot_b200/dist.py over NCCL, one process per GPU: the combined class-anchor pick equals
    the single-process pick over the concatenated batch (examples/segmentation/train.py:509-526)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _need2():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")


def test_two_devices_in_one_process():
    _need2()
    from geot_b200 import functional as F
    from geot_b200 import synth, transition as gt
    res = []
    xyz_h = torch.from_numpy(synth.tooth_batch(2, 16384))
    g = torch.Generator().manual_seed(0)
    feat_h = torch.randn((2, 8, 16384), generator=g)
    known_h = torch.randn((2, 8, 1024), generator=g)
    pred_h = torch.softmax(torch.randn((2, 33, 2048), generator=g), 1)
    W_h = torch.randn((33, 33, 66), generator=g) * 0.1
    cm_h = torch.softmax(torch.randn((33, 33), generator=g), 1)
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            xyz, feat = xyz_h.to(dev), feat_h.to(dev)
            idx = F.furthest_point_sampling(xyz, 1024)                                   # fps_mgr_kernel: 217 KB smem
            cen = F.gather_rows(xyz, idx)
            nbr = F.ball_query(cen, xyz, 0.1, 32)                                        # grid kernels
            grp = F.group_points(feat, nbr)                                              # group_rows_kernel: 64 KB rows
            gg = F.group_points_grad(torch.ones_like(grp), F.grouping_plan(nbr, 2, 16384), 16384)   # planrow: 226 KB
            d2, i3 = F.three_nn(xyz, cen)                                                # three_nn_grid_kernel
            w = torch.full((2, 16384, 3), 1.0 / 3, device=dev)
            out = F.three_interpolate(known_h.to(dev), i3, w)                            # rows4 / rows kernels
            go = F.three_interpolate_grad(torch.ones_like(out), w, F.ScatterPlan(i3, 2, 1024), 1024)   # wrows: 226 KB
            kidx, _ = F.knnquery(16, xyz.view(-1, 3), xyz.view(-1, 3), None, None, n_per=16384, m_per=16384)
            insT = F.sigt_forward(pred_h.to(dev), W_h.to(dev), cm_h.to(dev))             # sigt_kernel<33>: 144 KB
            torch.cuda.synchronize(dev)
            res.append([t.cpu() for t in (idx, nbr, grp, gg, i3, out, go, kidx, insT)])
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a, b), "cuda:1 differs from cuda:0"
    for a, b in zip(res[0], res[2]):
        assert torch.equal(a, b)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _nccl_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from geot_b200 import dist as gd
        from geot_b200 import transition as gt
        g = torch.Generator().manual_seed(11)
        B, C, N = 8, 33, 4096
        pred = torch.softmax(torch.randn(B, C, N, generator=g) * 2, 1)
        pred[1, 3, 5] = pred[6, 3, 9] = 2.0       # exact tie across ranks -> the lower rank's point must win
        pred = pred.to(dev)
        first, cnt = gd.shard_clouds(B)
        assert (first, cnt) == (rank * (B // world), B // world)
        ct_local, bv_local, _ = gt.class_T_pick(pred[first:first + cnt].contiguous())      # the CUDA kernel on the shard
        ct = gd.combine_class_T(ct_local, bv_local)                                        # NCCL all_gather
        ct_ref, _, bi_ref = gt.class_T_pick(pred)                                          # single-process pick
        assert torch.equal(ct, ct_ref), "combined anchor rows differ from the single-process pick"
        assert int(bi_ref[3]) == 1 * N + 5
        # the whole per-step state update is then identical on every rank
        st = gt.ClassTransitionState(C, device=dev)
        sigma = torch.full((C,), 0.4, device=dev)
        corr = st.step(pred[first:first + cnt].contiguous(), sigma)
        st1 = gt.ClassTransitionState(C, device=dev)
        c1, e1 = gt.class_T_step(ct_ref, st1.ema_t, sigma, st1.label_proj, st1.geo_lambma, st1.ema_t_decay)
        assert torch.equal(corr, c1) and torch.equal(st.ema_t, e1)
        cm = torch.full((C, C), float(rank + 1), device=dev)
        cn = torch.arange(C, dtype=torch.float32, device=dev) * (rank + 1)
        gd.allreduce_class_means(cm, cn)                                                   # NCCL all_reduce
        assert torch.equal(cn.cpu(), torch.arange(C, dtype=torch.float32) * 3)
        assert torch.allclose(cm[1:].cpu(), torch.full((C - 1, C), 5.0 / 3.0))
        p1 = torch.nn.Parameter(torch.zeros(33, 66, device=dev))
        p1.grad = torch.full((33, 66), float(rank + 1), device=dev)
        h = gd.GradAllReduce([p1], average=True).launch()                                  # NCCL all_reduce, side stream
        h.wait()
        torch.cuda.synchronize()
        assert torch.allclose(p1.grad.cpu(), torch.full((33, 66), 1.5))
        # the fused exchange (one packed all-gather + one combine kernel) must give what the three separate ones give
        cm2 = torch.full((C, C), float(rank + 1), device=dev)
        cn2 = torch.arange(C, dtype=torch.float32, device=dev) * (rank + 1)
        p2 = [torch.nn.Parameter(torch.zeros(33, 66, device=dev)) for _ in range(3)]
        for i, q in enumerate(p2):
            q.grad = torch.full((33, 66), float(rank + 1 + i), device=dev) if (rank + i) % 2 == 0 else None
        exp_g = [sum((r + 1 + i) if (r + i) % 2 == 0 else 0.0 for r in range(world)) / world for i in range(3)]
        ex = gd.StatsExchange(C, params=p2).launch(ct_local, bv_local, cm2, cn2)
        f_ct, f_cm, f_cn = ex.wait()
        torch.cuda.synchronize()
        assert torch.equal(f_ct, ct_ref), "fused exchange: anchor rows"
        assert torch.equal(f_cn, cn) and torch.allclose(f_cm, cm, rtol=1e-6, atol=0), "fused exchange: class means"
        for q, e in zip(p2, exp_g):
            assert torch.allclose(q.grad.cpu(), torch.full((33, 66), e)), "fused exchange: gradients"
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_statistics_exchange_over_nccl(tmp_path):
    _need2()
    import torch.multiprocessing as mp
    mp.spawn(_nccl_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
