"""Every callable of the three pybind-shaped facades (geot_b200/backends.py - what unmodified reference code binds
through compat.install()) executed with the reference's calling convention, next to the same call on the reference's
own kernels (tests/ref_backends.py over oracle/_ref/*.so).  Index results bit-equal; values / gradients within 1e-5
relative.  Covers the argument plumbing (sizes passed redundantly, caller-allocated and pre-zeroed outputs, n / m
derived from the gradient buffer's shape) that the op-level tests bypass."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from geot_b200 import backends, synth  # noqa: E402
from tests import refgpu  # noqa: E402


@pytest.fixture(scope="module")
def rb(cuda):
    for name in ("pn2batch", "pointops_full", "pointops", "pn2ext"):
        assert refgpu.available(name), f"oracle/_ref/libref_{name}.so missing"
    from tests import ref_backends
    return ref_backends


def _close(a, b, what):
    d = (a.double() - b.double()).abs().max().item()
    s = b.double().abs().max().item()
    assert d <= 1e-5 * s + 1e-12, (what, d, s)


def _cloud(B, N, seed=0):
    return torch.from_numpy(synth.tooth_batch(B, N, first_cloud=seed)).cuda()


def test_pointnet2_batch_cuda_all_nine(rb):
    B, N, M, ns, C = 3, 5000, 700, 24, 19
    g = torch.Generator(device="cuda").manual_seed(0)
    xyz = _cloud(B, N)
    feat = torch.randn((B, C, N), device="cuda", generator=g)
    res = {}
    for name, mod in (("geot", backends.pointnet2_batch_cuda), ("ref", rb.pointnet2_batch_cuda)):
        r = {}
        temp = torch.cuda.FloatTensor(B, N).fill_(1e10)                      # subsample.py:93-95
        idx = torch.cuda.IntTensor(B, M)
        mod.furthest_point_sampling_wrapper(B, N, M, xyz, temp, idx)
        r["fps"] = idx
        cen_t = torch.cuda.FloatTensor(B, 3, M)                              # group.py:156
        mod.gather_points_wrapper(B, 3, N, M, xyz.transpose(1, 2).contiguous(), idx, cen_t)
        r["gather"] = cen_t
        new_xyz = cen_t.transpose(1, 2).contiguous()
        bidx = torch.cuda.IntTensor(B, M, ns).zero_()                        # group.py:194
        mod.ball_query_wrapper(B, N, M, 0.12, ns, new_xyz, xyz, bidx)
        r["ball"] = bidx
        out = torch.cuda.FloatTensor(B, C, M, ns)                            # group.py:93
        mod.group_points_wrapper(B, C, N, M, ns, feat, bidx, out)
        r["group"] = out
        go = torch.randn((B, C, M, ns), device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
        gp = torch.zeros((B, C, N), device="cuda")                           # group.py:110
        mod.group_points_grad_wrapper(B, C, N, M, ns, go, bidx, gp)
        r["group_grad"] = gp
        gg = torch.zeros((B, C, N), device="cuda")
        go2 = torch.randn((B, C, M), device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
        mod.gather_points_grad_wrapper(B, C, N, M, go2, idx, gg)
        r["gather_grad"] = gg
        d2 = torch.cuda.FloatTensor(B, N, 3)                                 # upsampling.py:29-30
        nidx = torch.cuda.IntTensor(B, N, 3)
        mod.three_nn_wrapper(B, N, M, xyz, new_xyz, d2, nidx)
        r["nn_idx"], r["nn_d2"] = nidx, d2
        dist = torch.sqrt(d2)
        w = 1.0 / (dist + 1e-8)
        w = (w / w.sum(2, keepdim=True)).contiguous()
        known = torch.randn((B, C, M), device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
        o = torch.cuda.FloatTensor(B, C, N)                                  # upsampling.py:64
        mod.three_interpolate_wrapper(B, C, M, N, known, nidx, w, o)
        r["interp"] = o
        go3 = torch.randn((B, C, N), device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
        gk = torch.zeros((B, C, M), device="cuda")                           # upsampling.py:82
        mod.three_interpolate_grad_wrapper(B, C, N, M, go3, nidx, w, gk)
        r["interp_grad"] = gk
        torch.cuda.synchronize()
        res[name] = r
    a, b = res["geot"], res["ref"]
    for k in ("fps", "ball", "nn_idx"):
        assert torch.equal(a[k], b[k]), k
    for k in ("gather", "group", "nn_d2"):
        assert torch.equal(a[k], b[k]), k
    for k in ("group_grad", "gather_grad", "interp", "interp_grad"):
        _close(a[k], b[k], k)


def test_pointops_cuda_all_thirteen(rb):
    B, n_per, ns, c, w_c = 3, [3000, 4100, 2500], 12, 32, 8
    g = torch.Generator(device="cuda").manual_seed(5)
    xyz = torch.cat([torch.from_numpy(synth.tooth_cloud(n, seed=40 + i)).cuda() for i, n in enumerate(n_per)]).contiguous()
    n = xyz.shape[0]
    offset = torch.tensor(np.cumsum(n_per), dtype=torch.int32, device="cuda")
    m_per = [k // 4 for k in n_per]
    new_offset = torch.tensor(np.cumsum(m_per), dtype=torch.int32, device="cuda")
    m = int(new_offset[-1])
    feat = torch.randn((n, c), device="cuda", generator=g)
    weights = torch.rand((n,), device="cuda", generator=g) + 0.1
    res = {}
    for name, mod in (("geot", backends.pointops_cuda), ("ref", rb.pointops_cuda)):
        r = {}
        n_max = offset[0]                                                    # a 0-dim CUDA tensor, pointops.py:69-72
        for i in range(1, B):
            n_max = max(offset[i] - offset[i - 1], n_max)
        idx = torch.cuda.IntTensor(m).zero_()
        tmp = torch.cuda.FloatTensor(n).fill_(1e10)
        mod.furthestsampling_cuda(B, n_max, xyz, offset, new_offset, tmp, idx)
        r["fps"] = idx
        idxw = torch.cuda.IntTensor(m).zero_()
        tmp = torch.cuda.FloatTensor(n).fill_(1e10)
        mod.furthestsampling_weights_cuda(B, n_max, xyz, offset, new_offset, weights, tmp, idxw)
        r["fps_w"] = idxw
        new_xyz = xyz[idx.long()].contiguous()
        kidx = torch.cuda.IntTensor(m, ns).zero_()
        kd2 = torch.cuda.FloatTensor(m, ns).zero_()
        mod.knnquery_cuda(m, ns, xyz, new_xyz, offset, new_offset, kidx, kd2)
        r["knn"], r["knn_d2"] = kidx, kd2
        bidx = torch.cuda.IntTensor(m, ns).zero_()
        mod.ballquery_cuda(m, 0.15, ns, xyz, new_xyz, offset, new_offset, bidx)
        r["ball"] = bidx
        out = torch.cuda.FloatTensor(m, ns, c)
        mod.grouping_forward_cuda(m, ns, c, feat, kidx, out)
        r["group"] = out
        go = torch.randn((m, ns, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(6))
        gi = torch.cuda.FloatTensor(n, c).zero_()
        mod.grouping_backward_cuda(m, ns, c, go, kidx, gi)
        r["group_grad"] = gi
        # interpolation: known = the m sampled points, unknown = all n points, k = 3
        i3 = torch.cuda.IntTensor(n, 3).zero_()
        d3 = torch.cuda.FloatTensor(n, 3).zero_()
        mod.knnquery_cuda(n, 3, new_xyz, xyz, new_offset, offset, i3, d3)
        r["knn3"] = i3
        rec = 1.0 / (torch.sqrt(d3) + 1e-8)
        w3 = (rec / rec.sum(1, keepdim=True)).contiguous()
        fm = torch.randn((m, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(7))
        o = torch.cuda.FloatTensor(n, c).zero_()
        mod.interpolation_forward_cuda(n, c, 3, fm, i3, w3, o)
        r["interp"] = o
        go2 = torch.randn((n, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(8))
        gm = torch.cuda.FloatTensor(m, c).zero_()
        mod.interpolation_backward_cuda(n, c, 3, go2, i3, w3, gm)
        r["interp_grad"] = gm
        # vector-attention ops on the n points with their own ns neighbours
        sidx = torch.cuda.IntTensor(n, ns).zero_()
        sd = torch.cuda.FloatTensor(n, ns).zero_()
        mod.knnquery_cuda(n, ns, xyz, xyz, offset, offset, sidx, sd)
        in2 = torch.randn((n, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(9))
        so = torch.cuda.FloatTensor(n, ns, c).zero_()
        mod.subtraction_forward_cuda(n, ns, c, feat, in2, sidx, so)
        r["sub"] = so
        gso = torch.randn((n, ns, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(10))
        g1 = torch.cuda.FloatTensor(n, c).zero_()
        g2 = torch.cuda.FloatTensor(n, c).zero_()
        mod.subtraction_backward_cuda(n, ns, c, sidx, gso, g1, g2)
        r["sub_g1"], r["sub_g2"] = g1, g2
        pos = torch.randn((n, ns, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(11))
        aw = torch.randn((n, ns, w_c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(12))
        ao = torch.cuda.FloatTensor(n, c).zero_()
        mod.aggregation_forward_cuda(n, ns, c, w_c, feat, pos, aw, sidx, ao)
        r["agg"] = ao
        gao = torch.randn((n, c), device="cuda", generator=torch.Generator(device="cuda").manual_seed(13))
        gin = torch.cuda.FloatTensor(n, c).zero_()
        gpos = torch.cuda.FloatTensor(n, ns, c).zero_()
        gw = torch.cuda.FloatTensor(n, ns, w_c).zero_()
        mod.aggregation_backward_cuda(n, ns, c, w_c, feat, pos, aw, sidx, gao, gin, gpos, gw)
        r["agg_gi"], r["agg_gp"], r["agg_gw"] = gin, gpos, gw
        torch.cuda.synchronize()
        res[name] = r
    a, b = res["geot"], res["ref"]
    for k in ("fps", "fps_w", "knn", "ball", "knn3"):
        assert torch.equal(a[k], b[k]), k
    for k in ("knn_d2", "group", "sub"):
        assert torch.equal(a[k], b[k]), k
    for k in ("group_grad", "interp", "interp_grad", "sub_g1", "sub_g2", "agg", "agg_gi", "agg_gp"):
        _close(a[k], b[k], k)
    # grad_weight: each entry is a c/w_c-term dot product; bound it against the magnitude of its terms
    d = (a["agg_gw"].double() - b["agg_gw"].double()).abs().max().item()
    assert d <= 1e-5 * b["agg_gw"].abs().max().item(), d


def test_pointnet2_ext_all_nine(rb):
    B, N, M, ns, C = 2, 6000, 500, 16, 21
    g = torch.Generator(device="cuda").manual_seed(20)
    xyz = _cloud(B, N, seed=7)
    feat = torch.randn((B, C, N), device="cuda", generator=g)
    res = {}
    for name, mod in (("geot", backends.pointnet2_ext), ("ref", rb.pointnet2_ext)):
        r = {}
        idx = mod.furthest_point_sampling(xyz, M)
        r["fps"] = idx
        cen = mod.gather_points(xyz.transpose(1, 2).contiguous(), idx)
        r["gather"] = cen
        go = torch.randn((B, 3, M), device="cuda", generator=torch.Generator(device="cuda").manual_seed(21))
        r["gather_grad"] = mod.gather_points_grad(go, idx, N)
        new_xyz = cen.transpose(1, 2).contiguous()
        bidx = mod.ball_query(new_xyz, xyz, 0.1, ns)
        r["ball"] = bidx
        r["group"] = mod.group_points(feat, bidx)
        go2 = torch.randn((B, C, M, ns), device="cuda", generator=torch.Generator(device="cuda").manual_seed(22))
        r["group_grad"] = mod.group_points_grad(go2, bidx, N)
        d2, nidx = mod.three_nn(xyz, new_xyz)
        r["nn_d2"], r["nn_idx"] = d2, nidx
        w = 1.0 / (d2 + 1e-8)
        w = (w / w.sum(2, keepdim=True)).contiguous()
        known = torch.randn((B, C, M), device="cuda", generator=torch.Generator(device="cuda").manual_seed(23))
        r["interp"] = mod.three_interpolate(known, nidx, w)
        go3 = torch.randn((B, C, N), device="cuda", generator=torch.Generator(device="cuda").manual_seed(24))
        r["interp_grad"] = mod.three_interpolate_grad(go3, nidx, w, M)
        torch.cuda.synchronize()
        res[name] = r
    a, b = res["geot"], res["ref"]
    for k in ("fps", "ball", "nn_idx", "gather", "group", "nn_d2"):
        assert torch.equal(a[k], b[k]), k
    for k in ("gather_grad", "group_grad", "interp", "interp_grad"):
        _close(a[k], b[k], k)
    # the reference's dtype / device checks (include/utils.h:8-28) raise; so do the facades
    with pytest.raises(RuntimeError):
        backends.pointnet2_ext.group_points(feat.double(), res["geot"]["ball"])
    with pytest.raises(RuntimeError):
        backends.pointnet2_ext.three_nn(xyz.cpu(), xyz.cpu())
