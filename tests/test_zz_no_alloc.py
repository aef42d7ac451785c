"""geot_ball_query_static: the allocation-free ball query (scratch in a device array built into the library) against
geot_ball_query with a caller workspace, through the C ABI.  Indices, so the bar is bit-equality (the tolerance every
ball-query test here uses).  Named test_zz_* so pytest collects it last.

Covers: two calls in a row on different inputs (the array keeps the first call's bins - the second call must not see
them), the first input again, a second device's own copy of the array when there is one, the stated limit
(GEOT_ERR_WORKSPACE, nothing launched, output untouched) and that a warm call changes neither torch's nor the
driver's view of free device memory.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from geot_b200 import _lib, synth  # noqa: E402

STATIC_BYTES = 8 << 20      # GEOT_BALL_QUERY_STATIC_SCRATCH_BYTES (include/geot_b200.h)


def _inputs(B, N, M, first_cloud, device):
    xyz = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=first_cloud)).to(device)
    pick = torch.from_numpy(np.random.default_rng(B * 1000 + N).choice(N, M, replace=False)).to(device)
    return xyz, xyz[:, pick].contiguous()


def _with_workspace(xyz, new_xyz, radius, nsample, flags=0):
    B, N, M = xyz.shape[0], xyz.shape[1], new_xyz.shape[1]
    nbytes = int(_lib.lib().geot_ball_query_workspace_bytes(B, N))
    assert 0 < nbytes <= STATIC_BYTES
    ws = torch.empty(nbytes, dtype=torch.uint8, device=xyz.device)
    idx = torch.full((B, M, nsample), -7, dtype=torch.int32, device=xyz.device)
    _lib.call("geot_ball_query", B, N, M, float(radius), nsample, _lib.ptr(new_xyz), _lib.ptr(xyz), _lib.ptr(idx),
              _lib.ptr(ws), nbytes, flags, _lib.stream_ptr())
    return idx


def _static(xyz, new_xyz, radius, nsample, idx, flags=0):
    B, N, M = xyz.shape[0], xyz.shape[1], new_xyz.shape[1]
    _lib.call("geot_ball_query_static", B, N, M, float(radius), nsample, _lib.ptr(new_xyz), _lib.ptr(xyz), _lib.ptr(idx),
              flags, _lib.stream_ptr())
    return idx


CASES = [  # B, N, M, radius, nsample, first_cloud, flags     (flags 1 = GEOT_BQ_FORCE_GRID: the scratch is really read)
    (16, 16384, 4096, 0.05, 32, 0, 0),      # BASELINE config 3, stage 1: the largest shape the array is sized for
    (3, 5000, 700, 0.2, 16, 40, 1),         # fewer, smaller clouds: stale bins of the first call lie beyond and between
    (7, 65536, 512, 0.03, 64, 80, 1),       # N at the binned path's limit
    (16, 16384, 4096, 0.05, 32, 0, 1),      # the first input again
]


def test_static_scratch_ball_query_equals_workspace_ball_query(cuda):
    dev = cuda
    prepared = []
    for B, N, M, r, ns, first, flags in CASES:
        xyz, q = _inputs(B, N, M, first, dev)
        prepared.append((xyz, q, r, ns, flags, _with_workspace(xyz, q, r, ns, flags),
                         torch.full((B, M, ns), -7, dtype=torch.int32, device=dev)))
    torch.cuda.synchronize()
    for xyz, q, r, ns, flags, expect, out in prepared:          # back to back on one stream, nothing in between
        _static(xyz, q, r, ns, out, flags)
    torch.cuda.synchronize()
    for i, (xyz, q, r, ns, flags, expect, out) in enumerate(prepared):
        assert int(expect.min()) >= 0
        assert torch.equal(out, expect), f"case {i}: {(out != expect).sum().item()} of {out.numel()} indices differ"
    # a warm call allocates nothing: neither torch's allocator nor the driver's free-memory figure moves
    xyz, q, r, ns, flags, expect, out = prepared[1]
    out.fill_(-7)
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev), torch.cuda.mem_get_info(dev)[0])
    _static(xyz, q, r, ns, out, flags)
    torch.cuda.synchronize()
    after = (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev), torch.cuda.mem_get_info(dev)[0])
    assert after == before and torch.equal(out, expect)


def test_static_scratch_ball_query_rejects_sizes_beyond_the_array(cuda):
    xyz, q = _inputs(2, 1024, 64, 0, cuda)
    out = torch.full((2, 64, 8), -7, dtype=torch.int32, device=cuda)
    ws_bytes = _lib.lib().geot_ball_query_workspace_bytes
    assert int(ws_bytes(21, 16384)) <= STATIC_BYTES < int(ws_bytes(22, 16384))
    for B, N in ((22, 16384), (8, 65536), (1, 65537)):
        with pytest.raises(_lib.GeotError, match=r"ball_query_static.*\(-3\)|\(-3\).*ball_query_static"):
            _lib.call("geot_ball_query_static", B, N, 64, 0.1, 8, _lib.ptr(q), _lib.ptr(xyz), _lib.ptr(out), 0, _lib.stream_ptr())
    torch.cuda.synchronize()
    assert bool((out == -7).all())                               # nothing was launched
    _lib.call("geot_ball_query_static", 0, 1024, 64, 0.1, 8, None, None, None, 0, _lib.stream_ptr())   # empty batch: no-op
    _static(xyz, q, 0.1, 8, out)                                 # and the entry point still works after the refusals
    assert torch.equal(out, _with_workspace(xyz, q, 0.1, 8))


def test_static_scratch_ball_query_on_a_second_device(cuda):
    """Each GPU holds its own copy of the array and the address is looked up per device."""
    if torch.cuda.device_count() < 2:
        pytest.skip("one device")
    try:
        for d in (1, 0, 1):
            dev = torch.device("cuda", d)
            torch.cuda.set_device(dev)
            xyz, q = _inputs(4, 8192, 1024, 10 * d, dev)
            out = torch.full((4, 1024, 24), -7, dtype=torch.int32, device=dev)
            _static(xyz, q, 0.08, 24, out, 1)
            assert torch.equal(out, _with_workspace(xyz, q, 0.08, 24, 1))
    finally:
        torch.cuda.set_device(0)
