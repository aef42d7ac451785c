"""ctypes loader for libgeot_b200.so (include/geot_b200.h) — stream, device and error plumbing.

PyTorch is used for device memory and streams only.  There is NO fallback: if the CUDA library is
missing or a call fails this raises; nothing here (or anywhere in geot_b200/) imports oracle/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "_C", "libgeot_b200.so")
_lib = None

GEOT_FPS_CAP512 = 1
GEOT_FPS_MAGSKIP = 2
GEOT_FPS_FORCE_GENERIC = 4
GEOT_FPS_NO_PRUNE = 8
GEOT_FPS_SPEC = 16
GEOT_BQ_FORCE_GRID = 1
GEOT_BQ_BUILD_ONLY = 2
GEOT_BQ_PREBUILT = 4


def GEOT_BQ_LEVELS(n):
    """testing / A-B: at most n (1, 2, 4, 8) index levels in the ball-query grid (include/geot_b200.h)."""
    return (n & 15) << 8


GEOT_KNN_NO_GRID = 1
GEOT_KNN_TIES_BY_INDEX = 2

_vp, _i, _f, _i64, _sz = C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_size_t

# name -> argtypes (every function returns int unless listed in _RESTYPES)
_SIGNATURES = {
    "geot_abi_version": [],
    "geot_last_error": [],
    "geot_launch_count": [],
    "geot_furthest_point_sampling": [_i, _i, _i, _vp, _vp, _vp, _i, _vp],
    "geot_furthestsampling": [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "geot_ball_query_workspace_bytes": [_i, _i],
    "geot_ball_query": [_i, _i, _i, _f, _i, _vp, _vp, _vp, _vp, _sz, _i, _vp],
    "geot_ball_query_static": [_i, _i, _i, _f, _i, _vp, _vp, _vp, _i, _vp],
    "geot_ballquery_offset": [_i, _i, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_knnquery_workspace_bytes": [_i, _i, _i],
    "geot_knnquery": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _sz, _i, _vp],
    "geot_group_points": [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_gather_points": [_i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_group_points_config": [_i],
    "geot_gather_rows": [_i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_group_xyz_relative": [_i, _i, _i, _i, _vp, _vp, _vp, _i, _f, _vp, _vp],
    "geot_scatter_plan_bytes": [_i, _i, _i64],
    "geot_scatter_plan_build": [_i, _i, _i64, _vp, _vp, _sz, _vp],
    "geot_group_points_grad": [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_scatter_plan_pieces": [_i, _i64],
    "geot_group_points_grad_pieces": [_i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_gather_points_grad": [_i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_three_nn": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_three_interpolate": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_three_interpolate_grad": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_grouping_forward": [_i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_grouping_backward": [_i, _i, _i, _i, _vp, _vp, _vp, _vp],
    "geot_interpolation_forward": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_interpolation_backward": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_subtraction_forward": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_subtraction_backward": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_aggregation_forward": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_aggregation_backward": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_sigt_workspace_bytes": [_i],
    "geot_sigt_forward": [_i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _sz, _vp],
    "geot_sigt_backward_workspace_bytes": [_i, _i, _i],
    "geot_sigt_backward": [_i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp],
    "geot_transition_apply_forward": [_i, _i, _i, _vp, _vp, _vp, _f, _vp, _vp],
    "geot_transition_apply_backward": [_i, _i, _i, _vp, _vp, _vp, _f, _vp, _vp, _vp, _vp],
    "geot_manifold_workspace_bytes": [_i64],
    "geot_poly1_focal_workspace_bytes": [_i64],
    "geot_poly1_focal_forward": [_i, _i, _i, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp, _vp, _vp, _sz, _vp],
    "geot_poly1_focal_backward": [_i, _i, _i, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp, _vp, _vp, _vp],
    "geot_poly1_focal_t_forward": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp, _vp, _vp, _sz, _vp],
    "geot_poly1_focal_t_backward": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _f, _vp, _vp, _vp, _vp, _vp],
    "geot_spatial_order": [_i, _i, _i, _vp, _vp, _vp],
    "geot_manifold_loss_forward": [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _f, _i, _vp, _vp, _vp, _sz, _vp],
    "geot_manifold_loss_backward": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_class_pick_workspace_bytes": [_i],
    "geot_class_pick": [_i, _i, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp],
    "geot_class_t_step": [_i, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp],
    "geot_mean_feature_update": [_i64, _i, _vp, _vp, _vp, _vp, _vp, _sz, _vp],
    "geot_graph_feature": [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "geot_graph_feature_grad_q": [_i, _i, _i, _i, _vp, _vp, _vp],
    "geot_interp_weights": [_i64, _vp, _vp, _vp, _vp],
    "geot_interp_argmax": [_i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_seg_counts": [_i64, _i, _vp, _vp, _vp, _vp],
    "geot_stats_record_floats": [_i, _i64],
    "geot_stats_pack": [_i, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "geot_stats_combine": [_i, _i, _i64, _vp, _f, _vp, _vp, _vp, _vp, _vp],
}
_RESTYPES = {
    "geot_last_error": C.c_char_p,
    "geot_launch_count": C.c_uint64,
    "geot_scatter_plan_bytes": _sz,
    "geot_ball_query_workspace_bytes": _sz,
    "geot_group_points_config": None,
    "geot_knnquery_workspace_bytes": _sz,
    "geot_poly1_focal_workspace_bytes": _sz,
    "geot_sigt_workspace_bytes": _sz,
    "geot_sigt_backward_workspace_bytes": _sz,
    "geot_manifold_workspace_bytes": _sz,
    "geot_class_pick_workspace_bytes": _sz,
    "geot_stats_record_floats": _i64,
}


class GeotError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile every kernel for sm_100a into geot_b200/_C/libgeot_b200.so (nvcc, no torch headers)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise GeotError("building libgeot_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return SO_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise GeotError(
                f"{SO_PATH} is missing: the geot_b200 ops have no CPU or PyTorch fallback. "
                "Build it with `python -c 'import __graft_entry__ as g; g.build()'` or `make -C geot_b200/csrc`.")
        L = C.CDLL(SO_PATH)
        for name, argtypes in _SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header/library mismatch: fail loudly
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, C.c_int)
        _lib = L
    return _lib


def launch_count() -> int:
    return int(lib().geot_launch_count())


def ptr(t):
    """Device pointer of a tensor as an int (None -> NULL); ctypes converts it through the declared argtypes."""
    return None if t is None else t.data_ptr()


try:  # raw stream handle without building a torch.cuda.Stream object per call (private but long-stable torch API)
    _raw_stream = torch._C._cuda_getCurrentRawStream
except AttributeError:  # pragma: no cover
    _raw_stream = None


def stream_ptr():
    if _raw_stream is not None:
        return _raw_stream(_cur_dev())
    return torch.cuda.current_stream().cuda_stream


try:  # current device without the lazy-init / Python wrapper cost of torch.cuda.current_device() (0.25 -> 0.05 us)
    _cur_dev = torch._C._cuda_getDevice
except AttributeError:  # pragma: no cover
    _cur_dev = torch.cuda.current_device


class on_device:
    """`with on_device(dev):` = torch.cuda.device(dev) without the two set_device calls when dev is already current
    (the common case; the guard costs microseconds that matter for the 5-microsecond kernels)."""
    __slots__ = ("idx", "prev")

    def __init__(self, dev):
        self.idx = dev.index if (dev is not None and dev.index is not None) else _cur_dev()

    def __enter__(self):
        self.prev = _cur_dev()
        if self.prev != self.idx:
            torch.cuda.set_device(self.idx)

    def __exit__(self, *exc):
        if self.prev != self.idx:
            torch.cuda.set_device(self.prev)
        return False


def check_cuda(*f32, i32=(), i64=()):
    """Every tensor must live on one CUDA device, be contiguous and have the dtype the kernels read it as:
    positional tensors float32, `i32=` int32 (indices, offsets), `i64=` int64 (labels).  Returns the device.
    The C ABI takes raw pointers, so a float64 cloud, half features or int64 indices would otherwise be silently
    reinterpreted; the reference's extensions raise on these too (CHECK_IS_FLOAT / .data<float>(), utils.h:8-28)."""
    dev = None
    for group, want in ((f32, torch.float32), (i32, torch.int32), (i64, torch.int64)):
        for t in group:
            if t is None:
                continue
            if not t.is_cuda:
                raise GeotError("geot_b200 ops are CUDA-only (sm_100a); got a CPU tensor — there is no CPU fallback")
            if t.dtype is not want:
                raise GeotError(f"geot_b200: expected a {want} tensor, got {t.dtype} (shape {tuple(t.shape)}); "
                                "coordinates / features / weights are float32, indices and offsets int32, labels int64")
            if not t.is_contiguous():
                raise GeotError("geot_b200 ops need contiguous tensors (the reference asserts the same, group.py:88-89)")
            if dev is None:
                dev = t.device
            elif t.device != dev:
                raise GeotError(f"tensors on different devices: {dev} vs {t.device}")
    return dev


_fn_cache = {}


def call(name, *args):
    fn = _fn_cache.get(name)
    if fn is None:
        fn = _fn_cache[name] = getattr(lib(), name)
    rc = fn(*args)
    if rc != 0:
        msg = lib().geot_last_error()
        raise GeotError(f"{name} failed ({rc}): {msg.decode() if msg else ''}")
