"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol that
include/geot_b200.h declares (no compute calls — there is no GPU here), and the product has no fallback."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "geot_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(geot_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_symbols():
    syms = _declared_symbols()
    assert "geot_furthest_point_sampling" in syms and "geot_ball_query" in syms and len(syms) >= 25


def test_library_exports_every_declared_symbol():
    from geot_b200 import _lib
    if not os.path.exists(_lib.SO_PATH):
        _lib.build()
    L = ctypes.CDLL(_lib.SO_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(L, s)]
    assert not missing, f"declared in include/geot_b200.h but not exported: {missing}"


def test_python_signatures_cover_header():
    from geot_b200 import _lib
    assert sorted(_lib._SIGNATURES) == _declared_symbols()


def test_introspection_calls_work_without_gpu():
    from geot_b200 import _lib
    L = _lib.lib()
    assert L.geot_abi_version() == 1
    assert isinstance(_lib.launch_count(), int)
    assert L.geot_scatter_plan_bytes(2, 100, 50) >= 2 * 101 * 4 + 2 * 50 * 4


def test_argument_validation_returns_error_not_exit():
    from geot_b200 import _lib
    L = _lib.lib()
    rc = L.geot_furthest_point_sampling(1, 16, 4, None, None, None, 0, None)
    assert rc == -1 and b"NULL" in L.geot_last_error()
    rc = L.geot_knnquery(1, 4, 4, 0, None, None, None, None, 4, 4, None, None, None, 0, 0, None)
    assert rc == -1 and b"nsample" in L.geot_last_error()


def test_flag_constants_match_the_header():
    """Every integer #define GEOT_* of the header that the Python host mirrors has the same value there."""
    from geot_b200 import _lib
    src = open(os.path.join(ROOT, "include", "geot_b200.h")).read()
    defs = dict(re.findall(r"^#define\s+(GEOT_[A-Z0-9_]+)\s+(-?\d+)\b", src, flags=re.M))
    assert {"GEOT_FPS_CAP512", "GEOT_BQ_FORCE_GRID", "GEOT_BQ_BUILD_ONLY", "GEOT_BQ_PREBUILT", "GEOT_KNN_TIES_BY_INDEX"} <= set(defs)
    mirrored = [n for n in defs if hasattr(_lib, n) and isinstance(getattr(_lib, n), int)]
    assert len(mirrored) >= 8
    for n in mirrored:
        assert getattr(_lib, n) == int(defs[n]), n


def test_ball_query_split_flags_are_validated_before_any_launch():
    from geot_b200 import _lib
    L = _lib.lib()
    both = _lib.GEOT_BQ_BUILD_ONLY | _lib.GEOT_BQ_PREBUILT
    rc = L.geot_ball_query(1, 8192, 4, ctypes.c_float(0.1), 8, None, None, None, None, 0, both, None)
    assert rc == -1 and b"exclude" in L.geot_last_error()
    rc = L.geot_ball_query(1, 8192, 0, ctypes.c_float(0.1), 8, None, ctypes.c_void_p(256), None, None, 0, _lib.GEOT_BQ_BUILD_ONLY, None)
    assert rc == -1 and b"workspace" in L.geot_last_error()


def test_no_cpu_fallback():
    import torch
    from geot_b200 import _lib, layers
    with pytest.raises(_lib.GeotError):
        layers.furthest_point_sample(torch.zeros(1, 8, 3), 2)
    with pytest.raises(_lib.GeotError):
        layers.ball_query(0.1, 4, torch.zeros(1, 8, 3), torch.zeros(1, 2, 3))
    from geot_b200 import functional as F
    with pytest.raises(_lib.GeotError):
        F.ball_query_prepare(torch.zeros(1, 8, 3), 0.1, 4)
    with pytest.raises(_lib.GeotError):
        F.fps_ball_query(torch.zeros(1, 8, 3), 2, 0.1, 4)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "geot_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "libgeot_oracle" not in txt and "oracle/_ref" not in txt, f
