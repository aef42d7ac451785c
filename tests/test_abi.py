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


def test_no_cpu_fallback():
    import torch
    from geot_b200 import _lib, layers
    with pytest.raises(_lib.GeotError):
        layers.furthest_point_sample(torch.zeros(1, 8, 3), 2)
    with pytest.raises(_lib.GeotError):
        layers.ball_query(0.1, 4, torch.zeros(1, 8, 3), torch.zeros(1, 2, 3))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "geot_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "libgeot_oracle" not in txt and "oracle/_ref" not in txt, f
