"""geot_b200.compat: the reference's own op-API files import against the facades (authoring container only —
/root/reference does not exist on the GPU box, where this test skips)."""
import importlib.util
import os
import sys
import types

import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_reference_op_files_import_on_the_facades():
    import geot_b200.compat as compat
    from geot_b200 import backends
    saved = dict(sys.modules)
    try:
        compat.install()
        # bare package parents so `from openpoints.cpp[.pointnet2_batch] import pointnet2_cuda` resolves without
        # executing openpoints/__init__.py (which drags in datasets, timm, multimethod ...)
        op = types.ModuleType("openpoints"); op.__path__ = []
        cpp = types.ModuleType("openpoints.cpp"); cpp.__path__ = []
        cpp.pointnet2_cuda = backends.pointnet2_batch_cuda
        sys.modules["openpoints"], sys.modules["openpoints.cpp"] = op, cpp
        g = _load("ref_group", f"{REF}/openpoints/models/layers/group.py")
        s = _load("ref_subsample", f"{REF}/openpoints/models/layers/subsample.py")
        po = _load("ref_pointops", f"{REF}/pointops/functions/pointops.py")
        pof = _load("ref_pointops_full", f"{REF}/openpoints/cpp/pointops/functions/pointops.py")
        assert g.pointnet2_cuda is backends.pointnet2_batch_cuda and s.pointnet2_cuda is backends.pointnet2_batch_cuda
        assert po.pointops_cuda is backends.pointops_cuda and pof.pointops_cuda is backends.pointops_cuda
        # every callable those files use exists on the facades
        for fn in ("ball_query_wrapper", "group_points_wrapper", "group_points_grad_wrapper", "gather_points_wrapper",
                   "gather_points_grad_wrapper", "furthest_point_sampling_wrapper", "three_nn_wrapper",
                   "three_interpolate_wrapper", "three_interpolate_grad_wrapper"):
            assert callable(getattr(backends.pointnet2_batch_cuda, fn))
        import re
        used = set(re.findall(r"pointops_cuda\.(\w+)", open(f"{REF}/openpoints/cpp/pointops/functions/pointops.py").read()))
        used |= set(re.findall(r"pointops_cuda\.(\w+)", open(f"{REF}/pointops/functions/pointops.py").read()))
        assert used and all(callable(getattr(backends.pointops_cuda, u)) for u in used), used
        used = set(re.findall(r"_ext\.(\w+)", open(f"{REF}/pointnet2/pointnet2_utils.py").read()))
        assert used and all(callable(getattr(backends.pointnet2_ext, u)) for u in used), used
        import knn_cuda
        assert knn_cuda.KNN(4, transpose_mode=True).k == 4
        from pointnet2_ops import pointnet2_utils as p2u
        assert callable(p2u.three_nn) and callable(p2u.three_interpolate)
        # the reference's groupers construct on top of our ops
        assert isinstance(g.create_grouper({"NAME": "ballquery", "radius": 0.1, "nsample": 16}), g.QueryAndGroup)
    finally:
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        sys.modules.update(saved)
