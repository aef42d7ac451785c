"""Environment for running the reference's OWN Python (models, layers, op wrappers, losses) unchanged.

Test infrastructure (tests/, bench.py --config cfgN only).  What it does:

* puts the reference tree on sys.path: /root/reference here, oracle/_ref/src (staged by tools/stage_reference.sh,
  git-ignored, travels with gpurun) on the GPU box;
* installs stand-ins for the third-party imports the reference needs at import time but that are not arithmetic
  on the hot path: timm DropPath/Mlp, easydict, multimethod, termcolor, shortuuid,
  torch_scatter, h5py/open3d/trimesh/pyvista/matplotlib (import-time only);
* binds the three extension-module names the reference imports (`pointnet2_batch_cuda`, `pointops_cuda`,
  `pointnet2._ext`) plus `knn_cuda` / `pointnet2_ops` to a SWITCHABLE proxy, so the SAME model object can run
  once on geot_b200 (`use("geot")` -> geot_b200.backends, what compat.install() registers) and once on the
  reference's own kernels (`use("ref")` -> tests/ref_backends.py over oracle/_ref/*.so).
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_CANDIDATES = ["/root/reference", os.path.join(ROOT, "oracle", "_ref", "src")]

_state = {"backend": None, "installed": False, "root": None, "record": None, "timing": None}

# index-producing entry points: where the int32 result lives (argument position, or the return value)
_INDEX_OPS = {
    "ball_query_wrapper": 7, "furthest_point_sampling_wrapper": 5, "three_nn_wrapper": 6,          # pointnet2_batch_cuda
    "knnquery_cuda": 6, "furthestsampling_cuda": 6, "furthestsampling_weights_cuda": 7, "ballquery_cuda": 7,  # pointops_cuda
    "furthest_point_sampling": "ret", "three_nn": "ret1", "ball_query": "ret",                       # pointnet2._ext
}


def ref_root():
    for c in _CANDIDATES:
        if os.path.isdir(os.path.join(c, "openpoints", "models")):
            return c
    return None


def available():
    return ref_root() is not None


class _Switch:
    """Attribute lookups go to the currently selected backend's object of the same role."""

    def __init__(self, role):
        object.__setattr__(self, "_role", role)

    def __getattr__(self, name):
        b = _state["backend"]
        if b is None:
            raise RuntimeError("refenv.use('geot'|'ref') first")
        fn = getattr(_backend_objects(b)[self._role], name)
        if (_state["record"] is None and _state["timing"] is None) or not callable(fn):
            return fn
        return _instrument(name, fn)


def _instrument(name, fn):
    """Wrap one facade call: clone index results into the recording, CUDA-event time it into the timing table."""
    import torch

    def call(*a, **k):
        tm = _state["timing"]
        if tm is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        r = fn(*a, **k)
        if tm is not None:
            e1.record()
            tm.setdefault(name, []).append((e0, e1))
        rec, where = _state["record"], _INDEX_OPS.get(name)
        if rec is not None and where is not None:
            t = r if where == "ret" else (r[1] if where == "ret1" else a[where])
            rec.append((name, t.detach().clone()))
        return r
    return call


@contextlib.contextmanager
def recording():
    """`with refenv.recording() as rec:` collects (entry point, index tensor) for every index-producing call."""
    prev, _state["record"] = _state["record"], []
    try:
        yield _state["record"]
    finally:
        _state["record"] = prev


@contextlib.contextmanager
def timing():
    """`with refenv.timing() as tm:` -> after the block (and a device sync) tm[name] = [ms, ...] per facade call."""
    prev, _state["timing"] = _state["timing"], {}
    tm = _state["timing"]
    try:
        yield tm
    finally:
        _state["timing"] = prev
        import torch
        torch.cuda.synchronize()
        for k, v in tm.items():
            tm[k] = [a.elapsed_time(b) for a, b in v]


_cache = {}


def _backend_objects(name):
    if name not in _cache:
        if name == "geot":
            from geot_b200 import backends, compat
            from geot_b200 import pointnet2_utils as p2u
            _cache[name] = {"pn2batch": backends.pointnet2_batch_cuda, "pointops": backends.pointops_cuda,
                            "pn2ext": backends.pointnet2_ext, "knn": compat, "p2ops": p2u}
        elif name == "ref":
            from tests import ref_backends as rb
            _cache[name] = {"pn2batch": rb.pointnet2_batch_cuda, "pointops": rb.pointops_cuda,
                            "pn2ext": rb.pointnet2_ext, "knn": rb, "p2ops": rb.pointnet2_ops_utils}
        else:
            raise ValueError(name)
    return _cache[name]


def use(name):
    assert name in ("geot", "ref")
    _backend_objects(name)
    _state["backend"] = name


def backend():
    return _state["backend"]


@contextlib.contextmanager
def using(name):
    prev = _state["backend"]
    use(name)
    try:
        yield
    finally:
        _state["backend"] = prev


# ---------------------------------------------------------------------------------------------- stand-ins
def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []          # allow `import a.b`
    sys.modules[name] = m
    return m


def _install_standins():
    import torch
    import torch.nn as nn

    def _have(name):
        try:
            __import__(name)
            return True
        except Exception:
            return False

    if not _have("timm"):
        class DropPath(nn.Module):                     # timm.models.layers.DropPath (stochastic depth)
            def __init__(self, drop_prob=0.0, scale_by_keep=True):
                super().__init__()
                self.drop_prob, self.scale_by_keep = drop_prob, scale_by_keep

            def forward(self, x):
                if self.drop_prob == 0.0 or not self.training:
                    return x
                keep = 1 - self.drop_prob
                mask = x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(keep)
                if keep > 0.0 and self.scale_by_keep:
                    mask.div_(keep)
                return x * mask

        def trunc_normal_(t, mean=0.0, std=1.0, a=-2.0, b=2.0):
            return nn.init.trunc_normal_(t, mean, std, a, b)

        class Mlp(nn.Module):                          # timm.models.vision_transformer.Mlp
            def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.0):
                super().__init__()
                out_features = out_features or in_features
                hidden_features = hidden_features or in_features
                self.fc1 = nn.Linear(in_features, hidden_features)
                self.act = act_layer()
                self.fc2 = nn.Linear(hidden_features, out_features)
                self.drop = nn.Dropout(drop)

            def forward(self, x):
                return self.drop(self.fc2(self.drop(self.act(self.fc1(x)))))

        _mod("timm")
        _mod("timm.models")
        _mod("timm.models.layers", DropPath=DropPath, trunc_normal_=trunc_normal_)
        _mod("timm.models.vision_transformer", Mlp=Mlp)

    if not _have("easydict"):
        class EasyDict(dict):
            def __init__(self, d=None, **kw):
                super().__init__()
                for k, v in dict(d or {}, **kw).items():
                    setattr(self, k, v)

            def __setattr__(self, k, v):
                if isinstance(v, dict) and not isinstance(v, EasyDict):
                    v = EasyDict(v)
                super().__setitem__(k, v)

            __setitem__ = __setattr__

            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError:
                    raise AttributeError(k)
        _mod("easydict", EasyDict=EasyDict)

    if not _have("multimethod"):
        import typing

        class multimethod:                             # dispatch on the type of the first non-self argument
            def __init__(self, fn):
                self._impls = []
                self._name = fn.__name__
                self.register(fn)

            def register(self, fn):
                hints = [h for k, h in typing.get_type_hints(fn).items() if k != "return"]
                first = hints[0] if hints else object
                origin = typing.get_origin(first)
                if origin is typing.Union:
                    types_ = tuple(typing.get_origin(a) or a for a in typing.get_args(first))
                else:
                    types_ = (origin or first,)
                self._impls.insert(0, (types_, fn))
                return self

            def __get__(self, obj, cls=None):
                if obj is None:
                    return self
                return types.MethodType(self, obj)

            def __call__(self, obj, arg, *a, **kw):
                for types_, fn in self._impls:
                    if isinstance(arg, types_):
                        return fn(obj, arg, *a, **kw)
                raise TypeError(f"{self._name}: no implementation for {type(arg)}")

        _registry = {}

        def mm(fn):
            key = (fn.__module__, fn.__qualname__)
            if key in _registry:
                return _registry[key].register(fn)
            _registry[key] = multimethod(fn)
            return _registry[key]
        _mod("multimethod", multimethod=mm)

    if not _have("termcolor"):
        _mod("termcolor", colored=lambda s, *a, **k: s)
    if not _have("shortuuid"):
        import uuid
        _mod("shortuuid", uuid=lambda: uuid.uuid4().hex[:22])
    if not _have("torch_scatter"):
        def scatter(*a, **k):
            raise NotImplementedError("torch_scatter is an import-time stand-in (unused on the hot path)")
        _mod("torch_scatter", scatter=scatter)
    for name in ("h5py", "open3d", "trimesh", "pyvista"):
        if not _have(name):
            _mod(name)
    if not _have("matplotlib"):
        _mod("matplotlib", use=lambda *a, **k: None)
        _mod("matplotlib.pyplot")
        _mod("matplotlib.cm")
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
        sys.modules["matplotlib"].cm = sys.modules["matplotlib.cm"]


def install(backend="geot"):
    """Idempotent.  After this `import openpoints.models`, `pointops.functions.pointops`, `pointnet2.pointnet2_utils`,
    `utils.insT_loss` ... resolve to the reference's own files on top of the switchable backend."""
    root = ref_root()
    if root is None:
        raise RuntimeError("no reference tree: run tools/stage_reference.sh where /root/reference exists")
    if not _state["installed"]:
        _install_standins()
        for p in (os.path.join(root, "pointnet2"), root):   # pointnet2_utils does a bare `import pytorch_utils`
            if p not in sys.path:
                sys.path.insert(0, p)
        pn2b, po, ext = _Switch("pn2batch"), _Switch("pointops"), _Switch("pn2ext")
        sys.modules["pointnet2_batch_cuda"] = pn2b
        sys.modules["pointops_cuda"] = po
        sys.modules["pointnet2._ext"] = ext
        knn = types.ModuleType("knn_cuda")

        def KNN(k, transpose_mode=False):
            return _KNNSwitch(k, transpose_mode)
        knn.KNN = KNN
        sys.modules["knn_cuda"] = knn
        p2o = types.ModuleType("pointnet2_ops")
        p2o.__path__ = []
        p2o.pointnet2_utils = _Switch("p2ops")
        sys.modules["pointnet2_ops"] = p2o
        sys.modules["pointnet2_ops.pointnet2_utils"] = p2o.pointnet2_utils
        _state["installed"] = True
        _state["root"] = root
    use(backend)
    return root


def _make_knn_switch():
    import torch.nn as nn

    class KNNSwitch(nn.Module):
        """knn_cuda.KNN stand-in that resolves the implementation at call time (geot kernel / float64 definition)."""

        def __init__(self, k, transpose_mode=False):
            super().__init__()
            self.k, self._t = k, transpose_mode
            self._impl = {}

        def forward(self, ref, query):
            b = _state["backend"]
            if b not in self._impl:
                self._impl[b] = _backend_objects(b)["knn"].KNN(self.k, transpose_mode=self._t)
            r = self._impl[b](ref, query)
            if _state["record"] is not None:
                _state["record"].append(("knn_cuda.KNN", r[1].detach().clone()))
            return r
    return KNNSwitch


def _KNNSwitch(k, t):
    global _KNN_CLS
    try:
        cls = _KNN_CLS
    except NameError:
        cls = _KNN_CLS = _make_knn_switch()
    return cls(k, t)
