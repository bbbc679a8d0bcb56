"""Import the LIVE reference modules from ``/root/reference`` (read-only) for validating the
oracle restatement and for generating ``tests/golden/``.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.  Only usable in the build container: the
GPU box has no ``/root/reference``; ``available()`` is False there and every caller skips.

Stubs installed in ``sys.modules`` (all absent from this image / the snapshot):
  matplotlib(+pyplot)  -- ferro_class.py:6 imports it, never uses it on the hot path
  torchdiffeq          -- third-party, absent; ``odeint`` is routed to oracle.solvers.odeint
  kan_diffusion.kan    -- missing git submodule; ``KAN`` := efficientkan.KAN (SURVEY 8c rule)
  Time_MMD.numerical   -- missing git submodule (datasets only)
  efficient_kan        -- the shipped ``__init__`` imports a non-existent ``.kan``; replaced by
                          a stub package whose ``efficientkan`` is the real module + ``KANFET``
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("FETODE_REFERENCE_ROOT", "/root/reference")
_cache = {}


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "efficient_kan", "efficientkan.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _install_stubs():
    if "matplotlib" not in sys.modules:
        try:
            importlib.import_module("matplotlib")
        except Exception:
            class _Anything:
                def __getattr__(self, k):
                    return _Anything()

                def __call__(self, *a, **k):
                    return _Anything()
            def _lazy(k):                               # module-level __getattr__
                if k.startswith("__"):
                    raise AttributeError(k)
                return _Anything()
            mp = _stub("matplotlib", use=lambda *a, **k: None)
            pp = _stub("matplotlib.pyplot")
            pp.__getattr__ = _lazy
            mp.pyplot = pp
            for sub in ("cm", "colors", "ticker", "gridspec", "patches", "lines"):
                s = _stub("matplotlib." + sub)
                s.__getattr__ = _lazy
                setattr(mp, sub, s)
    if "torchdiffeq" not in sys.modules:
        from . import solvers
        _stub("torchdiffeq", odeint=solvers.odeint, odeint_adjoint=solvers.odeint)


def efficientkan():
    """The reference's efficient_kan/efficientkan.py, imported the only way that works
    (sys.path form, train_kanfet_node_predprey.py:11-12)."""
    if "efficientkan" in _cache:
        return _cache["efficientkan"]
    assert available(), "reference checkout not present"
    _install_stubs()
    p = os.path.join(REF_ROOT, "efficient_kan")
    if p not in sys.path:
        sys.path.insert(0, p)
    mod = importlib.import_module("efficientkan")
    _cache["efficientkan"] = mod
    return mod


def ferro_class():
    if "ferro_class" in _cache:
        return _cache["ferro_class"]
    assert available()
    _install_stubs()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    mod = importlib.import_module("ferro_class")
    _cache["ferro_class"] = mod
    return mod


def kanfet_ref_class():
    """``KANFET_ref`` -- the documented reconstruction of the class the reference forgot to
    ship: efficientkan.KAN whose per-layer ``logistic_basis`` slot holds a
    ferro_class.TwoDimensionFerroelectricBasis(in_features, num_basis).  Built purely from
    reference code.  PARITY UNPINNED for the composition itself."""
    ek, fc = efficientkan(), ferro_class()

    class KANFET(ek.KAN):
        def __init__(self, layers_hidden, grid_size=5, spline_order=3, **kw):
            super().__init__(layers_hidden, grid_size=grid_size, spline_order=spline_order, **kw)
            for layer in self.layers:
                layer.logistic_basis = fc.TwoDimensionFerroelectricBasis(layer.in_features, layer.num_basis)

        def reset_state(self):
            for layer in self.layers:
                layer.logistic_basis.reset_state()

    return KANFET


def script_module(name: str):
    """Import one of the ``__main__``-guarded scripts (train_ecg_kan_fet_nn_ode, train_kan_fet_ett,
    train_ecg, mnist_kuramoto_kan, ...) with the extra stubs it needs."""
    if name in _cache:
        return _cache[name]
    _install_stubs()                                    # another test may have removed a stub since the first call
    ek = efficientkan()
    ferro_class()
    if "kan_diffusion" not in sys.modules:
        kd = _stub("kan_diffusion")
        kd.kan = _stub("kan_diffusion.kan", KAN=ek.KAN, KANLinear=ek.KANLinear)
    if "Time_MMD" not in sys.modules:
        tm = _stub("Time_MMD")
        tm.numerical = _stub("Time_MMD.numerical", __all__=[])
    if "efficient_kan" not in sys.modules or not hasattr(sys.modules["efficient_kan"], "efficientkan"):
        pkg = _stub("efficient_kan")
        pkg.__path__ = []
        if not hasattr(ek, "KANFET"):
            ek.KANFET = kanfet_ref_class()
        pkg.efficientkan = ek
        sys.modules["efficient_kan.efficientkan"] = ek
    mod = importlib.import_module(name)
    _cache[name] = mod
    return mod


def params_of(module: torch.nn.Module) -> dict:
    """state_dict (parameters + registered buffers) as a plain dict of detached clones."""
    return {k: v.detach().clone() for k, v in module.state_dict().items()}


def kan_layer_params(kan_module) -> list:
    """Per-layer dicts in the key convention of oracle.layers.kanlinear_forward."""
    out = []
    for layer in kan_module.layers:
        d = params_of(layer)
        d = {k: v for k, v in d.items() if not k.endswith("prev_x") and not k.endswith("branch_sign")
             and not k.endswith("branch_state")}
        out.append(d)
    return out
