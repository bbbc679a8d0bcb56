"""Import shims: make the reference's own ``import`` lines resolve to fetode, so its scripts run on the sm_100a
kernels after ONE added line (``import fetode.compat; fetode.compat.install()``) and no other edit.

Names covered (reference checkout 704a0df4):
  ``import efficientkan`` after ``sys.path.append("efficient_kan/")``   train_kanfet_node_predprey.py:11-12
  ``from efficient_kan.efficientkan import KANFET``                      train_ecg.py:16, train_kan_fet_ett.py:10,
                                                                         kan_diffusion_ett.py:31
  ``from torchdiffeq import odeint``                                     train_kanfet_node_predprey.py:5,
                                                                         train_ecg_kan_fet_nn_ode.py:13
  ``from ferro_class import FerroelectricBasis, BatchedFerroelectricBasis``   train_ecg.py:17
  ``from kan_diffusion import kan`` / ``from kan_diffusion.kan import KAN``   train_ecg_kan_fet_nn_ode.py:14,
                                                                         train_kan_fet_ett.py:13 (missing submodule:
                                                                         substituted by efficientkan.KAN)

``torchdiffeq`` is an unvendored third-party package: where a real one is importable it is left alone unless
``install(force_torchdiffeq=True)``; the shim's ``odeint`` / ``odeint_adjoint`` are fetode.odeint (no CPU path: tensors
must be on a CUDA device).  Nothing here computes anything; it only edits ``sys.modules``.
"""
from __future__ import annotations

import importlib
import importlib.util
import sys
import types
from typing import Dict, List

_INSTALLED: Dict[str, types.ModuleType] = {}


def _module(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__fetode_shim__ = True
    return m


def _importable(name: str) -> bool:
    if name in sys.modules and not getattr(sys.modules[name], "__fetode_shim__", False):
        return True
    try:
        return importlib.util.find_spec(name) is not None
    except (ImportError, ValueError):
        return False


def install(force_torchdiffeq: bool = False, override: bool = True) -> List[str]:
    """Register the shims in ``sys.modules``; returns the names installed.  ``override=False`` keeps any module of the
    same name that is already imported (e.g. the reference's own ``efficientkan`` in a mixed comparison run)."""
    import fetode
    from . import efficientkan as ek
    from . import ferro_class as fc
    from .odeint import odeint

    def odeint_adjoint(func, y0, t, **kw):
        """The reference never calls the adjoint form; provided so that either import line works.  Gradients are
        discretise-then-optimise either way (fetode's hand-written reverse sweep)."""
        kw.pop("adjoint_params", None)
        return odeint(func, y0, t, **kw)

    shims = {
        "efficientkan": ek,
        "ferro_class": fc,
    }
    pkg = _module("efficient_kan", KAN=ek.KAN, KANLinear=ek.KANLinear, KANFET=ek.KANFET)
    pkg.__path__ = []                                          # a package, so ``efficient_kan.efficientkan`` imports
    pkg.efficientkan = ek
    shims["efficient_kan"] = pkg
    shims["efficient_kan.efficientkan"] = ek
    kd = _module("kan_diffusion")
    kd.__path__ = []
    kdk = _module("kan_diffusion.kan", KAN=ek.KAN, KANLinear=ek.KANLinear)
    kd.kan = kdk
    shims["kan_diffusion"] = kd
    shims["kan_diffusion.kan"] = kdk
    if force_torchdiffeq or not _importable("torchdiffeq"):
        shims["torchdiffeq"] = _module("torchdiffeq", odeint=odeint, odeint_adjoint=odeint_adjoint,
                                       __version__="fetode-" + getattr(fetode, "__version__", "0"))
    done = []
    for name, mod in shims.items():
        if not override and name in sys.modules and sys.modules[name] is not mod:
            continue
        sys.modules[name] = mod
        _INSTALLED[name] = mod
        done.append(name)
    return done


def uninstall() -> None:
    """Remove every shim this module installed (only if the entry is still ours)."""
    for name, mod in list(_INSTALLED.items()):
        if sys.modules.get(name) is mod:
            del sys.modules[name]
        del _INSTALLED[name]
