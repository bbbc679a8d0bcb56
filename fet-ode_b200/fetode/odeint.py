"""Solver front-end: drop-in for ``torchdiffeq.odeint`` as the reference calls it, and for the
reference's own fixed-step loops.

Reference call sites / loops mirrored here (reference checkout, commit 704a0df4):
  torchdiffeq.odeint(func, y0, t[, method=, rtol=, atol=])   train_kanfet_node_predprey.py:252,260,
      train_ecg_kan_fet_nn_ode.py:266-271,558-565,887-892,1034-1041, train_ecg.py:1055-1056, ...
  integrate_euler / integrate_rk2 / integrate_rk4            train_ecg_kan_fet_nn_ode.py:670-705
  odeint_rk4(func, z0, t, n_substeps)                         train_kan_fet_ett.py:51-83

torchdiffeq itself is absent from the reference snapshot and from this image (unpinned third-party
dependency); its fixed-grid semantics (grid == t, "rk4" == 3/8 rule) and dopri5 controller are
restated from the published algorithm -- "parity unpinned", see DESIGN.md.

Two execution paths, both on the GPU (there is no CPU path):
  * fused  -- ``func`` is recognised as a bare call of a small KAN / KANFET (``lambda t, X: model(X)``):
              every step of every interval runs inside ONE kernel launch (csrc/smallnet.cu), output
              slices are written by the kernel.
  * staged -- any other callable: ``func`` is called per stage and the stage algebra is one
              fetode_rk_combine launch per stage input (csrc/rkstage.cu).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Sequence

import torch

from . import ops
from ._lib import TABLEAU, check, lib, ptr, stream

_METHODS = {"euler": "euler", "midpoint": "midpoint", "rk4": "rk4_38", "rk4_classic": "rk4_classic",
            "dopri5": "dopri5"}


# --------------------------------------------------------------------------------------
# recognition of  func(t, y) == model(y)
# --------------------------------------------------------------------------------------
def recognise_kan_rhs(func: Callable, y0: torch.Tensor):
    """Return the KAN/KANFET module if ``func(t, y)`` is exactly ``model(y)`` for a model the fused
    kernel supports, else None.  Decided by a dry run: inside ``ops.tracing()`` no fetode module
    computes or touches its hysteresis state; a fused-capable KAN records its argument and
    returns a sentinel, anything else aborts the trace."""
    model = getattr(func, "__fetode_rhs__", None)
    if model is not None:
        return model if model.small_plan() is not None else None
    with ops.tracing() as tr:
        try:
            out = func(torch.zeros((), dtype=torch.float64, device=y0.device), y0)
        except ops.TraceAbort:
            return None
    if len(tr.calls) == 1 and tr.calls[0][1] is y0 and out is tr.calls[0][2]:
        return tr.calls[0][0]
    return None


def as_rhs(model) -> Callable:
    """``func(t, y) = model(y)`` tagged so that odeint skips the dry-run recognition."""
    def func(t, y):
        return model(y)
    func.__fetode_rhs__ = model
    return func


# --------------------------------------------------------------------------------------
# staged path
# --------------------------------------------------------------------------------------
class _Combine(torch.autograd.Function):
    """out = y + sum_i c_i k_i  as one launch; backward is the transposed scaling."""

    @staticmethod
    def forward(ctx, coefs, y, *ks):
        ctx.coefs = coefs
        return ops.rk_combine(y, ks, coefs)

    @staticmethod
    def backward(ctx, g):
        g = ops.f32c(g)
        outs = [None, g if ctx.needs_input_grad[1] else None]
        for i, c in enumerate(ctx.coefs):
            outs.append(ops.rk_combine(None, (g,), (c,)) if ctx.needs_input_grad[2 + i] else None)
        return tuple(outs)


def combine(y: Optional[torch.Tensor], ks: Sequence[torch.Tensor], coefs: Sequence[float]) -> torch.Tensor:
    ks = [ops.f32c(k) for k in ks]
    coefs = tuple(float(c) for c in coefs)
    if torch.is_grad_enabled() and ((y is not None and y.requires_grad) or any(k.requires_grad for k in ks)):
        if y is None:
            y = torch.zeros_like(ks[0])
        return _Combine.apply(coefs, y, *ks)
    return ops.rk_combine(y, ks, coefs)


def _step_staged(kind: str, f, t0, dt: float, y, tt):
    """One step of the named tableau; tt(x) turns a python float into the time tensor handed to f."""
    if kind == "euler":                                   # integrate_euler :676
        return combine(y, [f(tt(t0), y)], [dt])
    if kind == "midpoint":                                # integrate_rk2 :686-689
        k1 = f(tt(t0), y)
        k2 = f(tt(t0 + 0.5 * dt), combine(y, [k1], [0.5 * dt]))
        return combine(y, [k2], [dt])
    if kind == "rk4_classic":                             # integrate_rk4 :698-703
        k1 = f(tt(t0), y)
        k2 = f(tt(t0 + 0.5 * dt), combine(y, [k1], [0.5 * dt]))
        k3 = f(tt(t0 + 0.5 * dt), combine(y, [k2], [0.5 * dt]))
        k4 = f(tt(t0 + dt), combine(y, [k3], [dt]))
        return combine(y, [k1, k2, k3, k4], [dt / 6.0, dt / 3.0, dt / 3.0, dt / 6.0])
    if kind == "rk4_38":                                  # torchdiffeq rk4_alt_step_func
        k1 = f(tt(t0), y)
        k2 = f(tt(t0 + dt / 3.0), combine(y, [k1], [dt / 3.0]))
        k3 = f(tt(t0 + 2.0 * dt / 3.0), combine(y, [k2, k1], [dt, -dt / 3.0]))
        k4 = f(tt(t0 + dt), combine(y, [k1, k2, k3], [dt, -dt, dt]))
        return combine(y, [k1, k2, k3, k4], [dt / 8.0, 3.0 * dt / 8.0, 3.0 * dt / 8.0, dt / 8.0])
    raise ValueError(kind)


class _TimeTable:
    """Stage times of a fixed-grid solve as 0-dim views of ONE device tensor: a single asynchronous H2D copy per solve
    instead of a synchronising ``torch.tensor(x, device=...)`` per right-hand-side call (which would stop the host from
    running ahead of the GPU).  ``steps`` = [(t0, dt), ...]; the values are formed with the same float expressions
    as _step_staged uses, so the lookup is exact; anything else falls back to a fresh tensor."""

    def __init__(self, steps, dtype, device):
        vals = []
        for t0, dt in steps:
            vals += [t0, t0 + 0.5 * dt, t0 + dt, t0 + dt / 3.0, t0 + 2.0 * dt / 3.0]
        self.dtype, self.device = dtype, device
        self.index = {}
        for i, v in enumerate(vals):
            self.index.setdefault(v, i)
        self.dev = None
        if vals:
            host = torch.tensor(vals, dtype=dtype)
            if torch.device(device).type == "cuda":
                self._host = host.pin_memory()
                self.dev = self._host.to(device, non_blocking=True)
            else:
                self.dev = host.to(device)

    def __call__(self, x: float):
        i = self.index.get(x)
        if i is None or self.dev is None:
            return torch.tensor(x, dtype=self.dtype, device=self.device)
        return self.dev[i]


_TT_CACHE: dict = {}


def _time_table(steps, dtype, device) -> _TimeTable:
    """Training loops solve over the same grid every step: keep the table (its pinned source and the device copy)
    so that a repeat solve costs no copy at all, and so that a solve captured in a CUDA graph never refers to a
    pinned buffer that has been recycled."""
    key = (tuple(steps), dtype, str(device))
    tt = _TT_CACHE.get(key)
    if tt is None:
        if len(_TT_CACHE) >= 64:
            _TT_CACHE.pop(next(iter(_TT_CACHE)))
        tt = _TT_CACHE[key] = _TimeTable(steps, dtype, device)
    return tt


def _time_dtype(t: torch.Tensor):
    return t.dtype if t.is_floating_point() else torch.float32


def _time_tensor_factory(t: torch.Tensor, device):
    """Adaptive solvers (dopri5): the stage times are only known step by step, and the accept / reject decision
    synchronises with the device once per step anyway."""
    dtype = _time_dtype(t)

    def tt(x: float):
        return torch.tensor(x, dtype=dtype, device=device)
    return tt


# --------------------------------------------------------------------------------------
# fused path
# --------------------------------------------------------------------------------------
def _fused_fixed(model, y0: torch.Tensor, times: List[float], kind: str) -> torch.Tensor:
    plan = model.small_plan()
    y0c = ops.f32c(y0.detach())
    B, D = y0c.shape
    T = len(times)
    out = torch.empty(T, B, D, dtype=torch.float32, device=y0c.device)
    out[0].copy_(y0c)
    ferro = model._is_ferro()
    st, fresh = None, True
    if ferro:
        st = model.hysteresis_state(B, y0c.device)
        fresh = st is None
        if fresh:
            st = torch.empty(B, plan.state_floats, dtype=torch.float32, device=y0c.device)
    # runs of intervals with the same fp32 step share one launch
    dts = [times[i + 1] - times[i] for i in range(T - 1)]
    i = 0
    cur = y0c
    while i < T - 1:
        j = i + 1
        f32 = torch.tensor(dts[i], dtype=torch.float32).item()
        while j < T - 1 and torch.tensor(dts[j], dtype=torch.float32).item() == f32:
            j += 1
        n = j - i
        plan.rk_fwd(TABLEAU[kind], cur, dts[i], n, state=st, state_is_zero=fresh, traj=out[i + 1:j + 1],
                    dump_every=1, y_out=out[j])
        fresh = False
        cur = out[j]
        i = j
    if ferro:
        model.adopt_state(st)
    return out


# --------------------------------------------------------------------------------------
# public API
# --------------------------------------------------------------------------------------
def odeint(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, event_fn=None):
    """torchdiffeq-style ``odeint``: returns a tensor of shape (len(t), *y0.shape) whose first
    slice is y0.  ``method``: None/"dopri5", "rk4" (3/8 rule, one step per output interval),
    "midpoint", "euler", plus "rk4_classic" (the reference's own integrate_rk4 tableau)."""
    if event_fn is not None:
        raise NotImplementedError("fetode.odeint: event_fn is not used by the reference and is not supported")
    if not torch.is_tensor(y0):
        raise TypeError("fetode.odeint: y0 must be a tensor (the reference never passes tuples)")
    ops._require_cuda(y0, "y0")
    kind = _METHODS.get(method or "dopri5")
    if kind is None:
        raise ValueError(f"fetode.odeint: unsupported method {method!r}")
    times = [float(v) for v in t.detach().cpu().double().tolist()]
    if len(times) < 1:
        raise ValueError("t must have at least one element")
    if any(b <= a for a, b in zip(times, times[1:])) and any(b >= a for a, b in zip(times, times[1:])):
        raise ValueError("t must be strictly increasing or strictly decreasing")
    options = dict(options or {})
    if kind == "dopri5":
        from .dopri5 import dopri5_integrate
        return dopri5_integrate(func, y0, t, times, rtol, atol, options)

    needs_grad = torch.is_grad_enabled() and (y0.requires_grad or _func_requires_grad(func))
    if "step_size" not in options and y0.dim() == 2:
        model = recognise_kan_rhs(func, y0)
        if model is not None:
            if needs_grad:
                from .training import fused_fixed_autograd
                return fused_fixed_autograd(model, y0, times, kind)
            return _fused_fixed(model, y0, times, kind)
    return _staged_fixed(func, y0, t, times, kind, options)


def _func_requires_grad(func) -> bool:
    model = getattr(func, "__fetode_rhs__", None)
    mods = [model] if model is not None else []
    clo = getattr(func, "__closure__", None) or ()
    for c in clo:
        try:
            v = c.cell_contents
        except ValueError:
            continue
        if isinstance(v, torch.nn.Module):
            mods.append(v)
    g = getattr(func, "__globals__", {})
    for name in getattr(getattr(func, "__code__", None), "co_names", ()):
        v = g.get(name)
        if isinstance(v, torch.nn.Module):
            mods.append(v)
    if isinstance(func, torch.nn.Module):
        mods.append(func)
    if not mods:
        return True                                       # unknown callable: assume it may need grad
    return any(p.requires_grad for m in mods for p in m.parameters())


def _staged_fixed(func, y0, t, times, kind, options):
    y = ops.f32c(y0)
    sol = [y]
    step_size = options.get("step_size")
    if step_size is None:
        tt = _time_table([(a, b - a) for a, b in zip(times, times[1:])], _time_dtype(t), y0.device)
        for a, b in zip(times, times[1:]):
            y = _step_staged(kind, func, a, b - a, y, tt)
            sol.append(y)
    else:
        # torchdiffeq grid_constructor from step_size: uniform internal grid, linear interpolation
        import math
        t0, t1 = times[0], times[-1]
        n = int(math.ceil((t1 - t0) / step_size + 1))
        grid = [t0 + k * step_size for k in range(n)]
        grid[-1] = t1
        tt = _time_table([(a, b - a) for a, b in zip(grid, grid[1:])], _time_dtype(t), y0.device)
        j = 1
        for a, b in zip(grid, grid[1:]):
            y1 = _step_staged(kind, func, a, b - a, y, tt)
            while j < len(times) and b >= times[j]:
                if times[j] == b:
                    sol.append(y1)
                else:
                    w = (times[j] - a) / (b - a)
                    sol.append(combine(y, [y1, y], [w, -w]))
                j += 1
            y = y1
    return torch.stack(sol, 0)


# ---- the reference's own loops (same names / signatures) ------------------------------------
def _integrate(kind, odefunc, h0, t0, t1, n_steps):
    model = recognise_kan_rhs(odefunc, h0) if h0.dim() == 2 else None
    dt = (t1 - t0) / n_steps
    times = [t0 + i * dt for i in range(n_steps + 1)]
    times[-1] = t0 + n_steps * dt
    grad = torch.is_grad_enabled() and (h0.requires_grad or _func_requires_grad(odefunc))
    if model is not None and not grad:
        plan = model.small_plan()
        y0c = ops.f32c(h0.detach())
        st, fresh = None, True
        if model._is_ferro():
            st = model.hysteresis_state(y0c.shape[0], y0c.device)
            fresh = st is None
            if fresh:
                st = torch.empty(y0c.shape[0], plan.state_floats, dtype=torch.float32, device=y0c.device)
        out = plan.rk_fwd(TABLEAU[kind], y0c, dt, n_steps, state=st, state_is_zero=fresh)
        if st is not None:
            model.adopt_state(st)
        return out
    if model is not None and grad:
        from .training import fused_fixed_autograd
        return fused_fixed_autograd(model, h0, times, kind, uniform_dt=dt)[-1]
    h = ops.f32c(h0)
    tt = _time_table([(t0 + i * dt, dt) for i in range(n_steps)], h0.dtype, h0.device)
    for i in range(n_steps):
        h = _step_staged(kind, odefunc, t0 + i * dt, dt, h, tt)
    return h


def integrate_euler(odefunc, h0, t0=0.0, t1=1.0, n_steps=8):
    """train_ecg_kan_fet_nn_ode.py:670-679."""
    return _integrate("euler", odefunc, h0, t0, t1, n_steps)


def integrate_rk2(odefunc, h0, t0=0.0, t1=1.0, n_steps=8):
    """train_ecg_kan_fet_nn_ode.py:681-691."""
    return _integrate("midpoint", odefunc, h0, t0, t1, n_steps)


def integrate_rk4(odefunc, h0, t0=0.0, t1=1.0, n_steps=8):
    """train_ecg_kan_fet_nn_ode.py:693-705."""
    return _integrate("rk4_classic", odefunc, h0, t0, t1, n_steps)


def odeint_rk4(func, z0, t, n_substeps=4):
    """train_kan_fet_ett.py:51-83: classic RK4, n_substeps per output interval, returns [T, B, D]."""
    times = [float(v) for v in t.detach().cpu().double().tolist()]
    z = ops.f32c(z0)
    out = [z]
    tt = _time_table([(a + s * ((b - a) / float(n_substeps)), (b - a) / float(n_substeps))
                     for a, b in zip(times, times[1:]) for s in range(n_substeps)], z0.dtype, z0.device)
    for a, b in zip(times, times[1:]):
        h = (b - a) / float(n_substeps)
        for s in range(n_substeps):
            z = _step_staged("rk4_classic", func, a + s * h, h, z, tt)
        out.append(z)
    return torch.stack(out, 0)
