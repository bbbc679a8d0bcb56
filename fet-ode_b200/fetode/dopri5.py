"""Adaptive Dormand-Prince 5(4) driver behind ``fetode.odeint(method=None | "dopri5")``.

torchdiffeq (the reference's solver, call sites train_kanfet_node_predprey.py:252, train_ecg_kan_fet_nn_ode.py:1034-1041,
train_kan_fet_ett.py:702) is an unpinned third-party dependency absent from the snapshot; this follows its published
algorithm (SURVEY appendix B -- PARITY UNPINNED): FSAL tableau, Hairer initial step, one RMS error norm over the whole
state tensor, controller safety 0.9 / ifactor 10 / dfactor 0.2, 4th-order dense output, time kept in float64 on the host.

The right-hand side is called once per stage, in the order torchdiffeq calls it (stateful neurons mutate their
hysteresis snapshot on every call, including the initial-step probe and rejected attempts).  Stage sums, the error
estimate, the error norm and the dense-output polynomial are libfetode kernels (csrc/rkstage.cu): one launch per
stage input, one for the norm; the accept/reject decision reads one double back per attempted step, as torchdiffeq does.
Step sizes are constants of the autograd graph.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List

import torch

from . import ops
from ._lib import check, lib, ptr, stream

# Butcher tableau (Dormand & Prince 1980) as exact rationals
_ALPHA = (1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0)
_BETA = (
    (1 / 5,),
    (3 / 40, 9 / 40),
    (44 / 45, -56 / 15, 32 / 9),
    (19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729),
    (9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656),
    (35 / 384, 0.0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84),
)
_C_ERR = (35 / 384 - 1951 / 21600, 0.0, 500 / 1113 - 22642 / 50715, 125 / 192 - 451 / 720,
          -2187 / 6784 + 12231 / 42400, 11 / 84 - 649 / 6300, -1 / 60)
_C_MID = (6025192743 / 30085553152 / 2, 0.0, 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
          187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2)


def _sumsq(err, y0, y1, atol, rtol, sync=None) -> float:
    """mean over ALL elements of (err / (atol + rtol max(|y0|,|y1|)))^2  (one global norm, torchdiffeq semantics)."""
    e, a, b = (ops.f32c(v.detach()) for v in (err, y0, y1))
    acc = torch.empty(1, dtype=torch.float64, device=e.device)
    check(lib().fetode_rk_error_sumsq(ptr(e), ptr(a), ptr(b), C.c_float(atol), C.c_float(rtol),
                                      C.c_int64(e.numel()), C.c_void_p(acc.data_ptr()), stream()), "rk_error_sumsq")
    n = float(e.numel())
    if sync is not None:                                  # batch-sharded run: global norm = all ranks' rows
        pair = torch.stack([acc[0], torch.tensor(n, dtype=torch.float64, device=e.device)])
        sync(pair)
        s, n = pair.tolist()
        return s / max(n, 1.0)
    return float(acc.item()) / max(n, 1.0)


def dopri5_integrate(func, y0, t, times: List[float], rtol, atol, options) -> torch.Tensor:
    from .odeint import _time_tensor_factory, combine
    rtol, atol = float(rtol), float(atol)
    if times[-1] < times[0]:
        raise NotImplementedError("fetode.odeint(dopri5): decreasing t is not used by the reference")
    tt = _time_tensor_factory(t, y0.device)
    sync = options.get("norm_sync")                      # callable(tensor[2]) that sums across ranks, or None
    max_steps = int(options.get("max_num_steps", 2 ** 31 - 1))
    stats = options.get("stats")
    y = ops.f32c(y0)
    f0 = func(tt(times[0]), y)

    # ---- initial step (Hairer, Norsett, Wanner I, II.4)
    first = options.get("first_step")
    if first is not None:
        dt = float(first)
    else:
        d0 = _sumsq(y, y, y, atol, rtol, sync) ** 0.5
        d1 = _sumsq(f0, y, y, atol, rtol, sync) ** 0.5
        h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
        f1 = func(tt(times[0] + h0), combine(y, [f0], [h0]))
        d2 = _sumsq(combine(None, [f1, f0], [1.0, -1.0]), y, y, atol, rtol, sync) ** 0.5 / h0
        h1 = max(1e-6, h0 * 1e-3) if (d1 <= 1e-15 and d2 <= 1e-15) else (0.01 / max(d1, d2)) ** 0.2
        dt = min(100.0 * h0, h1)

    sol = [y]
    fy, ta, tb = f0, times[0], times[0]
    dense = None                                         # (ya, yb, ks, h) of the last accepted step
    n_acc = n_rej = 0
    for tq in times[1:]:
        while tq > tb:
            t_new = tb + dt
            # torchdiffeq's _adaptive_step asserts: a vanished or non-finite step must stop the solve, not spin
            if not (math.isfinite(dt) and t_new > tb):
                raise AssertionError(f"fetode.odeint(dopri5): underflow in dt {dt}")
            ks = [fy]
            yi = y
            for al, be in zip(_ALPHA, _BETA):
                ti = t_new if al == 1.0 else tb + al * dt
                nz = [(k, b * dt) for k, b in zip(ks, be) if b != 0.0]
                yi = combine(y, [k for k, _ in nz], [c for _, c in nz])
                ks.append(func(tt(ti), yi))
            y1 = yi
            err = combine(None, [k for k, c in zip(ks, _C_ERR) if c != 0.0], [c * dt for c in _C_ERR if c != 0.0])
            ratio = _sumsq(err, y, y1, atol, rtol, sync) ** 0.5
            if not math.isfinite(ratio):
                # min(10, max(nan, 0.2)) is 10 in Python: dt would GROW on every rejected attempt and the solve would
                # never end.  torchdiffeq dies here too (dt becomes nan -> "underflow in dt" / non-finite state assert)
                raise AssertionError("fetode.odeint(dopri5): non-finite values in the state / error estimate")
            if ratio <= 1.0:
                dense = (y, y1, ks, dt)
                ta, tb, y, fy = tb, t_new, y1, ks[-1]
                n_acc += 1
            else:
                n_rej += 1
            if ratio == 0.0:
                dt = dt * 10.0
            else:
                dfac = 1.0 if ratio < 1.0 else 0.2
                dt = dt * min(10.0, max(0.9 / ratio ** 0.2, dfac))
            if n_acc + n_rej > max_steps:
                raise RuntimeError("fetode.odeint(dopri5): max_num_steps exceeded")
        sol.append(_dense_eval(dense, ta, tb, tq, combine) if dense is not None else y)
    if stats is not None:
        stats.update(accepted=n_acc, rejected=n_rej)
    return torch.stack(sol, 0)


def _dense_eval(dense, ta, tb, tq, combine):
    """4th-order interpolant through (ya, fa), y_mid, (yb, fb) evaluated at tq in [ta, tb]: every monomial
    coefficient is a linear combination of ya, yb and the stage slopes, so the whole evaluation is ONE
    fetode_rk_combine launch."""
    ya, yb, ks, h = dense
    if tq == tb:
        return yb
    if tq == ta:
        return ya
    x = (tq - ta) / (tb - ta)
    # y_mid = ya + h sum cm_i k_i ;  a = 2h(fb-fa) - 8(yb+ya) + 16 ymid ; b = h(5fa-3fb) + 18ya + 14yb - 32ymid ;
    # c = h(fb-4fa) - 11ya - 5yb + 16ymid ;  p(x) = ya + x h fa + x^2 c + x^3 b + x^4 a
    x2, x3, x4 = x * x, x ** 3, x ** 4
    w_mid = 16 * x2 - 32 * x3 + 16 * x4
    w_ya = 1.0 - 11 * x2 + 18 * x3 - 8 * x4 + w_mid
    w_yb = -5 * x2 + 14 * x3 - 8 * x4
    w_k = [w_mid * h * c for c in _C_MID]
    w_k[0] += h * (x - 4 * x2 + 5 * x3 - 2 * x4)
    w_k[6] += h * (x2 - 3 * x3 + 2 * x4)
    terms = [(ya, w_ya - 1.0), (yb, w_yb)] + [(k, w) for k, w in zip(ks, w_k) if w != 0.0]
    # fetode_rk_combine takes <= 8 terms per launch
    out = combine(ya, [v for v, _ in terms[:8]], [w for _, w in terms[:8]])
    if len(terms) > 8:
        out = combine(out, [v for v, _ in terms[8:]], [w for _, w in terms[8:]])
    return out
