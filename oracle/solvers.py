"""Oracle: the RK stage loops around the right-hand side (torch CPU).

TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.

Two groups:

1. The reference's OWN fixed-step loops (pinned by live execution):
   ``rk4_classic`` / ``rk2_midpoint`` / ``euler_fixed``  = integrate_rk4/rk2/euler
   (train_ecg_kan_fet_nn_ode.py:670-705), ``rk4_substeps`` = odeint_rk4
   (train_kan_fet_ett.py:51-83), ``euler_rollout`` = rollout (train_kanfet_predprey.py:181-189).

2. A restatement of the third-party ``torchdiffeq.odeint`` the reference calls
   (train_kanfet_node_predprey.py:252, train_ecg_kan_fet_nn_ode.py:1034-1041, ...).
   torchdiffeq is NOT vendored, NOT pinned (no requirements file; only the comment
   ``# pip install torchdiffeq``) and NOT installed here, so this part is written from the
   published algorithm (torchdiffeq 0.2.x: fixed-grid euler/midpoint/rk4(3/8-rule) and the
   Dormand-Prince 5(4) adaptive solver with Hairer initial step and 4th-order dense output).
   PARITY UNPINNED for this group, with one independent anchor: the Dormand-Prince tableau equals
   scipy.integrate's RK45 tables and the embedded-error / mid-point weights satisfy their order
   conditions (tests/test_oracle_dopri5_scipy.py); the controller details remain restated.
"""
from __future__ import annotations

from typing import Callable, List, Sequence

import torch

Tensor = torch.Tensor
Rhs = Callable[[object, Tensor], Tensor]


# ------------------------------------------------------------------ reference's own loops
def euler_fixed(f: Rhs, h0: Tensor, t0=0.0, t1=1.0, n_steps=8) -> Tensor:
    """integrate_euler (train_ecg_kan_fet_nn_ode.py:670-679)."""
    h, dt = h0, (t1 - t0) / n_steps
    t = torch.as_tensor(t0, dtype=h0.dtype)
    for _ in range(n_steps):
        h = h + dt * f(t, h)
        t = t + dt
    return h


def rk2_midpoint(f: Rhs, h0: Tensor, t0=0.0, t1=1.0, n_steps=8) -> Tensor:
    """integrate_rk2 (train_ecg_kan_fet_nn_ode.py:681-691)."""
    h, dt = h0, (t1 - t0) / n_steps
    t = torch.as_tensor(t0, dtype=h0.dtype)
    for _ in range(n_steps):
        k1 = f(t, h)
        k2 = f(t + 0.5 * dt, h + 0.5 * dt * k1)
        h = h + dt * k2
        t = t + dt
    return h


def rk4_classic(f: Rhs, h0: Tensor, t0=0.0, t1=1.0, n_steps=8, trajectory=False):
    """integrate_rk4 (train_ecg_kan_fet_nn_ode.py:693-705).  ``trajectory=True`` also returns
    the state after every step (used by the LV ensemble checks)."""
    h, dt = h0, (t1 - t0) / n_steps
    t = torch.as_tensor(t0, dtype=h0.dtype)
    traj = [h0]
    for _ in range(n_steps):
        k1 = f(t, h)
        k2 = f(t + 0.5 * dt, h + 0.5 * dt * k1)
        k3 = f(t + 0.5 * dt, h + 0.5 * dt * k2)
        k4 = f(t + dt, h + dt * k3)
        h = h + (dt / 6.0) * (k1 + 2 * k2 + 2 * k3 + k4)
        t = t + dt
        if trajectory:
            traj.append(h)
    return torch.stack(traj, 0) if trajectory else h


def rk4_substeps(f: Rhs, z0: Tensor, t: Tensor, n_substeps: int = 4) -> Tensor:
    """odeint_rk4 (train_kan_fet_ett.py:51-83): returns [T, B, D]."""
    z, out = z0, [z0]
    for i in range(t.shape[0] - 1):
        h = (t[i + 1] - t[i]) / float(n_substeps)
        ti = t[i]
        for _ in range(n_substeps):
            k1 = f(ti, z)
            k2 = f(ti + 0.5 * h, z + 0.5 * h * k1)
            k3 = f(ti + 0.5 * h, z + 0.5 * h * k2)
            k4 = f(ti + h, z + h * k3)
            z = z + (h / 6.0) * (k1 + 2 * k2 + 2 * k3 + k4)
            ti = ti + h
        out.append(z)
    return torch.stack(out, 0)


def euler_rollout(model: Callable[[Tensor], Tensor], x0: Tensor, steps: int) -> Tensor:
    """rollout (train_kanfet_predprey.py:181-189): dt = 1/steps, returns [steps+1, ...]."""
    dt, x, traj = 1.0 / steps, x0, [x0]
    for _ in range(steps):
        x = x + dt * model(x)
        traj.append(x)
    return torch.stack(traj, 0)


# ------------------------------------------------------------------ torchdiffeq restatement
_DP_ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
_DP_BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
_DP_CSOL = [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0]
_DP_CERR = [
    35 / 384 - 1951 / 21600, 0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
    -2187 / 6784 - -12231 / 42400, 11 / 84 - 649 / 6300, -1.0 / 60.0,
]
_DP_CMID = [
    6025192743 / 30085553152 / 2, 0, 51252292925 / 65400821598 / 2,
    -2691868925 / 45128329728 / 2, 187940372067 / 1594534317056 / 2,
    -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2,
]


def _rms(x: Tensor) -> Tensor:
    return x.abs().pow(2).mean().sqrt()


def _lin_interp(t0, t1, y0, y1, t):
    if t == t0:
        return y0
    if t == t1:
        return y1
    return y0 + ((t - t0) / (t1 - t0)) * (y1 - y0)


def _fixed_grid(step, f: Rhs, y0: Tensor, t: Tensor) -> Tensor:
    sol = [y0]
    j, y = 1, y0
    for i in range(t.shape[0] - 1):
        t0, t1 = t[i], t[i + 1]
        dt = t1 - t0
        y1 = y + step(f, t0, dt, t1, y)
        while j < t.shape[0] and t1 >= t[j]:
            sol.append(_lin_interp(t0, t1, y, y1, t[j]))
            j += 1
        y = y1
    return torch.stack(sol, 0)


def _step_euler(f, t0, dt, t1, y):
    return dt * f(t0, y)


def _step_midpoint(f, t0, dt, t1, y):
    half = 0.5 * dt
    return dt * f(t0 + half, y + f(t0, y) * half)


def _step_rk4_38(f, t0, dt, t1, y):
    """torchdiffeq's ``rk4`` is the 3/8-rule variant, not the classic tableau."""
    third, two_thirds = 1.0 / 3.0, 2.0 / 3.0
    k1 = f(t0, y)
    k2 = f(t0 + dt * third, y + dt * k1 * third)
    k3 = f(t0 + dt * two_thirds, y + dt * (k2 - k1 * third))
    k4 = f(t1, y + dt * (k1 - k2 + k3))
    return (k1 + 3 * (k2 + k3) + k4) * dt * 0.125


def _dopri5(f: Rhs, y0: Tensor, t: Tensor, rtol: float, atol: float, max_steps=2 ** 31 - 1,
            stats: dict | None = None) -> Tensor:
    dt_dtype = t.dtype

    def initial_step(t0, f0):
        scale = atol + y0.abs() * rtol
        d0, d1 = _rms(y0 / scale), _rms(f0 / scale)
        if d0 < 1e-5 or d1 < 1e-5:
            h0 = torch.tensor(1e-6, dtype=y0.dtype)
        else:
            h0 = 0.01 * d0 / d1
        h0 = h0.abs()
        f1 = f(t0 + h0, y0 + h0 * f0)
        d2 = torch.abs(_rms((f1 - f0) / scale) / h0)
        if d1 <= 1e-15 and d2 <= 1e-15:
            h1 = torch.max(torch.tensor(1e-6, dtype=y0.dtype), h0 * 1e-3)
        else:
            h1 = (0.01 / max(d1, d2)) ** (1.0 / 5.0)
        return torch.min(100 * h0, h1.abs()).to(dt_dtype)

    def rk_step(ya, fa, ta, dt, tb):
        ta_, dt_, tb_ = ta.to(ya.dtype), dt.to(ya.dtype), tb.to(ya.dtype)
        ks = [fa]
        yi = ya
        for al, be in zip(_DP_ALPHA, _DP_BETA):
            ti = tb_ if al == 1.0 else ta_ + al * dt_
            yi = ya + sum(k * (b * dt_) for k, b in zip(ks, be))
            ks.append(f(ti, yi))
        y1 = yi                                  # FSAL: last beta row == c_sol
        err = sum(k * (dt_ * c) for k, c in zip(ks, _DP_CERR))
        return y1, ks[-1], err, ks

    def fit(ya, yb, ks, dt):
        dt_ = dt.to(ya.dtype)
        ymid = ya + sum(k * (dt_ * c) for k, c in zip(ks, _DP_CMID))
        fa, fb = ks[0], ks[-1]
        a = 2 * dt_ * (fb - fa) - 8 * (yb + ya) + 16 * ymid
        b = dt_ * (5 * fa - 3 * fb) + 18 * ya + 14 * yb - 32 * ymid
        c = dt_ * (fb - 4 * fa) - 11 * ya - 5 * yb + 16 * ymid
        return [ya, dt_ * fa, c, b, a]

    def evaluate(coef, ta, tb, tq):
        x = ((tq - ta) / (tb - ta)).to(coef[0].dtype)
        tot, xp = coef[0] + x * coef[1], x
        for c in coef[2:]:
            xp = xp * x
            tot = tot + xp * c
        return tot

    f0 = f(t[0], y0)
    dt = initial_step(t[0], f0)
    if stats is not None:
        stats["first_step"] = float(dt)
    y, fy, ta, tb, coef = y0, f0, t[0], t[0], [y0] * 5
    sol, n_acc, n_rej = [y0], 0, 0
    for i in range(1, t.shape[0]):
        while t[i] > tb:
            t_new = tb + dt
            y1, f1, err, ks = rk_step(y, fy, tb, dt, t_new)
            tol = atol + rtol * torch.max(y.abs(), y1.abs())
            ratio = _rms(err / tol)
            if ratio <= 1:
                coef = fit(y, y1, ks, dt)
                ta, tb, y, fy = tb, t_new, y1, f1
                n_acc += 1
            else:
                n_rej += 1
            # step-size controller: safety 0.9, ifactor 10, dfactor 0.2, order 5
            if ratio == 0:
                dt = dt * 10.0
            else:
                dfac = 1.0 if ratio < 1 else 0.2
                fac = min(10.0, max(0.9 / float(ratio) ** 0.2, dfac))
                dt = dt * fac
            if n_acc + n_rej > max_steps:
                raise RuntimeError("max_num_steps exceeded")
        sol.append(evaluate(coef, ta, tb, t[i]))
    if stats is not None:
        stats.update(accepted=n_acc, rejected=n_rej)
    return torch.stack(sol, 0)


def odeint(f: Rhs, y0: Tensor, t: Tensor, *, rtol=1e-7, atol=1e-9, method=None, stats=None) -> Tensor:
    """torchdiffeq.odeint restatement (see module docstring; PARITY UNPINNED)."""
    method = method or "dopri5"
    if method == "euler":
        return _fixed_grid(_step_euler, f, y0, t)
    if method == "midpoint":
        return _fixed_grid(_step_midpoint, f, y0, t)
    if method == "rk4":
        return _fixed_grid(_step_rk4_38, f, y0, t)
    if method == "dopri5":
        return _dopri5(f, y0, t, rtol, atol, stats=stats)
    raise ValueError(f"unsupported method {method!r}")
