"""Adaptive Dormand-Prince 5(4) driver behind ``fetode.odeint(method=None | "dopri5")``.

torchdiffeq (the reference's solver, call sites train_kanfet_node_predprey.py:252, train_ecg_kan_fet_nn_ode.py:1034-1041,
train_kan_fet_ett.py:702) is an unpinned third-party dependency absent from the snapshot; this follows its published
algorithm (SURVEY appendix B -- PARITY UNPINNED): FSAL tableau, Hairer initial step, one RMS error norm over the whole
state tensor, controller safety 0.9 / ifactor 10 / dfactor 0.2, 4th-order dense output, time kept in float64 on the host.

Two execution paths (both on the GPU): FUSED when ``func(t, y)`` is a bare call of a small KAN / KANFET the
hidden-units-on-lanes kernels support (the Lotka-Volterra script's model: for B <= 12 the whole accept / reject loop
runs inside one kernel, for larger batches one launch per attempted step and one per accepted step in the reverse sweep
-- see ``_FusedSolve`` below), STAGED for any other callable:

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
_C_ERR = (35 / 384 - 1951 / 21600, 0.0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
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
    model = fused_model(func, y0)
    if model is not None:
        return fused_dopri5(model, y0, times, rtol, atol, options)
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


def _dense_weights(ta, tb, tq, h):
    """4th-order interpolant through (ya, fa), y_mid, (yb, fb) at tq in [ta, tb] as weights of a linear combination:
    p = w_ya ya + w_yb yb + sum_s w_k[s] k_s.
    y_mid = ya + h sum cm_i k_i ;  a = 2h(fb-fa) - 8(yb+ya) + 16 ymid ; b = h(5fa-3fb) + 18ya + 14yb - 32ymid ;
    c = h(fb-4fa) - 11ya - 5yb + 16ymid ;  p(x) = ya + x h fa + x^2 c + x^3 b + x^4 a"""
    if tq == tb:
        return 0.0, 1.0, [0.0] * 7
    if tq == ta:
        return 1.0, 0.0, [0.0] * 7
    x = (tq - ta) / (tb - ta)
    x2, x3, x4 = x * x, x ** 3, x ** 4
    w_mid = 16 * x2 - 32 * x3 + 16 * x4
    w_ya = 1.0 - 11 * x2 + 18 * x3 - 8 * x4 + w_mid
    w_yb = -5 * x2 + 14 * x3 - 8 * x4
    w_k = [w_mid * h * c for c in _C_MID]
    w_k[0] += h * (x - 4 * x2 + 5 * x3 - 2 * x4)
    w_k[6] += h * (x2 - 3 * x3 + 2 * x4)
    return w_ya, w_yb, w_k


def _dense_eval(dense, ta, tb, tq, combine):
    """Every monomial coefficient of the interpolant is a linear combination of ya, yb and the stage slopes, so the
    whole evaluation is ONE fetode_rk_combine launch (two when more than 8 terms are non-zero)."""
    ya, yb, ks, h = dense
    if tq == tb:
        return yb
    if tq == ta:
        return ya
    w_ya, w_yb, w_k = _dense_weights(ta, tb, tq, h)
    terms = [(ya, w_ya - 1.0), (yb, w_yb)] + [(k, w) for k, w in zip(ks, w_k) if w != 0.0]
    # fetode_rk_combine takes <= 8 terms per launch
    out = combine(ya, [v for v, _ in terms[:8]], [w for _, w in terms[:8]])
    if len(terms) > 8:
        out = combine(out, [v for v, _ in terms[8:]], [w for _, w in terms[8:]])
    return out


# ------------------------------------------------------------------------------------------------------------------
# fused path: func(t, y) == model(y) for a small KAN / KANFET the hidden-units-on-lanes kernels support
# ------------------------------------------------------------------------------------------------------------------
# One launch per ATTEMPTED step (six right-hand sides, stage sums, solution, error norm: fetode_smallnet_attempt_fwd)
# and one 8-byte read-back for the accept / reject decision; the reverse sweep is one launch per ACCEPTED step
# (fetode_smallnet_attempt_bwd) walking the step checkpoints backwards.  Same controller, same call order of the
# right-hand side (initial value, initial-step probe, six per attempt -- hysteresis state advances on rejected
# attempts too) as the staged path above.
def _rk_struct(n_stages, first_eval, y1_last, dt, rtol=0.0, atol=0.0, rows=(), err=()):
    from ._lib import RkAttempt
    rk = RkAttempt()
    rk.n_stages, rk.first_eval, rk.y1_is_last_input = n_stages, first_eval, 1 if y1_last else 0
    rk.dt, rk.rtol, rk.atol = float(dt), float(rtol), float(atol)
    for i, row in enumerate(rows):
        for j, v in enumerate(row):
            rk.a[i + 1][j] = float(v)
    for i, v in enumerate(err):
        rk.e[i] = float(v)
    return rk


def fused_model(func, y0):
    """The KAN / KANFET module if ``func(t, y) == model(y)`` and the attempt kernels support its shape, else None."""
    import os
    if y0.dim() != 2 or os.environ.get("FETODE_DOPRI5_FUSED", "1") == "0":
        return None
    from .odeint import recognise_kan_rhs
    model = recognise_kan_rhs(func, y0)
    if model is None:
        return None
    plan = model.small_plan()
    if plan is None or lib().fetode_smallnet_wide_supported(C.byref(plan.net)) != 0:
        return None
    return model


class _FusedSolve:
    """Forward pass of the fused adaptive solve; ``keep`` retains what the reverse sweep needs."""

    def __init__(self, model, y0c, times, rtol, atol, options, keep):
        self.model, self.plan = model, model.small_plan()
        self.keep = keep
        self.B, self.D = y0c.shape
        self.dev = y0c.device
        self.img = self.plan.image()
        self.rtol, self.atol = rtol, atol
        self.steps = []                 # accepted: (dt, ckpt)
        self.rk = _rk_struct(7, 1, True, 0.0, rtol, atol, rows=_BETA, err=_C_ERR)   # one struct, dt set per attempt
        self.outs = []                  # (output index, accepted-step index, w_ya, w_yb, w_k[7])
        self.rhs0 = None                # (rk struct, ckpt) of f(t0, y0)
        self._acc = None
        self._slot = 0
        self._ck_floats = {}
        st, fresh = None, True
        if model._is_ferro():
            st = model.hysteresis_state(self.B, self.dev)
            fresh = st is None
            st = (torch.empty(self.B, self.plan.state_floats, dtype=torch.float32, device=self.dev) if fresh
                  else st.clone())
        self.st, self.fresh = st, fresh
        self.out = self._run(y0c, times, options)
        if st is not None:
            model.adopt_state(st)

    # -- one launch
    def _attempt(self, rk, y, k0, want_y1, want_err, want_ckpt):
        L = lib()
        B, D = self.B, self.D
        # ks | y1 | ckpt share one allocation (the caching allocator is the second largest host cost of an attempt)
        n_ks, n_y1, n_ck = rk.n_stages * B * D, (B * D if want_y1 else 0), 0
        if want_ckpt:
            key = (rk.n_stages, rk.first_eval)
            n_ck = self._ck_floats.get(key)
            if n_ck is None:
                n_ck = self._ck_floats[key] = int(L.fetode_smallnet_attempt_ckpt_floats(C.byref(self.plan.net),
                                                                                      C.byref(rk), C.c_int64(B)))
        pad = (-(n_ks + n_y1)) % 4                           # checkpoint records stay 16-byte aligned
        buf = torch.empty(n_ks + n_y1 + pad + n_ck, dtype=torch.float32, device=self.dev)
        ks = buf[:n_ks].view(rk.n_stages, B, D)
        y1 = buf[n_ks:n_ks + n_y1].view(B, D) if want_y1 else None
        ck = buf[n_ks + n_y1 + pad:] if want_ckpt else None
        err_ptr = C.c_void_p(0)
        if want_err:
            if self._acc is None or self._slot >= self._acc.numel():
                self._acc = torch.zeros(512, dtype=torch.float64, device=self.dev)
                self._slot = 0
            err_ptr = C.c_void_p(self._acc.data_ptr() + 8 * self._slot)
        check(L.fetode_smallnet_attempt_fwd(C.byref(self.plan.net), ptr(self.img), C.byref(rk), ptr(y), ptr(k0), ptr(y1),
                                            ptr(ks), ptr(self.st), C.c_int(1 if self.fresh else 0), C.c_int64(B),
                                            ptr(ck), err_ptr, stream()), "smallnet_attempt_fwd")
        self.fresh = False
        err = None
        if want_err:
            err = self._acc[self._slot]
            self._slot += 1
        return y1, ks, ck, err

    # -- controller on the device (B <= 12: one CTA holds the batch; the reference's LV run is one trajectory)
    CHUNK = 4096            # accepted steps per launch when the reverse sweep is wanted (checkpoint chunk)
    ATTEMPTS = 1 << 16      # attempts per launch: bounds the run time of one launch (~0.6 s)

    def _device_controller_ok(self, sync) -> bool:
        import os
        return (sync is None and os.environ.get("FETODE_DOPRI5_DEVICE", "1") != "0"
                and self.B <= int(lib().fetode_smallnet_dp5_max_rows()))

    def _run_on_device(self, y, f0, times, dt, max_steps, stats):
        L = lib()
        B, D, dev, T = self.B, self.D, self.dev, len(times)
        out = torch.empty(T, B, D, dtype=torch.float32, device=dev)
        out[0].copy_(y)
        if T == 1:
            if stats is not None:
                stats.update(accepted=0, rejected=0, fused=True, device_controller=True)
            return out
        times_d = torch.tensor(times, dtype=torch.float64).to(dev)
        ctrl = torch.tensor([times[0], times[0], dt, 1.0, 0.0, 0.0, 0.0, 0.0], dtype=torch.float64).to(dev)
        ycur, fy = y.clone(), f0.clone()
        self.chunks = []
        self.rec_step = self.rec_w = None
        if self.keep:
            self.rec_step = torch.full((T,), -1, dtype=torch.int32, device=dev)
            self.rec_w = torch.zeros(T, 9, dtype=torch.float32, device=dev)
            n_ck = int(L.fetode_smallnet_dp5_ckpt_floats(C.byref(self.plan.net), C.c_int64(B), C.c_int32(self.CHUNK)))
        base = 0
        while True:
            ck = dts = None
            if self.keep:
                ck = torch.empty(n_ck, dtype=torch.float32, device=dev)
                dts = torch.empty(self.CHUNK, dtype=torch.float32, device=dev)
            budget = int(min(self.ATTEMPTS, max(1, max_steps + 1)))
            check(L.fetode_smallnet_dp5_solve_fwd(
                C.byref(self.plan.net), ptr(self.img), ptr(ycur), ptr(fy), ptr(self.st), C.c_int(1 if self.fresh else 0),
                C.c_int64(B), C.c_void_p(times_d.data_ptr()), C.c_int32(T), ptr(out), C.c_void_p(ctrl.data_ptr()),
                C.c_double(self.rtol), C.c_double(self.atol), ptr(ck), ptr(dts), C.c_int32(self.CHUNK),
                C.c_void_p(self.rec_step.data_ptr() if self.keep else 0), ptr(self.rec_w), C.c_int32(budget), stream()),
                "smallnet_dp5_solve_fwd")
            self.fresh = False
            c = ctrl.tolist()                                  # the one synchronisation of this launch
            n_chunk, status = int(c[7]), int(c[6])
            if self.keep and n_chunk > 0:
                self.chunks.append((ck, dts, n_chunk, base))
                base += n_chunk
            if status == 2:
                raise AssertionError(f"fetode.odeint(dopri5): underflow in dt {c[2]}")
            if status == 3:
                raise AssertionError("fetode.odeint(dopri5): non-finite values in the state / error estimate")
            if c[4] + c[5] > max_steps:
                raise RuntimeError("fetode.odeint(dopri5): max_num_steps exceeded")
            if status == 0:
                break
        if stats is not None:
            stats.update(accepted=int(c[4]), rejected=int(c[5]), fused=True, device_controller=True)
        return out

    def _run(self, y, times, options):
        from .odeint import combine
        rtol, atol = self.rtol, self.atol
        sync = options.get("norm_sync")
        max_steps = int(options.get("max_num_steps", 2 ** 31 - 1))
        stats = options.get("stats")
        n_el = float(self.B * self.D)

        def ratio_of(err):
            if sync is not None:
                pair = torch.stack([err, torch.tensor(n_el, dtype=torch.float64, device=self.dev)])
                sync(pair)
                s, n = pair.tolist()
                return (s / max(n, 1.0)) ** 0.5
            return (float(err.item()) / max(n_el, 1.0)) ** 0.5

        rk0 = _rk_struct(1, 0, False, 0.0)
        _, ks0, ck0, _ = self._attempt(rk0, y, None, False, False, self.keep)
        f0 = ks0[0]
        self.rhs0 = (rk0, ck0)
        first = options.get("first_step")
        if first is not None:
            dt = float(first)
        else:
            d0 = _sumsq(y, y, y, atol, rtol, sync) ** 0.5
            d1 = _sumsq(f0, y, y, atol, rtol, sync) ** 0.5
            h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
            _, ksp, _, _ = self._attempt(_rk_struct(2, 1, False, h0, rows=((1.0,),)), y, f0, False, False, False)
            d2 = _sumsq(combine(None, [ksp[1], f0], [1.0, -1.0]), y, y, atol, rtol, sync) ** 0.5 / h0
            h1 = max(1e-6, h0 * 1e-3) if (d1 <= 1e-15 and d2 <= 1e-15) else (0.01 / max(d1, d2)) ** 0.2
            dt = min(100.0 * h0, h1)

        if self._device_controller_ok(sync):
            return self._run_on_device(y, f0, times, dt, max_steps, stats)
        sol = [y]
        fy, ta, tb = f0, times[0], times[0]
        dense = None
        n_acc = n_rej = 0
        for oi, tq in enumerate(times[1:], start=1):
            while tq > tb:
                t_new = tb + dt
                if not (math.isfinite(dt) and t_new > tb):
                    raise AssertionError(f"fetode.odeint(dopri5): underflow in dt {dt}")
                self.rk.dt = dt
                y1, ks, ck, err = self._attempt(self.rk, y, fy, True, True, self.keep)
                ratio = ratio_of(err)
                if not math.isfinite(ratio):
                    raise AssertionError("fetode.odeint(dopri5): non-finite values in the state / error estimate")
                if ratio <= 1.0:
                    dense = (y, y1, ks, dt)
                    ta, tb, y, fy = tb, t_new, y1, ks[6]
                    n_acc += 1
                    if self.keep:
                        self.steps.append((dt, ck))
                else:
                    n_rej += 1
                if ratio == 0.0:
                    dt = dt * 10.0
                else:
                    dfac = 1.0 if ratio < 1.0 else 0.2
                    dt = dt * min(10.0, max(0.9 / ratio ** 0.2, dfac))
                if n_acc + n_rej > max_steps:
                    raise RuntimeError("fetode.odeint(dopri5): max_num_steps exceeded")
            if dense is None:
                sol.append(y)
                continue
            sol.append(_dense_eval((dense[0], dense[1], list(dense[2]), dense[3]), ta, tb, tq, combine))
            if self.keep:
                self.outs.append((oi, n_acc - 1) + _dense_weights(ta, tb, tq, dense[3]))
        if stats is not None:
            stats.update(accepted=n_acc, rejected=n_rej, fused=True)
        return torch.stack(sol, 0)


class FusedDopri5Fn(torch.autograd.Function):
    """[T, B, D] dopri5 solution with autograd edges to y0 and the model's parameters.  The reverse sweep differentiates
    the accepted steps and the dense-output combinations; step sizes are constants (discretise-then-optimise, what
    torchdiffeq's plain ``odeint`` backward does through its recorded operations)."""

    @staticmethod
    def forward(ctx, model, times, rtol, atol, options, y0, *params):
        solve = _FusedSolve(model, ops.f32c(y0.detach()), times, rtol, atol, options, keep=True)
        ctx.solve = solve
        return solve.out

    @staticmethod
    def backward(ctx, g_out):
        from .training import unfold_param_grads
        sv = ctx.solve
        plan, L = sv.plan, lib()
        B, D, dev = sv.B, sv.D, sv.dev
        g_out = ops.f32c(g_out)
        gfold = torch.zeros(int(L.fetode_smallnet_gfold_floats(C.byref(plan.net))), dtype=torch.float32, device=dev)
        # adjoints of (ya, yb, k_0..k_6) of every step that produced outputs, from the dense-output weights
        G = {}
        if sv.outs:
            W = torch.tensor([[w_ya, w_yb] + list(w_k) for (_, _, w_ya, w_yb, w_k) in sv.outs],
                             dtype=torch.float32).to(dev)
            for j, (oi, n, *_w) in enumerate(sv.outs):
                term = W[j].view(9, 1, 1) * g_out[oi].unsqueeze(0)
                if n in G:
                    G[n] += term
                else:
                    G[n] = term.contiguous()
        net, img = C.byref(plan.net), ptr(sv.img)

        def sweep(rk, ck, g_y1, g_ks, x_y1, x_klast, x_y, want_k0):
            g_y = torch.empty(B, D, dtype=torch.float32, device=dev)
            g_k0 = torch.empty(B, D, dtype=torch.float32, device=dev) if want_k0 else None
            check(L.fetode_smallnet_attempt_bwd(net, img, C.byref(rk), ptr(ck), ptr(g_y1), ptr(g_ks), ptr(x_y1),
                                                ptr(x_klast), ptr(x_y), ptr(g_y), ptr(g_k0), ptr(gfold), C.c_int64(B),
                                                stream()), "smallnet_attempt_bwd")
            return g_y, g_k0

        carry_y = carry_k = None
        if getattr(sv, "chunks", None):
            T = g_out.shape[0]
            cursor = torch.full((1,), T - 1, dtype=torch.int32, device=dev)
            carry_y = torch.zeros(B, D, dtype=torch.float32, device=dev)
            carry_k = torch.zeros(B, D, dtype=torch.float32, device=dev)
            for ck, dts, n, base in reversed(sv.chunks):
                check(L.fetode_smallnet_dp5_solve_bwd(net, img, ptr(ck), ptr(dts), C.c_int32(n), C.c_int32(base),
                                                      C.c_void_p(sv.rec_step.data_ptr()), ptr(sv.rec_w), ptr(g_out),
                                                      C.c_void_p(cursor.data_ptr()), ptr(carry_y), ptr(carry_k),
                                                      ptr(gfold), C.c_int64(B), stream()), "smallnet_dp5_solve_bwd")
        for n in range(len(sv.steps) - 1, -1, -1):
            dt, ck = sv.steps[n]
            rk = sv.rk
            rk.dt = dt
            Gn = G.get(n)
            carry_y, carry_k = sweep(rk, ck, None if Gn is None else Gn[1], None if Gn is None else Gn[2:9], carry_y,
                                     carry_k, None if Gn is None else Gn[0], True)
        rk0, ck0 = sv.rhs0
        gy0, _ = sweep(rk0, ck0, None, None, None, carry_k, carry_y, False)
        gy0 = gy0 + g_out[0]
        outs = unfold_param_grads(plan, gfold, ctx.needs_input_grad, 6)
        return (None, None, None, None, None, gy0 if ctx.needs_input_grad[5] else None, *outs)


def fused_dopri5(model, y0, times, rtol, atol, options) -> torch.Tensor:
    plan = model.small_plan()
    params = [t for v in plan.views for n, t in v.tensors() if t is not None and n != "grid"]
    if torch.is_grad_enabled() and (y0.requires_grad or any(p.requires_grad for p in params)):
        return FusedDopri5Fn.apply(model, times, float(rtol), float(atol), options, y0, *params)
    return _FusedSolve(model, ops.f32c(y0.detach()), times, float(rtol), float(atol), options, keep=False).out
