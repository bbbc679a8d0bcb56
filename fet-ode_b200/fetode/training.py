"""Differentiable fused fixed-step integration of a small KAN / KANFET right-hand side.

What the reference does for ``loss.backward()`` through ``odeint(calDeriv, X0, t)``
(train_kanfet_node_predprey.py:252-256, train_kanfet_mlp_node_predprey.py:265-272) is plain autograd through every
solver stage (no adjoint).  Here the forward pass is the fused Runge-Kutta kernel with per-step checkpoints
(csrc/smallnet.cu) and the backward pass is the reverse-sweep kernel (csrc/smallnet_bwd.cu) followed by the
chain rule from folded to reference parameters (fetode_smallnet_unfold).  All arithmetic is in libfetode.so.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import ops
from ._lib import TABLEAU, KanLayer, KanLayerGrads, check, lib, ptr, stream


def _runs(times: List[float], uniform_dt: Optional[float]):
    """[(first_interval, n_intervals, dt)] -- consecutive intervals with the same fp32 step share a launch."""
    T = len(times)
    if uniform_dt is not None:
        return [(0, T - 1, float(uniform_dt))] if T > 1 else []
    dts = [times[i + 1] - times[i] for i in range(T - 1)]
    f32 = [torch.tensor(d, dtype=torch.float32).item() for d in dts]
    out, i = [], 0
    while i < T - 1:
        j = i + 1
        while j < T - 1 and f32[j] == f32[i]:
            j += 1
        out.append((i, j - i, dts[i]))
        i = j
    return out


WIDE_MAX_ROWS = 20480


def _wide_ok(plan, B: int) -> bool:
    """Use the hidden-units-on-lanes kernels (csrc/smallnet_wide.cu) for this solve?  They win while the batch does
    not fill the machine -- measured fwd+bwd+Adam of the LV step (34 RK4 steps, B200): B = 1: 0.76 ms vs 4.3 ms,
    B = 4096: 1.4 vs 4.7, B = 16384: 4.0 vs 4.8, B = 32768: ~7.7 vs 5.7, B = 65536: 15.0 vs 9.8 -- the
    lane-per-trajectory pair takes over from ~20 000 rows.  FETODE_SN_WIDE=0 / 1 forces either (A/B timing; read once)."""
    global _WIDE_ENV
    if _WIDE_ENV is None:
        import os
        _WIDE_ENV = os.environ.get("FETODE_SN_WIDE", "")
    if lib().fetode_smallnet_wide_supported(C.byref(plan.net)) != 0 or _WIDE_ENV == "0":
        return False
    return _WIDE_ENV == "1" or B <= WIDE_MAX_ROWS


_WIDE_ENV = None


def unfold_param_grads(plan, gfold: torch.Tensor, needs_input_grad, first: int) -> list:
    """Chain rule from the folded gradient buffer of the small-net reverse sweeps to the reference's parameters
    (fetode_smallnet_unfold); returns one tensor (or None) per ``fused_params`` entry, ``needs_input_grad[first:]``
    saying which are wanted."""
    L = lib()
    dev = gfold.device
    views = plan.views
    structs = (KanLayer * 2)(*[v.struct() for v in views])
    grads = (KanLayerGrads * 2)()
    outs = []
    k = first
    for li, v in enumerate(views):
        for name, t in v.tensors():
            if t is None or name == "grid":
                continue
            gt = torch.zeros(t.shape, dtype=torch.float32, device=dev) if needs_input_grad[k] else None
            k += 1
            outs.append(gt)
            if name.startswith("bp"):
                grads[li].bp[int(name[2:])] = ptr(gt).value
            else:
                setattr(grads[li], name, ptr(gt).value)
    check(L.fetode_smallnet_unfold(C.byref(plan.net), structs, ptr(gfold), grads, stream()), "smallnet_unfold")
    # no host synchronisation: any f32c() temporary behind `structs` was allocated on this stream, and the caching
    # allocator only hands its block to work queued later on the same stream (this also keeps the step capturable)
    return outs


class FusedRKFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, times, kind, uniform_dt, y0, *params):
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
            else:
                st = st.clone()
        runs = _runs(times, uniform_dt)
        ckpts = []
        cur = y0c
        NC = D + plan.state_floats
        wide = _wide_ok(plan, B)
        img = plan.image()
        for (i, n, dt) in runs:
            if wide:
                nck = int(lib().fetode_smallnet_wide_ckpt_floats(C.byref(plan.net), C.c_int64(B), C.c_int32(n)))
                ck = torch.empty(max(nck, 1), dtype=torch.float32, device=y0c.device)
                check(lib().fetode_smallnet_wide_fwd(C.byref(plan.net), ptr(img), C.c_int(TABLEAU[kind]), ptr(cur),
                                                     ptr(out[i + n]), ptr(st), C.c_int(1 if fresh else 0), C.c_int64(B),
                                                     C.c_double(dt), C.c_int32(n), ptr(out[i + 1:i + n + 1]), C.c_int32(1),
                                                     ptr(ck), stream()), "smallnet_wide_fwd")
            else:
                ck = torch.empty(n, NC, B, dtype=torch.float32, device=y0c.device)
                plan.rk_fwd(TABLEAU[kind], cur, dt, n, state=st, state_is_zero=fresh, traj=out[i + 1:i + n + 1],
                            dump_every=1, ckpt=ck, y_out=out[i + n])
            ckpts.append(ck)
            fresh = False
            cur = out[i + n]
        if ferro:
            model.adopt_state(st)
        ctx.model, ctx.plan, ctx.kind, ctx.runs, ctx.ckpts, ctx.wide = model, plan, kind, runs, ckpts, wide
        ctx.image = img
        ctx.shape = (T, B, D)
        return out

    @staticmethod
    def backward(ctx, g_out):
        plan, L = ctx.plan, lib()
        T, B, D = ctx.shape
        g_out = ops.f32c(g_out)
        dev = g_out.device
        nfold = int(L.fetode_smallnet_gfold_floats(C.byref(plan.net)))
        gfold = torch.zeros(nfold, dtype=torch.float32, device=dev)
        carry = None
        for (i, n, dt), ck in zip(reversed(ctx.runs), reversed(ctx.ckpts)):
            g_y0 = torch.empty(B, D, dtype=torch.float32, device=dev)
            fn = L.fetode_smallnet_wide_bwd if ctx.wide else L.fetode_smallnet_rk_bwd
            check(fn(C.byref(plan.net), ptr(ctx.image), C.c_int(TABLEAU[ctx.kind]), ptr(ck),
                     ptr(carry), ptr(g_out[i + 1:i + n + 1]), C.c_int32(1), ptr(g_y0),
                     ptr(gfold), C.c_int64(B), C.c_double(dt), C.c_int32(n), stream()),
                  "smallnet_wide_bwd" if ctx.wide else "smallnet_rk_bwd")
            carry = g_y0
        gy0 = g_out[0] if carry is None else carry + g_out[0]
        outs = unfold_param_grads(plan, gfold, ctx.needs_input_grad, 5)
        return (None, None, None, None, gy0 if ctx.needs_input_grad[4] else None, *outs)


def fused_fixed_autograd(model, y0: torch.Tensor, times: List[float], kind: str,
                         uniform_dt: Optional[float] = None) -> torch.Tensor:
    """[T, B, D] solution at ``times`` with autograd edges to y0 and to the model's parameters."""
    plan = model.small_plan()
    params = [t for v in plan.views for n, t in v.tensors() if t is not None and n != "grid"]
    return FusedRKFn.apply(model, times, kind, uniform_dt, y0, *params)
