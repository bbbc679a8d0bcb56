"""Thin callers of the C ABI (include/fetode.h): pack caches, launches, autograd glue.

Everything here is plumbing -- tensors are allocated with torch, the arithmetic happens in
libfetode.so.  Nothing in this file computes a KAN / ferro value on the host or with torch ops.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import (BRANCH_FERRO2D, BRANCH_LOGISTIC, BRANCH_NONE, KanLayer, KanLayerGrads, SmallNet,
                   check, f32c, lib, ptr, stream)

_BP_NAMES = {
    BRANCH_LOGISTIC: ("a", "b"),
    BRANCH_FERRO2D: ("k", "Ec", "Ps", "bias", "coef"),
}


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"fetode: {what} lives on {t.device}; the KAN-FET path runs on CUDA only "
            "(there is no CPU fallback) -- move the module and its inputs with .to('cuda')")


class LayerView:
    """The tensors of one KANLinear-shaped module in the order the C struct wants them.

    ``module`` must expose the reference's attribute names (efficientkan.py:49-90 /
    mnist_kuramoto_kan.py:40-78): grid, base_weight, spline_weight, [spline_scaler],
    [logistic_basis.{a,b}|{k,Ec,Ps,bias,coef}], [logistic_weight], [logistic_scaler], [logistic_bias].
    """

    def __init__(self, module):
        self.m = module

    @property
    def branch(self) -> int:
        m = self.m
        d = m.__dict__
        if not d.get("enable_logistic_basis", d.get("use_logistic_basis", False)):
            return BRANCH_NONE
        lb = m._modules.get("logistic_basis")
        if lb is None:
            lb = m.logistic_basis
        return BRANCH_FERRO2D if "Ec" in lb._parameters or hasattr(lb, "Ec") else BRANCH_LOGISTIC

    @staticmethod
    def _get(m, name):
        """Parameter / buffer `name` of module m without nn.Module.__getattr__'s fallback chain (hot path)."""
        t = m._parameters.get(name)
        if t is None:
            t = m._buffers.get(name)
        if t is None:
            t = getattr(m, name, None)
        return t

    def tensors(self) -> List[Tuple[str, Optional[torch.Tensor]]]:
        m = self.m
        g = self._get
        br = self.branch
        d = m.__dict__
        out = [("grid", g(m, "grid")), ("base_weight", g(m, "base_weight")), ("spline_weight", g(m, "spline_weight")),
               ("spline_scaler", g(m, "spline_scaler") if d.get("enable_standalone_scale_spline", True) else None)]
        if br == BRANCH_NONE:
            out += [("branch_weight", None), ("branch_scaler", None), ("branch_bias", None)]
            out += [(f"bp{i}", None) for i in range(5)]
        else:
            out.append(("branch_weight", g(m, "logistic_weight")))
            out.append(("branch_scaler", g(m, "logistic_scaler") if d.get("enable_standalone_scale_logistic", False) else None))
            out.append(("branch_bias", g(m, "logistic_bias")))
            names = _BP_NAMES[br]
            lb = m._modules.get("logistic_basis")
            if lb is None:
                lb = m.logistic_basis
            for i in range(5):
                out.append((f"bp{i}", g(lb, names[i]) if i < len(names) else None))
        return out

    def struct(self, tensors=None) -> KanLayer:
        m = self.m
        br = self.branch
        ts = dict(tensors if tensors is not None else self.tensors())
        s = KanLayer()
        s.in_features, s.out_features = int(m.in_features), int(m.out_features)
        s.grid_size, s.spline_order = int(m.grid_size), int(m.spline_order)
        s.num_basis = 0 if br == BRANCH_NONE else int(m.num_basis)
        s.branch = br
        for name in ("grid", "base_weight", "spline_weight", "spline_scaler", "branch_weight",
                     "branch_scaler", "branch_bias"):
            t = ts[name]
            if t is not None:
                _require_cuda(t, name)
            setattr(s, name, ptr(None if t is None else _keep(s, f32c(t.detach()))).value)
        for i in range(5):
            t = ts[f"bp{i}"]
            s.bp[i] = ptr(None if t is None else _keep(s, f32c(t.detach()))).value
        # mnist_kuramoto_kan.KANLinear (use_logistic_basis) never multiplies by scale_logistic in forward
        # (mnist_kuramoto_kan.py:139-140); efficientkan.KANLinear does (efficientkan.py:149-158)
        s.scale_logistic = 1.0 if "use_logistic_basis" in m.__dict__ else float(getattr(m, "scale_logistic", 1.0))
        lb = getattr(m, "logistic_basis", None)
        s.gate_slope = float(getattr(lb, "gate_slope", 10.0))
        s.alpha = float(getattr(lb, "alpha", 0.8))
        return s

    def key(self, tensors=None):
        return tuple(None if t is None else (t.data_ptr(), t._version)
                     for _, t in (tensors if tensors is not None else self.tensors()))


def _keep(owner, t: torch.Tensor) -> torch.Tensor:
    """Pin a (possibly temporary) tensor's lifetime to the ctypes struct that points at it."""
    refs = getattr(owner, "_refs", None)
    if refs is None:
        refs = []
        owner._refs = refs
    refs.append(t)
    return t


# --------------------------------------------------------------------------------------
# any-width KANLinear
# --------------------------------------------------------------------------------------
class LayerPack:
    """Device image of one layer's folded parameters; rebuilt when any parameter's version
    counter or storage changes (i.e. after every optimiser step)."""

    def __init__(self, module):
        self.view = LayerView(module)
        self._key = None
        self.blob: Optional[torch.Tensor] = None
        self.struct: Optional[KanLayer] = None
        self._handle: Optional[torch.Tensor] = None
        self._handle_key = None

    def __deepcopy__(self, memo):
        # copy.deepcopy(model) after a forward: the copy gets a fresh, empty pack bound to the copied module (the
        # ctypes struct holds raw pointers and the fold handle is a non-leaf tensor: neither may be copied)
        import copy
        return LayerPack(copy.deepcopy(self.view.m, memo))

    def __reduce__(self):
        return (LayerPack, (self.view.m,))

    def get(self, tensors=None) -> Tuple[KanLayer, torch.Tensor]:
        k = self.view.key(tensors)
        if k != self._key:
            s = self.view.struct(tensors)
            n = lib().fetode_kanlinear_wcat_floats(C.byref(s))
            if n < 0:
                raise RuntimeError("fetode: KANLinear configuration not supported by the sm_100a kernels "
                                   f"(in={s.in_features}, out={s.out_features}, grid={s.grid_size}, "
                                   f"order={s.spline_order}, nb={s.num_basis})")
            dev = self.view.m.base_weight.device
            if self.blob is None or self.blob.numel() != n or self.blob.device != dev:
                self.blob = torch.empty(n, dtype=torch.float32, device=dev)
            check(lib().fetode_kanlinear_pack(C.byref(s), ptr(self.blob), stream()), "kanlinear_pack")
            self.struct, self._key = s, k
        return self.struct, self.blob

    def fold(self, tensors, params) -> torch.Tensor:
        """The FoldFn handle of the current parameter versions: one autograd node per (layer, parameter version),
        shared by every call until the parameters change."""
        k = (self.view.key(tensors), tuple(p.requires_grad for p in params))
        h = self._handle
        if h is None or k != self._handle_key or h.grad_fn is None:
            # drop the stale node (every reference, the local one too) BEFORE building the new one: while it lives it
            # pins the parameters' AccumulateGrad nodes and the stream they were created on
            h = self._handle = None
            h = self._handle = FoldFn.apply(self, tensors, *params)
            h._fetode_shared = h.grad_fn.shared
            self._handle_key = k
        return h


def kanlinear_fwd(pack: LayerPack, x2d: torch.Tensor, prev: Optional[torch.Tensor], tensors=None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """y = KANLinear(x2d); returns (y, new_prev) where new_prev (FERRO2D only) is x2d itself."""
    _require_cuda(x2d, "input")
    s, blob = pack.get(tensors)
    B = x2d.shape[0]
    y = torch.empty(B, s.out_features, dtype=torch.float32, device=x2d.device)
    ferro = s.branch == BRANCH_FERRO2D
    if B == 0:
        return y, (x2d if ferro else None)
    check(lib().fetode_kanlinear_fwd(C.byref(s), ptr(blob), ptr(x2d), ptr(prev) if ferro else None,
                                      C.c_int(1 if prev is None else 0), None, ptr(y),
                                      C.c_int64(B), stream()), "kanlinear_fwd")
    return y, (x2d if ferro else None)


class FoldFn(torch.autograd.Function):
    """The parameter fold as an autograd node: reference parameters -> packed image (``LayerPack.blob``), represented
    in the graph by a data-less ``handle`` of the folded gradient's shape.  A layer that is applied many times between
    two optimiser steps (once per Runge-Kutta stage: 32 times per ECG step) shares ONE fold node; every
    ``KanLinearFn.backward`` adds its folded-weight gradient into one accumulator, and this node's backward applies the
    chain rule to the reference's parameters (fetode_kanlinear_unfold) ONCE -- instead of an unfold plus a dozen
    gradient-accumulation launches per call."""

    @staticmethod
    def forward(ctx, pack, tens, *params):
        s, _ = pack.get(tens)
        ctx.pack, ctx.tens = pack, tens
        ctx.names = [n for n, t in tens if t is not None and n != "grid"]
        ctx.shared = {"acc": None}
        ctx.set_materialize_grads(False)
        n = max(int(lib().fetode_kanlinear_bwd_ws_floats(C.byref(s))), 4) + 2 * _bp_floats(s)
        return torch.empty(n, dtype=torch.float32, device=params[0].device)

    @staticmethod
    def backward(ctx, g_handle):
        acc = ctx.shared["acc"]
        n_out = 2 + len(ctx.names)
        if g_handle is None or acc is None:
            return (None,) * n_out
        s, _ = ctx.pack.get(ctx.tens)
        tens = dict(ctx.tens)
        n_ws = acc.numel() - 2 * _bp_floats(s)
        want = [ctx.needs_input_grad[2 + idx] for idx in range(len(ctx.names))]
        direct = {"bp0": 0, "bp1": 1}                        # accumulated by the input-gradient kernel, not unfolded
        sizes = [(tens[n].numel() + 3) // 4 * 4 if (w and n not in direct) else 0 for n, w in zip(ctx.names, want)]
        flat = torch.zeros(max(sum(sizes), 1), dtype=torch.float32, device=acc.device)
        g = KanLayerGrads()
        outs, off = [], 0
        for name, w, n in zip(ctx.names, want, sizes):
            t = tens[name]
            if not w:
                outs.append(None)
                continue
            if name in direct:
                o = n_ws + direct[name] * _bp_floats(s)
                outs.append(acc[o:o + t.numel()].view(t.shape))
                continue
            gt = flat[off:off + t.numel()].view(t.shape)
            off += n
            outs.append(gt)
            if name.startswith("bp"):
                g.bp[int(name[2:])] = ptr(gt).value
            else:
                setattr(g, name, ptr(gt).value)
        check(lib().fetode_kanlinear_unfold(C.byref(s), ptr(acc), C.byref(g), stream()), "kanlinear_unfold")
        return (None, None, *outs)


def _bp_floats(s: KanLayer) -> int:
    return (int(s.in_features) * int(s.num_basis) + 3) // 4 * 4


class KanLinearFn(torch.autograd.Function):
    """autograd node of one KANLinear call: (x, fold handle) -> y.  The hysteresis state ``prev`` is a constant of the
    graph, as in the reference (ferro_class.py:553-554 detaches the snapshots).  ``handle`` is FoldFn's output (None
    when no parameter requires a gradient): the parameters' gradient flows through it."""

    @staticmethod
    def forward(ctx, pack: LayerPack, prev, tens, x2d, handle):
        y, _ = kanlinear_fwd(pack, x2d, prev, tens)
        ctx.pack, ctx.prev, ctx.tens = pack, prev, tens
        ctx.shared = None if handle is None else handle._fetode_shared
        ctx.save_for_backward(x2d)
        return y

    @staticmethod
    def backward(ctx, gy):
        (x2d,) = ctx.saved_tensors
        pack = ctx.pack
        s, blob = pack.get(ctx.tens)
        L = lib()
        gy = f32c(gy)
        B = x2d.shape[0]
        need_x = ctx.needs_input_grad[3]
        gx = torch.empty_like(x2d) if need_x else None
        ferro = s.branch == BRANCH_FERRO2D
        shared = ctx.shared if ctx.needs_input_grad[4] else None
        if shared is None:                                   # frozen parameters: input gradient only
            if not need_x:
                return (None, None, None, None, None)
            ws = torch.empty(max(int(L.fetode_kanlinear_bwd_ws_floats(C.byref(s))), 1), dtype=torch.float32,
                             device=x2d.device)
            check(L.fetode_kanlinear_bwd(C.byref(s), ptr(blob), ptr(x2d), ptr(ctx.prev) if ferro else None,
                                         C.c_int(1 if ctx.prev is None else 0), ptr(gy), ptr(gx),
                                         C.byref(KanLayerGrads()), ptr(ws), C.c_int64(B), stream()), "kanlinear_bwd")
            return (None, None, None, gx, None)
        first = shared["acc"] is None
        if first:
            n_ws = max(int(L.fetode_kanlinear_bwd_ws_floats(C.byref(s))), 4)
            shared["acc"] = torch.zeros(n_ws + 2 * _bp_floats(s), dtype=torch.float32, device=x2d.device)
            # whatever this backward pass reaches (a partial torch.autograd.grad never runs the fold node), the
            # accumulator never survives the pass
            torch.autograd.Variable._execution_engine.queue_callback(lambda: shared.__setitem__("acc", None))
        acc = shared["acc"]
        n_ws = acc.numel() - 2 * _bp_floats(s)
        has_bp = s.branch != BRANCH_NONE
        gbp0 = acc[n_ws:] if has_bp else None
        gbp1 = acc[n_ws + _bp_floats(s):] if has_bp else None
        check(L.fetode_kanlinear_bwd_acc(C.byref(s), ptr(blob), ptr(x2d), ptr(ctx.prev) if ferro else None,
                                         C.c_int(1 if ctx.prev is None else 0), ptr(gy), ptr(gx), ptr(gbp0), ptr(gbp1),
                                         ptr(acc), C.c_int64(B), stream()), "kanlinear_bwd_acc")
        # the accumulator is handed to autograd once; the later calls add into it in place
        return (None, None, None, gx, acc if first else None)


def kanlinear_apply(pack: LayerPack, x2d: torch.Tensor, prev: Optional[torch.Tensor]) -> torch.Tensor:
    tens = pack.view.tensors()
    if torch.is_grad_enabled():
        params = [t for n, t in tens if t is not None and n != "grid"]
        if any(p.requires_grad for p in params):
            return KanLinearFn.apply(pack, prev, tens, x2d, pack.fold(tens, params))
        if x2d.requires_grad:
            return KanLinearFn.apply(pack, prev, tens, x2d, None)
    return kanlinear_fwd(pack, x2d, prev, tens)[0]


# --------------------------------------------------------------------------------------
# one gradient accumulator per (module, parameter version) for parameters applied many times per step
# --------------------------------------------------------------------------------------
class ParamGroupFn(torch.autograd.Function):
    """(parameters...) -> data-less handle.  A module that is applied once per Runge-Kutta stage (32 times per ECG
    step: the hard-gated basis, the LayerNorm, the dense head) routes its parameter gradients through ONE such node:
    every call's backward kernel accumulates (+=) into one shared buffer, zero-filled once per backward pass, and this
    node hands the buffer's views to the parameters -- instead of a zero-fill per parameter and call plus an
    AccumulateGrad add per parameter and call (~1 200 adds and ~700 fills per ECG step)."""

    @staticmethod
    def forward(ctx, shared, *params):
        ctx.shared = shared
        ctx.shapes = [tuple(p.shape) for p in params]
        ctx.set_materialize_grads(False)
        return torch.empty(max(_group_floats(ctx.shapes), 4), dtype=torch.float32, device=params[0].device)

    @staticmethod
    def backward(ctx, g):
        acc = ctx.shared["acc"]
        n = 1 + len(ctx.shapes)
        if g is None or acc is None:
            return (None,) * n
        outs, off = [], 0
        for i, shp in enumerate(ctx.shapes):
            numel = 1
            for d in shp:
                numel *= d
            outs.append(acc[off:off + numel].view(shp) if ctx.needs_input_grad[1 + i] else None)
            off += (numel + 3) // 4 * 4
        return (None, *outs)


def _group_floats(shapes) -> int:
    tot = 0
    for shp in shapes:
        numel = 1
        for d in shp:
            numel *= d
        tot += (numel + 3) // 4 * 4
    return tot


class ParamGroup:
    """Cache of the ParamGroupFn handle of one module (rebuilt when a parameter's version / storage changes)."""

    def __init__(self):
        self._key, self._handle = None, None

    def __deepcopy__(self, memo):
        return ParamGroup()                                   # the handle is a non-leaf tensor: never copied

    def __reduce__(self):
        return (ParamGroup, ())

    def handle(self, params) -> Optional[torch.Tensor]:
        """Handle for ``params`` (list of Parameters), or None when autograd is off / nothing requires a gradient."""
        if not torch.is_grad_enabled() or not any(p.requires_grad for p in params):
            return None
        key = tuple((p.data_ptr(), p._version, p.requires_grad) for p in params)
        h = self._handle
        if h is None or key != self._key or h.grad_fn is None:
            h = self._handle = None
            shared = {"acc": None}
            h = self._handle = ParamGroupFn.apply(shared, *params)
            h._fetode_shared = shared
            h._fetode_shapes = [tuple(p.shape) for p in params]
            self._key = key
        return h


def group_grad_views(handle, device):
    """Inside a consumer's backward: (views of the shared accumulator, one per parameter of the group; first) where
    ``first`` says this call created (zero-filled) the buffer and must return it as the handle's gradient."""
    shared, shapes = handle._fetode_shared, handle._fetode_shapes
    first = shared["acc"] is None
    if first:
        shared["acc"] = torch.zeros(max(_group_floats(shapes), 4), dtype=torch.float32, device=device)
        torch.autograd.Variable._execution_engine.queue_callback(lambda: shared.__setitem__("acc", None))
    acc = shared["acc"]
    views, off = [], 0
    for shp in shapes:
        numel = 1
        for d in shp:
            numel *= d
        views.append(acc[off:off + numel].view(shp))
        off += (numel + 3) // 4 * 4
    return views, acc, first


# --------------------------------------------------------------------------------------
# LayerNorm + tanh bound (prologue of the ECG right-hand side)
# --------------------------------------------------------------------------------------
class LnTanhFn(torch.autograd.Function):
    """y = h_bound * tanh(LayerNorm(x) / h_bound), train_ecg_kan_fet_nn_ode.py:459-461, one launch each way."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, h_bound, handle=None):
        """handle: ParamGroup handle of (gamma, beta) -- their gradients then go through it (gamma / beta are passed
        detached), one accumulator per backward pass."""
        B, D = x.shape
        y = torch.empty_like(x)
        stats = torch.empty(2, B, dtype=torch.float32, device=x.device)
        g, b = f32c(gamma.detach()), f32c(beta.detach())
        check(lib().fetode_ln_tanh_fwd(ptr(x), ptr(g), ptr(b), ptr(y), ptr(stats[0]), ptr(stats[1]), C.c_int64(B),
                                       C.c_int32(D), C.c_float(eps), C.c_float(h_bound), stream()), "ln_tanh_fwd")
        ctx.save_for_backward(x, g, stats, y)
        ctx.h_bound, ctx.handle = h_bound, handle
        return y

    @staticmethod
    def backward(ctx, gy):
        x, g, stats, y = ctx.saved_tensors
        B, D = x.shape
        gy = f32c(gy)
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        h = ctx.handle if (ctx.handle is not None and ctx.needs_input_grad[5]) else None
        if h is not None:
            (gg, gb), acc, first = group_grad_views(h, x.device)
            want_g = want_b = True
        else:
            gp = torch.zeros(2, D, dtype=torch.float32, device=x.device)
            gg, gb = gp[0], gp[1]
            want_g, want_b = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        check(lib().fetode_ln_tanh_bwd(ptr(x), ptr(g), ptr(stats[0]), ptr(stats[1]), ptr(y), ptr(gy), ptr(gx),
                                       ptr(gg) if want_g else None, ptr(gb) if want_b else None, C.c_int64(B), C.c_int32(D),
                                       C.c_float(ctx.h_bound), stream()), "ln_tanh_bwd")
        if h is not None:
            return (gx, None, None, None, None, acc if first else None)
        return (gx, gg if want_g else None, gb if want_b else None, None, None, None)


def ln_tanh(x: torch.Tensor, ln: torch.nn.LayerNorm, h_bound: float) -> torch.Tensor:
    """Fused when the kernel has the case (CUDA fp32 [B, D], D <= 256, affine LayerNorm over the last axis); other
    configurations keep the reference's two torch ops (on the device: CPU tensors are refused like everywhere else)."""
    _require_cuda(x, "input")
    if (x.is_cuda and x.dim() == 2 and x.dtype == torch.float32 and ln.elementwise_affine and ln.bias is not None
            and tuple(ln.normalized_shape) == (x.shape[1],) and x.shape[1] <= 256 and h_bound > 0):
        grp = ln.__dict__.get("_fetode_pgroup")
        if grp is None:
            grp = ln.__dict__["_fetode_pgroup"] = ParamGroup()
        h = grp.handle([ln.weight, ln.bias])
        if h is not None:
            return LnTanhFn.apply(f32c(x), ln.weight.detach(), ln.bias.detach(), float(ln.eps), float(h_bound), h)
        return LnTanhFn.apply(f32c(x), ln.weight, ln.bias, float(ln.eps), float(h_bound))
    h = ln(x)
    return torch.tanh(h) if h_bound == 1.0 else h_bound * torch.tanh(h / h_bound)


class LinearFn(torch.autograd.Function):
    """y = x W^T + b in fp32 FMA arithmetic (csrc/linear.cu): the dense head of the ECG right-hand side."""

    @staticmethod
    def forward(ctx, x, weight, bias, handle=None):
        """handle: ParamGroup handle of (weight[, bias]) -- see LnTanhFn.forward."""
        B, K = x.shape
        O = weight.shape[0]
        y = torch.empty(B, O, dtype=torch.float32, device=x.device)
        w = f32c(weight.detach())
        b = None if bias is None else f32c(bias.detach())
        check(lib().fetode_linear_fwd(ptr(x), ptr(w), ptr(b), ptr(y), C.c_int64(B), C.c_int32(K), C.c_int32(O),
                                      stream()), "linear_fwd")
        ctx.save_for_backward(x, w)
        ctx.has_bias, ctx.handle = bias is not None, handle
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        B, K = x.shape
        O = w.shape[0]
        gy = f32c(gy)
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        h = ctx.handle if (ctx.handle is not None and ctx.needs_input_grad[3]) else None
        if h is not None:
            views, acc, first = group_grad_views(h, x.device)
            gw, gb = views[0], (views[1] if ctx.has_bias else None)
            check(lib().fetode_linear_bwd(ptr(x), ptr(w), ptr(gy), ptr(gx), ptr(gw), ptr(gb), C.c_int64(B), C.c_int32(K),
                                          C.c_int32(O), stream()), "linear_bwd")
            return gx, None, None, (acc if first else None)
        want_w = ctx.needs_input_grad[1]
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        flat = torch.zeros(O * K + O, dtype=torch.float32, device=x.device) if (want_w or want_b) else None
        gw = flat[:O * K].view(O, K) if want_w else None
        gb = flat[O * K:] if want_b else None
        check(lib().fetode_linear_bwd(ptr(x), ptr(w), ptr(gy), ptr(gx), ptr(gw), ptr(gb), C.c_int64(B), C.c_int32(K),
                                      C.c_int32(O), stream()), "linear_bwd")
        return gx, gw, gb, None


def linear(x: torch.Tensor, lin: torch.nn.Linear) -> torch.Tensor:
    """nn.Linear on a CUDA fp32 [B, in] input through csrc/linear.cu; other shapes / dtypes keep the module's own call
    (on the device: CPU tensors are refused like everywhere else)."""
    _require_cuda(x, "input")
    if x.is_cuda and x.dim() == 2 and x.dtype == torch.float32 and lin.weight.dtype == torch.float32:
        grp = lin.__dict__.get("_fetode_pgroup")
        if grp is None:
            grp = lin.__dict__["_fetode_pgroup"] = ParamGroup()
        params = [lin.weight] + ([lin.bias] if lin.bias is not None else [])
        h = grp.handle(params)
        if h is not None and all(p.requires_grad for p in params):
            return LinearFn.apply(f32c(x), lin.weight.detach(), None if lin.bias is None else lin.bias.detach(), h)
        return LinearFn.apply(f32c(x), lin.weight, lin.bias)
    return lin(x)


class HeadFn(torch.autograd.Function):
    """dh = gain * (silu(z) W^T + b): the tail of MLPKANODEFunc.forward (train_ecg_kan_fet_nn_ode.py:471-480) as one
    launch (csrc/linear.cu fetode_head_fwd); backward = three launches, the gradients of (W, b, gain) accumulate in the
    ParamGroup buffer behind ``handle``."""

    @staticmethod
    def forward(ctx, z, weight, bias, gain, handle=None):
        B, K = z.shape
        O = weight.shape[0]
        y = torch.empty(B, O, dtype=torch.float32, device=z.device)
        u = torch.empty(B, O, dtype=torch.float32, device=z.device) if handle is not None else None
        w, b, g = f32c(weight.detach()), f32c(bias.detach()), f32c(gain.detach()).reshape(1)
        check(lib().fetode_head_fwd(ptr(z), ptr(w), ptr(b), ptr(g), ptr(u), ptr(y), C.c_int64(B), C.c_int32(K),
                                    C.c_int32(O), stream()), "head_fwd")
        if handle is not None:
            ctx.save_for_backward(z, w, g, u)
        ctx.handle = handle
        return y

    @staticmethod
    def backward(ctx, gy):
        z, w, g, u = ctx.saved_tensors
        B, K = z.shape
        O = w.shape[0]
        gy = f32c(gy)
        gz = torch.empty_like(z) if ctx.needs_input_grad[0] else None
        (gw, gb, gg), acc, first = group_grad_views(ctx.handle, z.device)
        check(lib().fetode_head_bwd(ptr(z), ptr(w), ptr(g), ptr(u), ptr(gy), ptr(gz), ptr(gw), ptr(gb), ptr(gg.reshape(1)),
                                    C.c_int64(B), C.c_int32(K), C.c_int32(O), stream()), "head_bwd")
        return gz, None, None, None, (acc if first else None)


def head(z: torch.Tensor, act, lin: torch.nn.Linear, gain: torch.Tensor) -> torch.Tensor:
    """gain * lin(act(z)) -- fused when act is SiLU, the input a CUDA fp32 [B, in] tensor, the layer has a bias and either
    nothing or every one of (weight, bias, gain) requires a gradient; otherwise the three steps separately."""
    _require_cuda(z, "input")
    if (isinstance(act, torch.nn.SiLU) and z.is_cuda and z.dim() == 2 and z.dtype == torch.float32
            and lin.weight.dtype == torch.float32 and lin.bias is not None and gain.numel() == 1
            and gain.dtype == torch.float32):
        params = [lin.weight, lin.bias, gain]
        if not torch.is_grad_enabled() or not (z.requires_grad or any(p.requires_grad for p in params)):
            return HeadFn.apply(f32c(z), lin.weight, lin.bias, gain, None)
        if all(p.requires_grad for p in params):
            grp = lin.__dict__.get("_fetode_hgroup")
            if grp is None:
                grp = lin.__dict__["_fetode_hgroup"] = ParamGroup()
            h = grp.handle(params)
            if h is not None:
                return HeadFn.apply(f32c(z), lin.weight.detach(), lin.bias.detach(), gain.detach(), h)
    return gain * linear(act(z), lin)


# --------------------------------------------------------------------------------------
# fused small-width network + Runge-Kutta loop
# --------------------------------------------------------------------------------------
class SmallNetPlan:
    """Packed image + launch helpers for a two-layer KAN/KANFET whose widths fit the fused
    kernel (fetode_smallnet_supported).  ``layers`` are KANLinear-shaped modules."""

    def __init__(self, layers: Sequence):
        self.views = [LayerView(m) for m in layers]
        self.net = SmallNet()
        self.net.n_layers = len(layers)
        widths = [int(layers[0].in_features)] + [int(m.out_features) for m in layers]
        for i, w in enumerate(widths[:3]):
            self.net.widths[i] = w
        self.net.grid_size = int(layers[0].grid_size)
        self.net.spline_order = int(layers[0].spline_order)
        br = self.views[0].branch
        self.net.branch = br
        self.net.num_basis = 0 if br == BRANCH_NONE else int(layers[0].num_basis)
        self.D, self.H = widths[0], widths[1] if len(widths) > 1 else 0
        self._key = None
        self.packed: Optional[torch.Tensor] = None

    def __deepcopy__(self, memo):
        import copy
        return SmallNetPlan([copy.deepcopy(v.m, memo) for v in self.views])

    def __reduce__(self):
        return (SmallNetPlan, ([v.m for v in self.views],))

    @staticmethod
    def supported(layers: Sequence) -> bool:
        if len(layers) != 2:
            return False
        v = [LayerView(m) for m in layers]
        if v[0].branch != v[1].branch:
            return False
        net = SmallNet()
        net.n_layers = 2
        net.widths[0], net.widths[1], net.widths[2] = (int(layers[0].in_features), int(layers[0].out_features),
                                                      int(layers[1].out_features))
        if int(layers[1].in_features) != net.widths[1]:
            return False
        net.grid_size, net.spline_order = int(layers[0].grid_size), int(layers[0].spline_order)
        if (int(layers[1].grid_size), int(layers[1].spline_order)) != (net.grid_size, net.spline_order):
            return False
        net.branch = v[0].branch
        net.num_basis = 0 if net.branch == BRANCH_NONE else int(layers[0].num_basis)
        if net.branch != BRANCH_NONE and int(layers[1].num_basis) != net.num_basis:
            return False
        return lib().fetode_smallnet_supported(C.byref(net)) == 0

    @property
    def state_floats(self) -> int:
        return int(lib().fetode_smallnet_state_floats(C.byref(self.net)))

    def image(self) -> torch.Tensor:
        k = tuple(v.key() for v in self.views)
        if k != self._key:
            structs = (KanLayer * 2)(*[v.struct() for v in self.views])
            n = lib().fetode_smallnet_packed_floats(C.byref(self.net))
            if n < 0:
                raise RuntimeError("fetode: network not supported by the fused small-width kernel")
            dev = self.views[0].m.base_weight.device
            if self.packed is None or self.packed.numel() != n or self.packed.device != dev:
                self.packed = torch.empty(n, dtype=torch.float32, device=dev)
            check(lib().fetode_smallnet_pack(C.byref(self.net), structs, ptr(self.packed), stream()),
                  "smallnet_pack")
            # no host synchronisation: f32c() temporaries were allocated on this stream, and the caching allocator
            # reuses their blocks only for work queued after the pack kernel on the same stream
            self._key = k
        return self.packed

    def rhs(self, y: torch.Tensor, state: Optional[torch.Tensor], state_is_zero: bool) -> torch.Tensor:
        _require_cuda(y, "state vector")
        img = self.image()
        dy = torch.empty_like(y)
        if y.shape[0] == 0:
            return dy
        check(lib().fetode_smallnet_rhs(C.byref(self.net), ptr(img), ptr(y), ptr(dy), ptr(state),
                                        C.c_int(1 if state_is_zero else 0), C.c_int64(y.shape[0]), stream()),
              "smallnet_rhs")
        return dy

    def rk_fwd(self, tableau: int, y0: torch.Tensor, dt: float, n_steps: int, *, state=None,
               state_is_zero=True, traj: Optional[torch.Tensor] = None, dump_every: int = 1,
               ckpt: Optional[torch.Tensor] = None, y_out: Optional[torch.Tensor] = None) -> torch.Tensor:
        _require_cuda(y0, "initial state")
        img = self.image()
        if y_out is None:
            y_out = torch.empty_like(y0)
        if y0.shape[0] == 0:
            return y_out
        check(lib().fetode_smallnet_rk_fwd(C.byref(self.net), ptr(img), C.c_int(tableau), ptr(y0), ptr(y_out),
                                           ptr(state), C.c_int(1 if state_is_zero else 0),
                                           C.c_int64(y0.shape[0]), C.c_double(dt), C.c_int32(n_steps),
                                           ptr(traj), C.c_int32(dump_every), ptr(ckpt), stream()),
              "smallnet_rk_fwd")
        return y_out


# --------------------------------------------------------------------------------------
# dry-run tracing used by odeint to recognise  func(t, y) == model(y)
# --------------------------------------------------------------------------------------
class TraceAbort(Exception):
    """Raised by any fetode module that is evaluated during a dry run but cannot be fused."""


class _Trace:
    def __init__(self):
        self.calls = []          # (module, input, sentinel)


_ACTIVE_TRACE: Optional[_Trace] = None


class tracing:
    def __enter__(self):
        global _ACTIVE_TRACE
        self._saved = _ACTIVE_TRACE
        _ACTIVE_TRACE = _Trace()
        return _ACTIVE_TRACE

    def __exit__(self, *exc):
        global _ACTIVE_TRACE
        _ACTIVE_TRACE = self._saved
        return False


def active_trace() -> Optional[_Trace]:
    return _ACTIVE_TRACE


def abort_if_tracing() -> None:
    if _ACTIVE_TRACE is not None:
        raise TraceAbort()


# --------------------------------------------------------------------------------------
# Runge-Kutta stage algebra
# --------------------------------------------------------------------------------------
def rk_combine(y: Optional[torch.Tensor], ks: Sequence[torch.Tensor], coefs: Sequence[float]) -> torch.Tensor:
    """out = y + sum_i coefs[i] * ks[i] in one launch (csrc/rkstage.cu)."""
    ref = y if y is not None else ks[0]
    _require_cuda(ref, "stage tensor")
    ks = [f32c(k.detach()) for k in ks]
    yc = None if y is None else f32c(y.detach())
    out = torch.empty_like(ref, dtype=torch.float32, memory_format=torch.contiguous_format)
    n = len(ks)
    karr = (C.c_void_p * max(n, 1))(*[k.data_ptr() for k in ks])
    carr = (C.c_float * max(n, 1))(*[float(c) for c in coefs])
    check(lib().fetode_rk_combine(ptr(yc), karr, carr, C.c_int32(n), ptr(out), C.c_int64(out.numel()), stream()),
          "rk_combine")
    return out


FERRO_FULL, FERRO_TWO_D, FERRO_HARDGATE = _lib.FERRO_FULL, _lib.FERRO_TWO_D, _lib.FERRO_HARDGATE
FERRO_LOGISTIC = 3


# --------------------------------------------------------------------------------------
# stand-alone ferroelectric / logistic neurons (csrc/ferro.cu)
# --------------------------------------------------------------------------------------
def _ferro_struct(cfg: dict, k, Ec, Ps, bias, coef) -> _lib.Ferro:
    s = _lib.Ferro()
    s.variant, s.in_dim, s.out_dim, s.num_basis = cfg["variant"], cfg["I"], cfg["O"], cfg["K"]
    for name, t in (("k", k), ("Ec", Ec), ("Ps", Ps), ("bias", bias), ("coef", coef)):
        setattr(s, name, ptr(None if t is None else _keep(s, f32c(t.detach()))).value)
    s.gate_slope, s.alpha = float(cfg["gate_slope"]), float(cfg["alpha"])
    s.apply_sigmoid = int(cfg.get("apply_sigmoid", 0))
    s.pre_bound, s.post_op = float(cfg.get("pre_bound", 0.0)), int(cfg.get("post_op", 0))
    return s


class FerroFn(torch.autograd.Function):
    """One call of a ferroelectric neuron.  ``prev`` / ``br`` are constants of the graph
    (ferro_class.py:381-382).  Side outputs (basis, new branch_sign, gate bits) are left in cfg["side"]."""

    @staticmethod
    def forward(ctx, cfg, prev, br, x, k, Ec, Ps, bias, coef, handle=None):
        """handle: ParamGroup handle of the (non-None) parameters in the order k, Ec, Ps, bias, coef -- their gradients
        then go through it (the parameters are passed detached): one accumulator per backward pass for a neuron that
        is applied once per Runge-Kutta stage."""
        _require_cuda(x, "input")
        B = x.shape[0]
        I, O, K = cfg["I"], cfg["O"], cfg["K"]
        full = cfg["variant"] == FERRO_FULL
        out = torch.empty((B, O) if full else (B, I, K), dtype=torch.float32, device=x.device)
        side = {}
        elem_shape = (B, I, O, K) if full else (B, I, K)
        basis = torch.empty(elem_shape, dtype=torch.float32, device=x.device) if cfg.get("want_basis") else None
        br_out = torch.empty(elem_shape, dtype=torch.float32, device=x.device) if cfg.get("want_br_out") else None
        gate = torch.empty(elem_shape, dtype=torch.uint8, device=x.device) if cfg.get("want_gate") else None
        prev_out = cfg.get("prev_out")
        s = _ferro_struct(cfg, k, Ec, Ps, bias, coef)
        if br_out is not None and br is None:
            br = torch.ones(elem_shape, dtype=torch.float32, device=x.device)
        # prev_mode: 0 read prev, 1 all zero, 2 (FULL) prev == effective input (first call with a batch, :373-378)
        mode = int(cfg.get("prev_mode", 1 if prev is None else 0))
        if B > 0:
            check(lib().fetode_ferro_fwd(C.byref(s), ptr(x), ptr(prev if mode == 0 else None), C.c_int(mode), ptr(br),
                                         ptr(out), ptr(basis), ptr(br_out), ptr(gate), ptr(prev_out),
                                         C.c_int64(B), stream()), "ferro_fwd")
        side.update(basis=basis, br_out=br_out, gate=gate)
        cfg["side"] = side
        ctx.cfg, ctx.prev, ctx.br, ctx.mode, ctx.handle = cfg, (prev if mode == 0 else None), br, mode, handle
        ctx.save_for_backward(x, k, Ec, Ps, bias, coef if coef is not None else x.new_empty(0),
                              out if cfg.get("post_op", 0) else x.new_empty(0))
        ctx.has_coef = coef is not None
        return out

    @staticmethod
    def backward(ctx, gout):
        x, k, Ec, Ps, bias, coef, yout = ctx.saved_tensors
        if not ctx.has_coef:
            coef = None
        cfg = ctx.cfg
        if not cfg.get("post_op", 0):
            yout = None
        gout = f32c(gout)
        B = x.shape[0]
        s = _ferro_struct(cfg, k, Ec, Ps, bias, coef)
        g = _lib.FerroGrads()
        grads = []
        h = ctx.handle if (ctx.handle is not None and ctx.needs_input_grad[9]) else None
        named = (("k", k), ("Ec", Ec), ("Ps", Ps), ("bias", bias), ("coef", coef))
        acc, first = None, False
        if h is not None:
            views, acc, first = group_grad_views(h, x.device)
            vi = iter(views)
            for name, t in named:
                setattr(g, name, ptr(next(vi) if t is not None else None).value)
            grads = [None] * 5
        else:
            for idx, (name, t) in enumerate(named):
                gt = None
                if t is not None and ctx.needs_input_grad[4 + idx]:
                    gt = torch.zeros(t.shape, dtype=torch.float32, device=x.device)
                setattr(g, name, ptr(gt).value)
                grads.append(gt)
        gx = torch.empty_like(x) if ctx.needs_input_grad[3] else None
        check(lib().fetode_ferro_bwd_fused(C.byref(s), ptr(x), ptr(ctx.prev), C.c_int(ctx.mode), ptr(ctx.br), ptr(gout),
                                           ptr(yout), ptr(gx), C.byref(g), C.c_int64(B), stream()), "ferro_bwd")
        return (None, None, None, gx, *grads, (acc if first else None))


def ferro_apply(module, variant: int, x2: torch.Tensor, prev: Optional[torch.Tensor], *, br=None,
                want_basis=False, want_br_out=False, want_gate=False, apply_sigmoid=False, prev_out=None,
                gate_slope=None, alpha=None, pre_bound=0.0, post_op=0, prev_mode=None):
    """Launch the neuron kernel for ``module`` (parameter names k, Ec, Ps, bias, coef | a, b).
    Returns (out, side) where side holds the optional basis / br_out / gate tensors."""
    if variant == FERRO_LOGISTIC:
        k, Ec, Ps, bias, coef = module.a, module.b, None, None, None
        I, K, O = module.a.shape[0], module.a.shape[1], 1
    else:
        k, Ec, Ps, bias, coef = module.k, module.Ec, module.Ps, module.bias, getattr(module, "coef", None)
        if variant == FERRO_FULL:
            I, O, K = k.shape
        else:
            (I, K), O = k.shape, 1
    cfg = dict(variant=variant, I=int(I), O=int(O), K=int(K),
               gate_slope=getattr(module, "gate_slope", 0.0) if gate_slope is None else gate_slope,
               alpha=getattr(module, "alpha", 0.5) if alpha is None else alpha,
               apply_sigmoid=int(apply_sigmoid), want_basis=want_basis, want_br_out=want_br_out,
               want_gate=want_gate, prev_out=prev_out, pre_bound=float(pre_bound), post_op=int(post_op))
    if prev_mode is not None:
        cfg["prev_mode"] = int(prev_mode)
    if variant == FERRO_LOGISTIC:
        cfg["gate_slope"], cfg["alpha"] = 0.0, 0.0
    params = [t for t in (k, Ec, Ps, bias, coef) if t is not None]
    handle = None
    if all(isinstance(t, torch.nn.Parameter) for t in params):
        grp = module.__dict__.get("_fetode_pgroup")
        if grp is None:
            grp = module.__dict__["_fetode_pgroup"] = ParamGroup()
        handle = grp.handle(params)
        if handle is not None and not all(t.requires_grad for t in params):
            handle = None                                     # mixed frozen / trainable: the plain route
    if handle is not None:
        det = [None if t is None else t.detach() for t in (k, Ec, Ps, bias, coef)]
        out = FerroFn.apply(cfg, prev, br, x2, *det, handle)
    else:
        out = FerroFn.apply(cfg, prev, br, x2, k, Ec, Ps, bias, coef)
    return out, cfg["side"]


def logistic_basis_apply(module, x2: torch.Tensor) -> torch.Tensor:
    """efficientkan.LogisticBasis.forward :20-24 as a stand-alone call."""
    return ferro_apply(module, FERRO_LOGISTIC, x2, None)[0]


def bspline_basis(module, x2: torch.Tensor) -> torch.Tensor:
    """KANLinear.b_splines efficientkan.py:117-131 for CUDA inputs -> [B, I, G+k]."""
    _require_cuda(x2, "input")
    B, I = x2.shape
    S = int(module.grid_size) + int(module.spline_order)
    out = torch.empty(B, I, S, dtype=torch.float32, device=x2.device)
    if B:
        check(lib().fetode_bspline_basis(ptr(f32c(module.grid)), ptr(x2), ptr(out), C.c_int64(B), C.c_int32(I),
                                         C.c_int32(int(module.grid_size)), C.c_int32(int(module.spline_order)),
                                         stream()), "bspline_basis")
    return out


# --------------------------------------------------------------------------------------
# Kuramoto lattice (csrc/kuramoto.cu)
# --------------------------------------------------------------------------------------
class KuramotoFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, K, omega, H, W, steps, dt):
        _require_cuda(x, "input")
        B = x.shape[0]
        x2 = f32c(x.reshape(B, H * W))
        feat = torch.empty(B, 2 * H * W, dtype=torch.float32, device=x.device)
        need = any(ctx.needs_input_grad[:3])
        ckpt = torch.empty(steps, B, H * W, dtype=torch.float32, device=x.device) if need else None
        Kc, om = f32c(K.detach().reshape(1)), f32c(omega.detach().reshape(-1))
        if B:
            check(lib().fetode_kuramoto_fwd(ptr(x2), ptr(Kc), ptr(om), C.c_int32(H), C.c_int32(W), C.c_int32(steps),
                                            C.c_float(dt), ptr(feat), ptr(ckpt), C.c_int64(B), stream()), "kuramoto_fwd")
        ctx.dims = (H, W, steps, dt, tuple(x.shape), tuple(K.shape), tuple(omega.shape))
        ctx.save_for_backward(ckpt if ckpt is not None else feat.new_empty(0), feat, Kc, om)
        return feat

    @staticmethod
    def backward(ctx, gfeat):
        ckpt, feat, Kc, om = ctx.saved_tensors
        H, W, steps, dt, xs, Ks, oms = ctx.dims
        B = feat.shape[0]
        gfeat = f32c(gfeat)
        gx = torch.empty(B, H * W, dtype=torch.float32, device=feat.device) if ctx.needs_input_grad[0] else None
        gK = torch.zeros(1, dtype=torch.float32, device=feat.device)
        gom = torch.zeros(H * W, dtype=torch.float32, device=feat.device)
        if B:
            check(lib().fetode_kuramoto_bwd(ptr(ckpt if steps else None), ptr(feat), ptr(Kc), ptr(om), C.c_int32(H),
                                            C.c_int32(W), C.c_int32(steps), C.c_float(dt), ptr(gfeat), ptr(gx), ptr(gK),
                                            ptr(gom), C.c_int64(B), stream()), "kuramoto_bwd")
        return (None if gx is None else gx.reshape(xs), gK.reshape(Ks), gom.reshape(oms), None, None, None, None)


def kuramoto_apply(x, K, omega, H, W, steps, dt):
    return KuramotoFn.apply(x, K, omega, int(H), int(W), int(steps), float(dt))


# --------------------------------------------------------------------------------------
# FerroelectricBasisConv2d (csrc/ferro.cu, implicit im2col around the FULL neuron kernel)
# --------------------------------------------------------------------------------------
class FerroConvFn(torch.autograd.Function):
    """One call of the convolutional ferroelectric neuron.  Parameters arrive in the kernel layout [Cin*kH*kW, Cout, K]
    (the module permutes the reference's [Cout, Cin, K, kH, kW] with torch views: layout only).  ``prev`` (previous
    image) and ``br`` are constants of the graph (ferro_class.py:719-722 detaches the snapshots)."""

    @staticmethod
    def forward(ctx, cfg, prev, br, x, k, Ec, Ps, bias, coef, out_bias):
        _require_cuda(x, "input")
        B = x.shape[0]
        g = cfg["geom"]
        L = g.Hout * g.Wout
        I, O, K = cfg["I"], cfg["O"], cfg["K"]
        out = torch.empty(B, O, g.Hout, g.Wout, dtype=torch.float32, device=x.device)
        rows = B * L
        basis = torch.empty(rows, I, O, K, dtype=torch.float32, device=x.device) if cfg.get("want_basis") else None
        br_out = torch.empty(rows, I, O, K, dtype=torch.float32, device=x.device) if cfg.get("want_br_out") else None
        if br_out is not None and br is None:
            br = torch.ones(rows, I, O, K, dtype=torch.float32, device=x.device)
        s = _ferro_struct(dict(variant=FERRO_FULL, I=I, O=O, K=K, gate_slope=cfg["gate_slope"], alpha=cfg["alpha"]),
                          k, Ec, Ps, bias, coef)
        ob = None if out_bias is None else f32c(out_bias.detach())
        if B > 0:
            check(lib().fetode_ferro_conv2d_fwd(C.byref(s), C.byref(g), ptr(x), ptr(prev), C.c_int(cfg["prev_mode"]),
                                                ptr(br), ptr(ob), ptr(out), ptr(basis), ptr(br_out),
                                                C.c_int(1 if cfg.get("br_momentum") else 0), C.c_int64(B), stream()),
                  "ferro_conv2d_fwd")
        cfg["side"] = dict(basis=basis, br_out=br_out)
        ctx.cfg, ctx.prev, ctx.br = cfg, prev, br
        ctx.has_bias = out_bias is not None
        ctx.save_for_backward(x, k, Ec, Ps, bias, coef)
        return out

    @staticmethod
    def backward(ctx, gout):
        x, k, Ec, Ps, bias, coef = ctx.saved_tensors
        cfg = ctx.cfg
        gout = f32c(gout)
        B = x.shape[0]
        s = _ferro_struct(dict(variant=FERRO_FULL, I=cfg["I"], O=cfg["O"], K=cfg["K"], gate_slope=cfg["gate_slope"],
                               alpha=cfg["alpha"]), k, Ec, Ps, bias, coef)
        g = _lib.FerroGrads()
        grads = []
        for idx, (name, t) in enumerate((("k", k), ("Ec", Ec), ("Ps", Ps), ("bias", bias), ("coef", coef))):
            gt = torch.zeros(t.shape, dtype=torch.float32, device=x.device) if ctx.needs_input_grad[4 + idx] else None
            setattr(g, name, ptr(gt).value)
            grads.append(gt)
        gob = torch.zeros(cfg["O"], dtype=torch.float32, device=x.device) if (ctx.has_bias and ctx.needs_input_grad[9]) else None
        gx = torch.empty_like(x) if ctx.needs_input_grad[3] else None
        check(lib().fetode_ferro_conv2d_bwd(C.byref(s), C.byref(cfg["geom"]), ptr(x), ptr(ctx.prev),
                                            C.c_int(cfg["prev_mode"]), ptr(ctx.br), ptr(gout), ptr(gx), C.byref(g),
                                            ptr(gob), C.c_int64(B), stream()), "ferro_conv2d_bwd")
        return (None, None, None, gx, *grads, gob)
