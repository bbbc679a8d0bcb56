"""Drop-in for the reference's ``ferro_class`` module: the ferroelectric hysteresis neurons.

Same class names, constructor arguments, parameter / buffer names and call semantics as
ferro_class.py (reference checkout, commit 704a0df4); the arithmetic runs in libfetode.so
(csrc/ferro.cu, and inside the fused KAN kernels for the 2-D variant).  Class names matter:
``reset_stateful_ferro_buffers`` (train_ecg.py:371-378) matches modules by
``__class__.__name__`` and then zeroes ``prev_x`` / fills ``branch_sign`` in place.

State representation.  The reference keeps ``prev_x`` as the previous input *expanded* over the
basis axes ([B,I,O,K] / [B,I,K]).  Here the state is the compact previous input ``[B, I]`` and
the ``prev_x`` buffer is a stride-0 expanded view of it, so ``prev_x.zero_()`` from user code
still resets the real state without ever materialising B*I*O*K floats in HBM.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops


class _FerroState:
    """Mixin: compact previous-input rows + the reference-shaped buffer views."""

    def _init_state(self, shape_tail):
        self._tail = tuple(shape_tail)
        self.register_buffer("prev_x", torch.zeros(1, *self._tail))
        self.register_buffer("branch_sign", torch.ones(1, *self._tail))
        self._rows: Optional[torch.Tensor] = None       # [B, I]; None == reference's fresh [1,...] zeros

    def _set_rows(self, rows: torch.Tensor) -> None:
        """Record x as the new previous input (ferro_class.py:409 / :584-590).  ``rows`` must be OWNED by the module
        (the reference stores ``x.detach().clone()``): callers pass a copy of the user's input or a state tensor a
        fused kernel wrote, never the input itself -- ``prev_x.zero_()`` from user code then cannot reach the data."""
        self._rows = rows
        view = rows
        for _ in range(len(self._tail) - 1):
            view = view.unsqueeze(-1)
        self.prev_x = view.expand(rows.shape[0], *self._tail)

    def _prev_rows(self, B: int, device) -> Optional[torch.Tensor]:
        """Previous input to use for a batch of B rows, or None for 'all zeros'."""
        r = self._rows
        if r is None:
            # fresh / reset buffers: [1,...] zeros broadcast against any B.  A user may also have
            # written into the registered buffer directly (load_state_dict): honour a non-zero
            # [1,...] buffer by broadcasting its first basis column.
            return None
        if r.shape[0] != B:
            if r.shape[0] == 1:
                return r.expand(B, -1).contiguous()
            raise RuntimeError(
                f"ferro state holds {r.shape[0]} rows but the input has {B}: call reset_state() "
                "when the batch size changes (the reference fails on the same broadcast)")
        return r if r.is_contiguous() else r.contiguous()

    def reset_state(self):
        """ferro_class.py:422-424 / :594-596."""
        dev = self.prev_x.device
        self._rows = None
        self.prev_x = torch.zeros(1, *self._tail, device=dev)
        self.branch_sign = torch.ones(1, *self._tail, device=dev)

    def _apply(self, fn, *a, **k):                      # keep the compact rows with the module on .to()
        out = super()._apply(fn, *a, **k)
        if self._rows is not None:
            self._set_rows(fn(self._rows))
        return out

    # ---- checkpoint restore (train_kan_fet_ett.py:353-358 reloads the best state_dict, buffers included) ----
    # ``prev_x`` is a stride-0 view after any forward: copy_() into it is illegal, and the incoming buffer has the
    # reference's expanded shape [B, *tail].  Receive it in a real tensor of the incoming shape, then fold it back
    # into the compact rows (its first basis column).
    def _load_from_state_dict(self, state_dict, prefix, *args):
        for name in ("prev_x", "branch_sign"):
            t = state_dict.get(prefix + name)
            cur = self._buffers.get(name)
            if t is not None and cur is not None:
                self._buffers[name] = torch.empty(tuple(t.shape), dtype=cur.dtype, device=cur.device)
        super()._load_from_state_dict(state_dict, prefix, *args)
        self._adopt_loaded_buffers()

    def _adopt_loaded_buffers(self):
        p = self._buffers["prev_x"]
        if p.dim() != len(self._tail) + 1 or tuple(p.shape[1:]) != self._tail:
            return                                       # size mismatch: the base class has reported it
        rows = p[(slice(None), slice(None)) + (0,) * (len(self._tail) - 1)]
        if p.shape[0] == 1 and not bool((rows != 0).any()):
            self._rows = None                            # the reference's fresh [1, ...] zeros
            self.prev_x = torch.zeros(1, *self._tail, device=p.device)
        else:
            self._set_rows(rows.clone())
        br = self._buffers["branch_sign"]
        if hasattr(self, "_br_live"):
            self._br_live = bool(getattr(self, "_update_branch", False) and br.dim() == len(self._tail) + 1
                                 and (br.shape[0] > 1 or bool((br != 1).any())))


class TwoDimensionFerroelectricBasis(_FerroState, nn.Module):
    """ferro_class.py:526-596.  Used in the ``logistic_basis`` slot of KANLinear to form KANFET;
    returns ``basis * coef`` of shape [B, in_dim, num_basis]."""

    def __init__(self, in_dim, num_basis, use_noise=False, gate_slope=10.0, alpha=0.8, noise_std=0.05):
        super().__init__()
        self.in_dim, self.num_basis = in_dim, num_basis
        self.use_noise, self.noise_std = use_noise, noise_std
        self.gate_slope, self.alpha = gate_slope, alpha
        self.k = nn.Parameter(torch.rand(in_dim, num_basis) * 2 + 0.5)
        self.Ec = nn.Parameter(torch.rand(in_dim, num_basis) * 2 + 0.5)
        self.Ps = nn.Parameter(torch.rand(in_dim, num_basis) * 1.5 + 0.5)
        self.bias = nn.Parameter(torch.randn(in_dim, num_basis) * 0.1)
        self.coef = nn.Parameter(torch.randn(in_dim, num_basis))
        self._init_state((in_dim, num_basis))

    def forward(self, x, return_activations=False):
        ops.abort_if_tracing()
        if x.dim() > 2:
            x = x.view(x.shape[0], -1)
        x2 = ops.f32c(x)
        prev = self._prev_rows(x2.shape[0], x2.device)
        out, _ = ops.ferro_apply(self, ops.FERRO_TWO_D, x2, prev)
        if self.use_noise:
            out = out + (torch.randn_like(out) * self.noise_std).detach() * self.coef
        self._set_rows(x2.detach().clone())
        return out


class _FullFerro(_FerroState, nn.Module):
    """Shared body of the [in, out, num_basis] neurons (ferro_class.py:329-523)."""

    _always_noise = False
    _update_branch = False

    def __init__(self, in_dim, out_dim, num_basis, use_noise=False, gate_slope=10.0, alpha=0.8, noise_std=0.05):
        super().__init__()
        self.in_dim, self.out_dim, self.num_basis = in_dim, out_dim, num_basis
        self.use_noise, self.noise_std = use_noise, noise_std
        self.gate_slope, self.alpha = gate_slope, alpha
        shape = (in_dim, out_dim, num_basis)
        self.k = nn.Parameter(torch.rand(*shape) * 2 + 0.5)
        self.Ec = nn.Parameter(torch.rand(*shape) * 2 + 0.5)
        self.Ps = nn.Parameter(torch.rand(*shape) * 1.5 + 0.5)
        self.bias = nn.Parameter(torch.randn(*shape) * 0.1)
        self.coef = nn.Parameter(torch.randn(*shape))
        self._init_state(shape)

    # first-call rule of the reference (:373-378): buffers whose shape is not [B,I,O,K] are
    # re-created as prev = x (dx = 0), branch_sign = 1; the initial [1,...] zeros match only B == 1.
    def _full_prev(self, x2: torch.Tensor):
        B = x2.shape[0]
        r = self._rows
        if r is None:
            return None if B == 1 else x2
        if r.shape[0] != B:
            return x2
        return r if r.is_contiguous() else r.contiguous()

    def _full_branch(self, B: int):
        br = self.branch_sign
        if br.dim() == 4 and br.shape[0] == B and (B != 1 or self._br_live):
            return br if br.is_contiguous() else br.contiguous()
        return None                                     # all ones

    def reset_state(self):
        """ferro_class.py:422-424: zero_() / fill_(1) IN PLACE -- the batch shape of the buffers survives."""
        if self._rows is not None:
            self._set_rows(torch.zeros_like(self._rows))
        else:
            self.prev_x = torch.zeros(1, *self._tail, device=self.prev_x.device)
        if self._br_live:
            self.branch_sign.fill_(1.0)
        else:
            self.branch_sign = torch.ones(1, *self._tail, device=self.prev_x.device)

    _br_live = False

    def forward(self, x, return_activations=False):
        ops.abort_if_tracing()
        if x.dim() > 2:
            x = x.view(x.shape[0], -1)
        x2 = ops.f32c(x)
        B = x2.shape[0]
        prev = self._full_prev(x2.detach())
        br = self._full_branch(B) if self._update_branch else None
        noisy = (self._always_noise or self.use_noise) and self.noise_std != 0.0
        want_basis = return_activations or noisy
        out, side = ops.ferro_apply(self, ops.FERRO_FULL, x2, prev, br=br, want_basis=want_basis,
                                    want_br_out=self._update_branch)
        basis = side["basis"]
        if noisy:
            # basis + noise.detach()  (:410-412 / :508-509): the noise enters the output through coef only
            noise = torch.randn_like(basis) * self.noise_std
            out = out + (noise * self.coef).sum(dim=(1, 3))
            basis = basis + noise
        self._set_rows(x2.detach().clone())
        if self._update_branch:                          # :512-514
            self.branch_sign = side["br_out"]
            self._br_live = True
        if return_activations:
            return out, basis, self.coef.detach()
        return out


class FerroelectricBasis(_FullFerro):
    """ferro_class.py:329-424.  branch_sign is never written by the reference => it stays 1."""


class NoisyFerroelectricBasis(_FullFerro):
    """ferro_class.py:427-523: noise is added unconditionally (:508-509) and branch_sign <- target_sign
    (:512-514), a real [B,I,O,K] hysteresis state."""
    _always_noise = True
    _update_branch = True

    def __init__(self, in_dim, out_dim, num_basis, use_noise=True, gate_slope=10.0, alpha=0.8, noise_std=0.2):
        super().__init__(in_dim, out_dim, num_basis, use_noise, gate_slope, alpha, noise_std)


class BatchedFerroelectricBasis(_FullFerro):
    """ferro_class.py:107-214: per-sample loop in which row b sees row b-1 as its previous input
    (prev_x is [1,I,O,K] and is overwritten after every sample); branch_sign is never written.
    One launch: prev = [stored row ; x[:-1]]."""

    def forward(self, x, return_activations=False):
        ops.abort_if_tracing()
        if x.dim() > 2:
            x = x.view(x.shape[0], -1)
        x2 = ops.f32c(x)
        B = x2.shape[0]
        if B == 0:
            return x2.new_empty(0, self.out_dim)
        first = self._rows if self._rows is not None else torch.zeros(1, self.in_dim, device=x2.device)
        prev = torch.cat([first[-1:].to(x2.device), x2.detach()[:-1]], dim=0).contiguous()
        noisy = (self._always_noise or self.use_noise) and self.noise_std != 0.0
        out, side = ops.ferro_apply(self, ops.FERRO_FULL, x2, prev, want_basis=return_activations or noisy)
        basis = side["basis"]
        if noisy:
            noise = torch.randn_like(basis) * self.noise_std
            out = out + (noise * self.coef).sum(dim=(1, 3))
            basis = basis + noise
        self._set_rows(x2.detach()[-1:].clone())
        if return_activations:
            return out, basis, self.coef.detach()
        return out


class NoisyBatchedFerroelectricBasis(BatchedFerroelectricBasis):
    """ferro_class.py:220-326: BatchedFerroelectricBasis with unconditional noise."""
    _always_noise = True

    def __init__(self, in_dim, out_dim, num_basis, use_noise=True, gate_slope=10.0, alpha=0.8, noise_std=0.2):
        super().__init__(in_dim, out_dim, num_basis, use_noise, gate_slope, alpha, noise_std)


class FerroelectricBasisConv2d(nn.Module):
    """ferro_class.py:601-779: the ferroelectric neuron over the kH x kW patches of an image.

    Same constructor arguments and parameter names / shapes ([Cout, Cin, K, kH, kW], ``out_bias`` [Cout]) as the
    reference.  One kernel launch gathers the patches from the image (implicit im2col, zero padding) -- neither
    ``F.unfold``'s (B, Cin*kH*kW, L) matrix nor the (B,Cout,Cin,K,kH,kW,L) broadcast exists in HBM.
    State: stateless by default (prev = x, branch = +1 on every call, :724-727).  ``stateful=True`` keeps the previous
    IMAGE (the reference keeps its unfolded expansion) and, for this class, the branch momentum; as in the reference
    (whose branch buffer has no Cout axis, :686 vs :767) that only works for ``out_channels == 1`` and raises
    RuntimeError otherwise."""

    _mem_efficient = False

    def __init__(self, in_channels, out_channels, kernel_size, num_basis=3, stride=1, padding=0, use_noise=False,
                 noise_std=0.2, gate_slope=10.0, alpha=0.8, stateful=False):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size = kernel_size if isinstance(kernel_size, tuple) else (kernel_size, kernel_size)
        self.num_basis, self.stride, self.padding = num_basis, stride, padding
        self.use_noise, self.noise_std = use_noise, noise_std
        self.gate_slope, self.alpha, self.stateful = gate_slope, alpha, stateful
        kH, kW = self.kernel_size
        shape = (out_channels, in_channels, num_basis, kH, kW)
        self.k = nn.Parameter(torch.rand(*shape) * 2.0 + 0.5)
        self.Ec = nn.Parameter(torch.rand(*shape) * 2.0 + 0.5)
        self.Ps = nn.Parameter(torch.rand(*shape) * 1.5 + 0.5)
        self.bias = nn.Parameter(torch.randn(*shape) * 0.1)
        self.coef = nn.Parameter(torch.randn(*shape))
        self.out_bias = nn.Parameter(torch.zeros(out_channels))
        self.register_buffer("_prev_x", torch.zeros(1))          # compact: the previous image once stateful
        self.register_buffer("_branch_sign", torch.ones(1))      # [B*L, Cin*kH*kW, Cout, K] once stateful (plain class)
        self._state_ready = False

    def reset_state(self):
        self._state_ready = False

    def _kernel_params(self):
        """[Cout, Cin, K, kH, kW] -> [(Cin, kH, kW), Cout, K]: layout only, autograd carries the gradients back."""
        I = self.in_channels * self.kernel_size[0] * self.kernel_size[1]
        return [p.permute(1, 3, 4, 0, 2).reshape(I, self.out_channels, self.num_basis).contiguous()
                for p in (self.k, self.Ec, self.Ps, self.bias, self.coef)]

    def forward(self, x, return_activations=False):
        from ._lib import Conv2d
        ops.abort_if_tracing()
        B, Cin, H, W = x.shape
        assert Cin == self.in_channels
        kH, kW = self.kernel_size
        x = ops.f32c(x)
        g = Conv2d(Cin, H, W, kH, kW, self.stride, self.padding,
                   (H + 2 * self.padding - kH) // self.stride + 1, (W + 2 * self.padding - kW) // self.stride + 1)
        L = g.Hout * g.Wout
        I, O, K = Cin * kH * kW, self.out_channels, self.num_basis
        prev, br, mode = None, None, 2
        track_br = self.stateful and not self._mem_efficient
        if self.stateful:
            if track_br and O != 1:
                raise RuntimeError(f"output with shape {[B, 1, Cin, K, kH, kW, L]} doesn't match the broadcast shape "
                                   f"{[B, O, Cin, K, kH, kW, L]} (the reference's stateful branch buffer has no Cout axis)")
            ready = self._state_ready and tuple(self._prev_x.shape) == tuple(x.shape) and self._prev_x.device == x.device
            if ready:
                prev, mode = self._prev_x, 0
                br = self._branch_sign if track_br else None
        noisy = self.use_noise and self.noise_std != 0.0
        want_basis = (return_activations and not self._mem_efficient) or noisy
        cfg = dict(geom=g, I=I, O=O, K=K, gate_slope=float(self.gate_slope), alpha=float(self.alpha), prev_mode=mode,
                   want_basis=want_basis, want_br_out=track_br, br_momentum=True)
        out = ops.FerroConvFn.apply(cfg, prev, br, x, *self._kernel_params(), self.out_bias)
        side = cfg["side"]
        basis = side["basis"]
        if noisy:
            # basis + noise.detach() (:753-755): the noise reaches the output through coef only
            noise = torch.randn_like(basis) * self.noise_std
            coef_k = self._kernel_params()[4]
            out = out + (noise * coef_k).sum(dim=(1, 3)).view(B, L, O).permute(0, 2, 1).reshape(B, O, g.Hout, g.Wout)
            basis = basis + noise
        if self.stateful:
            self._prev_x = x.detach().clone()
            if track_br:
                self._branch_sign = side["br_out"]
            self._state_ready = True
        if return_activations:
            if self._mem_efficient:
                return out, None, None
            # [B*L, (Cin,kH,kW), Cout, K] -> the reference's (B, Cout, Cin, K, kH, kW, L)
            b7 = basis.detach().view(B, L, Cin, kH, kW, O, K).permute(0, 5, 2, 6, 3, 4, 1)
            return out, b7, self.coef.detach().unsqueeze(0).unsqueeze(-1)
        return out


class MemEfficient_FerroelectricBasisConv2d(FerroelectricBasisConv2d):
    """ferro_class.py:783-944: the same function evaluated in Cout chunks by the reference; here it IS the same
    kernel (nothing is materialised to begin with, so ``out_chunk`` is accepted and ignored).  Stateful mode keeps the
    previous image only: the reference's Cout-free branch buffer stays +1 (:927-931)."""

    _mem_efficient = True

    def __init__(self, in_channels, out_channels, kernel_size, num_basis=3, stride=1, padding=0, use_noise=False,
                 noise_std=0.05, gate_slope=10.0, alpha=0.8, stateful=False, out_chunk=8):
        super().__init__(in_channels, out_channels, kernel_size, num_basis, stride, padding, use_noise, noise_std,
                         gate_slope, alpha, stateful)
        self.out_chunk = out_chunk


class _FerroSeqFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, prev0, br0, k, Ec, Ps, bias, coef, want_basis):
        import ctypes as C
        from ._lib import check, lib, ptr, stream
        ops._require_cuda(x, "input")
        B, I = x.shape
        _, O, K = k.shape
        need = any(ctx.needs_input_grad)
        out = torch.empty(B, O, dtype=torch.float32, device=x.device)
        basis = torch.empty(B, I, O, K, dtype=torch.float32, device=x.device) if want_basis else None
        hist = torch.empty(B, I, O, K, dtype=torch.float32, device=x.device) if need else None
        br_final = torch.empty(I, O, K, dtype=torch.float32, device=x.device)
        s = ops._ferro_struct(dict(variant=ops.FERRO_FULL, I=I, O=O, K=K, gate_slope=mod.gate_slope, alpha=mod.alpha),
                              k, Ec, Ps, bias, coef)
        if B:
            check(lib().fetode_ferro_seq_fwd(C.byref(s), ptr(x), ptr(prev0), ptr(br0), ptr(out), ptr(basis), ptr(hist),
                                             ptr(br_final), C.c_int64(B), stream()), "ferro_seq_fwd")
        else:
            br_final = br0.clone() if br0 is not None else torch.ones(I, O, K, device=x.device)
        ctx.mod, ctx.prev0 = mod, prev0
        ctx.save_for_backward(x, k, Ec, Ps, bias, coef, hist if hist is not None else x.new_empty(0))
        ctx.mark_non_differentiable(br_final)
        if basis is not None:
            ctx.mark_non_differentiable(basis)
            return out, br_final, basis
        return out, br_final, x.new_empty(0)

    @staticmethod
    def backward(ctx, gout, _gbr, _gbasis):
        import ctypes as C
        from . import _lib
        from ._lib import check, lib, ptr, stream
        x, k, Ec, Ps, bias, coef, hist = ctx.saved_tensors
        B, I = x.shape
        _, O, K = k.shape
        mod = ctx.mod
        gout = ops.f32c(gout)
        s = ops._ferro_struct(dict(variant=ops.FERRO_FULL, I=I, O=O, K=K, gate_slope=mod.gate_slope, alpha=mod.alpha),
                              k, Ec, Ps, bias, coef)
        g = _lib.FerroGrads()
        grads = []
        for idx, (name, t) in enumerate((("k", k), ("Ec", Ec), ("Ps", Ps), ("bias", bias), ("coef", coef))):
            gt = torch.zeros(t.shape, dtype=torch.float32, device=x.device) if ctx.needs_input_grad[4 + idx] else None
            setattr(g, name, ptr(gt).value)
            grads.append(gt)
        gx = torch.empty_like(x) if ctx.needs_input_grad[1] else None
        check(lib().fetode_ferro_seq_bwd(C.byref(s), ptr(x), ptr(ctx.prev0), ptr(hist), ptr(gout), ptr(gx), C.byref(g),
                                         C.c_int64(B), stream()), "ferro_seq_bwd")
        return (None, gx, None, None, *grads, None)


class OriginalBatchedFerroelectricBasis(nn.Module):
    """ferro_class.py:9-104: the per-sample loop in which sample b sees the previous ROW and the branch momentum it
    left, and that momentum -- not a detached snapshot -- enters the shifted field, so gradients flow through the
    recurrence over the batch.  One launch runs the scan (csrc/ferro.cu, ferro_seq_*).  Buffers ``prev_x`` /
    ``branch_sign`` keep the reference's [1, I, O, K] shapes; the state is detached between calls (the reference leaves
    the previous call's graph attached to the buffer, which only makes a second backward fail)."""

    def __init__(self, in_dim, out_dim, num_basis, use_noise=False, gate_slope=10.0, alpha=0.8, noise_std=0.05):
        super().__init__()
        self.in_dim, self.out_dim, self.num_basis = in_dim, out_dim, num_basis
        self.use_noise, self.noise_std = use_noise, noise_std
        self.gate_slope, self.alpha = gate_slope, alpha
        shape = (in_dim, out_dim, num_basis)
        self.k = nn.Parameter(torch.rand(*shape) * 2 + 0.5)
        self.Ec = nn.Parameter(torch.rand(*shape) * 2 + 0.5)
        self.Ps = nn.Parameter(torch.rand(*shape) * 1.5 + 0.5)
        self.bias = nn.Parameter(torch.randn(*shape) * 0.1)
        self.coef = nn.Parameter(torch.randn(*shape))
        self.register_buffer("prev_x", torch.zeros(1, *shape))
        self.register_buffer("branch_sign", torch.ones(1, *shape))

    def forward(self, x, return_activations=False):
        ops.abort_if_tracing()
        if x.dim() > 2:
            x = x.view(x.shape[0], -1)
        x2 = ops.f32c(x)
        B = x2.shape[0]
        if B == 0:
            return x2.new_empty(0, self.out_dim)
        prev0 = self.prev_x[0, :, 0, 0].contiguous()                 # the buffer is x expanded over (O, K)
        br0 = self.branch_sign[0].contiguous()
        noisy = self.use_noise and self.noise_std != 0.0
        out, br_final, basis = _FerroSeqFn.apply(self, x2, prev0, br0, self.k, self.Ec, self.Ps, self.bias, self.coef,
                                                 bool(return_activations or noisy))
        if noisy:
            noise = torch.randn_like(basis) * self.noise_std         # basis + noise.detach()  (:82-84)
            out = out + (noise * self.coef).sum(dim=(1, 3))
            basis = basis + noise
        with torch.no_grad():
            self.prev_x = x2.detach()[-1].view(1, -1, 1, 1).expand(1, self.in_dim, self.out_dim, self.num_basis).clone()
            self.branch_sign = br_final.unsqueeze(0)
        if return_activations:
            return out, basis.detach(), self.coef.detach()
        return out

    def reset_state(self):
        self.prev_x.zero_()
        self.branch_sign.fill_(1.0)
