"""Fused Adam / AdamW on one flat fp32 buffer (csrc/optim.cu).

Drop-in for the optimisers the reference's training loops build,
    torch.optim.Adam(model.parameters(), lr=2e-3)                       train_kanfet_node_predprey.py:171
    torch.optim.AdamW(...) + torch.nn.utils.clip_grad_norm_(.., 1.0)    train_ecg_kan_fet_nn_ode.py:1093, 1110
with the same hyper-parameter names and update rule (torch/optim/adam.py, amsgrad=False).  Parameters, gradients and
moments are re-homed into flat buffers (every ``p.data`` / ``p.grad`` becomes a view), so that

  * ``zero_grad()`` is one fill, ``step()`` is one launch (two with clipping), whatever the number of tensors;
  * the data-parallel gradient all-reduce (fetode.dist) runs on ``flat_grad`` directly, no gather / scatter copies;
  * ``clip_grad_norm_`` costs no host synchronisation: the norm stays on the device and the update kernel reads it.

Plumbing only: the arithmetic is in libfetode.so; there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Optional

import torch

from ._lib import check, lib, ptr, stream


class FlatAdam:
    """The flat buffers are built at the first ``step()`` / ``clip_grad_norm_()``, from the parameters that actually
    received a gradient in the first backward pass: torch's Adam never touches a parameter whose ``.grad`` is None
    (e.g. the unused ``coef`` of the hard-gated LogisticBasis, train_ecg_kan_fet_nn_ode.py:91), and neither does this."""

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, decoupled: bool = False, capturable: bool = False):
        self.all_params = [p for p in params if p.requires_grad]
        if not self.all_params:
            raise ValueError("FlatAdam: no parameters to optimise")
        dev = self.all_params[0].device
        if dev.type != "cuda":
            raise RuntimeError("fetode.optim: parameters must live on a CUDA device (there is no CPU path)")
        for p in self.all_params:
            if p.dtype != torch.float32 or p.device != dev:
                raise ValueError("FlatAdam: all parameters must be fp32 on one device")
        self.device = dev
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.weight_decay, self.decoupled = float(weight_decay), bool(decoupled)
        self.step_count = 0
        # capturable: the step counter and bias corrections live on the device (fetode_adam_flat_graph), so a captured
        # step() replays correctly in a CUDA graph (fetode.graphs.GraphedTrainStep); lr / betas are baked in at capture
        self.capturable = bool(capturable)
        self.params = None                 # active parameters, once built
        self.flat_param = self.flat_grad = self.exp_avg = self.exp_avg_sq = None
        self._max_norm: Optional[float] = None

    @property
    def built(self) -> bool:
        return self.params is not None

    def _build(self) -> None:
        self.params = [p for p in self.all_params if p.grad is not None]
        if not self.params:
            raise RuntimeError("FlatAdam.step(): no parameter has a gradient (call backward() first)")
        self._idle = [p for p in self.all_params if p.grad is None]
        self._offsets, off = [], 0
        for p in self.params:
            self._offsets.append(off)
            off += (p.numel() + 3) // 4 * 4                      # keep every view 16-byte aligned
        self.n = off
        dev = self.device
        self.flat_param = torch.zeros(off, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)
        self._sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
        self._dev_state = torch.zeros(2, dtype=torch.int64, device=dev)      # {step; (step_size, bc2_sqrt) as 2 x f32}
        if self.capturable and self.step_count:
            self._dev_state[0] = self.step_count
        with torch.no_grad():
            for p, o in zip(self.params, self._offsets):
                pv = self.flat_param[o:o + p.numel()].view(p.shape)
                pv.copy_(p.data)
                p.data = pv
                gv = self.flat_grad[o:o + p.numel()].view(p.shape)
                gv.copy_(p.grad)
                p.grad = gv

    def _sync_grads(self) -> None:
        """Gradients normally accumulate straight into the attached views; anything else (grads dropped by
        model.zero_grad(set_to_none=True), replaced tensors) is copied in, and a parameter outside the flat set that
        suddenly has a gradient is an error."""
        if not self.built:
            self._build()
            return
        base = self.flat_grad.data_ptr()
        with torch.no_grad():
            for p, o in zip(self.params, self._offsets):
                g = p.grad
                if g is None:
                    self.flat_grad[o:o + p.numel()].zero_()
                    p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)
                elif g.data_ptr() != base + 4 * o:
                    gv = self.flat_grad[o:o + p.numel()].view(p.shape)
                    gv.copy_(g)
                    p.grad = gv
        for p in self._idle:
            if p.grad is not None:
                raise RuntimeError("FlatAdam: a parameter that had no gradient at the first step now has one; "
                                   "build a new optimiser (the flat layout is fixed at the first step)")

    def zero_grad(self, set_to_none: bool = True) -> None:
        """Before the first step: drops the gradients (torch's default).  Afterwards: ONE fill of the flat gradient
        buffer; the views stay attached so that the next backward accumulates straight into it."""
        if not self.built:
            for p in self.all_params:
                p.grad = None
            return
        self.flat_grad.zero_()
        base = self.flat_grad.data_ptr()
        for p, o in zip(self.params, self._offsets):
            if p.grad is None or p.grad.data_ptr() != base + 4 * o:
                p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)

    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        """clip_grad_norm_(params, max_norm) semantics (L2 norm over all gradients); the scaling itself is applied by the
        next step() launch.  Returns the total norm as a 0-dim device tensor (no synchronisation)."""
        self._sync_grads()
        check(lib().fetode_sumsq_f32(ptr(self.flat_grad), C.c_int64(self.n), C.c_void_p(self._sumsq.data_ptr()),
                                     stream()), "sumsq_f32")
        self._max_norm = float(max_norm)
        return self._sumsq[0].sqrt().float()

    @torch.no_grad()
    def step(self) -> None:
        if self._max_norm is None:
            self._sync_grads()
        self.step_count += 1
        clip = self._max_norm is not None
        if self.capturable:
            check(lib().fetode_adam_flat_graph(
                ptr(self.flat_param), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq), C.c_int64(self.n),
                C.c_float(self.lr), C.c_float(self.betas[0]), C.c_float(self.betas[1]), C.c_float(self.eps),
                C.c_float(self.weight_decay), C.c_int(1 if self.decoupled else 0),
                C.c_void_p(self._dev_state.data_ptr()), C.c_float(self._max_norm if clip else 0.0),
                C.c_void_p(self._sumsq.data_ptr() if clip else 0), stream()), "adam_flat_graph")
            self._max_norm = None
            for p in self.params:
                torch.autograd.graph.increment_version(p)
            return
        check(lib().fetode_adam_flat(ptr(self.flat_param), ptr(self.flat_grad), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                     C.c_int64(self.n), C.c_float(self.lr), C.c_float(self.betas[0]),
                                     C.c_float(self.betas[1]), C.c_float(self.eps), C.c_float(self.weight_decay),
                                     C.c_int(1 if self.decoupled else 0), C.c_int64(self.step_count),
                                     C.c_float(self._max_norm if clip else 0.0),
                                     C.c_void_p(self._sumsq.data_ptr() if clip else 0), stream()), "adam_flat")
        self._max_norm = None
        # the kernel wrote through raw pointers: bump the version counters so that the packed layer images
        # (ops.LayerPack / SmallNetPlan keys) are rebuilt on the next forward
        for p in self.params:
            torch.autograd.graph.increment_version(p)


def Adam(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, capturable=False) -> FlatAdam:
    """torch.optim.Adam semantics (weight_decay is an L2 term added to the gradient)."""
    return FlatAdam(params, lr, betas, eps, weight_decay, decoupled=False, capturable=capturable)


def AdamW(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, capturable=False) -> FlatAdam:
    """torch.optim.AdamW semantics (decoupled weight decay)."""
    return FlatAdam(params, lr, betas, eps, weight_decay, decoupled=True, capturable=capturable)
