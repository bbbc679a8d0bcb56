"""Whole-step CUDA-graph capture for the launch-bound training loops.

The ECG step (train_ecg_kan_fet_nn_ode.py:1100-1112: forward through 8 RK4 steps, cross-entropy, backward,
``clip_grad_norm_``, AdamW) is ~1 500 kernel launches of which the wide-layer tcgen05 kernels are the only long ones:
the GPU is busy ~30 ms of a 37 ms step and the host is the limit (more so with one process per GPU on a shared
host).  ``GraphedTrainStep`` captures ONE step -- forward, loss, backward, optional flat-gradient all-reduce, clipping,
optimiser -- into a ``torch.cuda.CUDAGraph`` and replays it: no Python, no per-launch driver cost.

What makes the capture sound here:
  * every libfetode launch goes to the current (capturing) stream and allocates nothing;
  * ``fetode.optim`` with ``capturable=True`` keeps the Adam step counter on the device (fetode_adam_flat_graph) and
    the clipping norm never leaves the device;
  * solver time grids are cached tables (odeint._time_table), no per-step host->device copy;
  * the hysteresis state of the hard-gated basis (``ecg.LogisticBasis._row``: the previous call's last row, carried
    ACROSS steps in the reference, :131-132) is copied back into the tensor the captured step started from, inside the
    graph, so replay n+1 starts from the state replay n left -- exactly the eager sequence;
  * all parameter versions are bumped before capture, so the packed weight images are rebuilt INSIDE the graph
    (a replayed step must repack after the optimiser changed the parameters).

Not supported under capture (they synchronise with the host by construction): dopri5 (adaptive accept / reject) and
``update_grid``.  The soft ferroelectric bases work when the step resets their state itself (the Lotka-Volterra loop calls
``reset_state()`` every iteration, so no state crosses a replay boundary); state that must survive from one step to the
next is only carried for the hard-gated ``ecg.LogisticBasis``.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from . import dist as fdist
from .optim import FlatAdam


def _stateful(model):
    from .ecg import LogisticBasis
    return [m for m in model.modules() if isinstance(m, LogisticBasis)]


class GraphedTrainStep:
    """step = GraphedTrainStep(model, opt, loss_fn, x, y, max_norm=1.0); loss = step(x, y)

    ``opt`` must be a ``fetode.optim`` optimiser built with ``capturable=True``.  ``x`` / ``y`` fix the shapes; later
    calls copy their arguments into the captured input buffers.  The returned loss is a device tensor that the next
    call overwrites.  ``forward(model, x)`` replaces ``model(x)`` when the step is not a plain module call (e.g. the
    Lotka-Volterra loop: ``reset_state()`` followed by ``odeint`` over the model, train_kanfet_node_predprey.py:250-257;
    keep the time grid on the host).  The warm-up iterations are real optimiser steps on (x, y).  ``allreduce=True`` puts fetode.dist's flat-gradient all-reduce inside the graph (NCCL)."""

    def __init__(self, model: torch.nn.Module, opt: FlatAdam, loss_fn: Callable, x: torch.Tensor, y: torch.Tensor,
                 max_norm: Optional[float] = None, allreduce: bool = False, warmup: int = 3,
                 forward: Optional[Callable] = None):
        if not isinstance(opt, FlatAdam) or not opt.capturable:
            raise ValueError("GraphedTrainStep needs a fetode.optim optimiser built with capturable=True")
        if not x.is_cuda:
            raise RuntimeError("GraphedTrainStep: inputs must live on a CUDA device (there is no CPU path)")
        self.model, self.opt, self.loss_fn = model, opt, loss_fn
        self.forward = forward if forward is not None else (lambda m, inp: m(inp))
        self.max_norm, self.allreduce = max_norm, bool(allreduce)
        self.x, self.y = x.clone(), y.clone()
        self._params = [p for p in model.parameters()]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                       # eager steps: builds the flat buffers, warms every cache
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        for p in model.parameters():                        # force the weight images to be repacked inside the graph
            torch.autograd.graph.increment_version(p)
        mods = _stateful(model)
        start = [m._row for m in mods]
        if any(r is None for r in start):
            raise RuntimeError("GraphedTrainStep: a hard-gated basis was never called during warm-up")
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._body()
            for m, r0 in zip(mods, start):                  # carry the hysteresis state to the next replay
                r0.copy_(m._row)
        self.opt.step_count -= 1                            # capturing ran step()'s host side, not its kernels
        for m, r0 in zip(mods, start):
            m._row = r0
            m.prev_x = r0.view(1, -1, 1).expand(1, m.in_dim, m.num_basis)
        self.replays = 0

    def _body(self) -> torch.Tensor:
        self.opt.zero_grad()
        loss = self.loss_fn(self.forward(self.model, self.x), self.y)
        loss.backward()
        if self.allreduce:
            fdist.allreduce_gradients(self.opt)
        if self.max_norm is not None:
            self.opt.clip_grad_norm_(self.max_norm)
        self.opt.step()
        return loss.detach()

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if x is not self.x:
            self.x.copy_(x, non_blocking=True)
        if y is not self.y:
            self.y.copy_(y, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        self.opt.step_count += 1
        # the replay changed the parameters behind autograd's back: bump their version counters (host only) so that
        # every version-keyed cache (LayerPack.blob, SmallNetPlan.packed, MLPKANODEFunc._gain) is rebuilt by the next
        # EAGER call, e.g. a validation pass between replays
        torch.autograd.graph.increment_version(self._params)
        return self.loss


def _ferro_holders(modules):
    """(module, attribute) pairs that hold hysteresis state tensors: ferro_class neurons (``_rows``) and the hard-gated
    basis (``_row``)."""
    out = []
    for root in modules:
        for m in root.modules():
            if hasattr(m, "_rows") and hasattr(m, "_set_rows"):
                out.append((m, "_rows"))
            elif hasattr(m, "_row") and hasattr(m, "branch_breaking_point"):
                out.append((m, "_row"))
    return out


class GraphedCall:
    """``out = fn(*inputs)`` under ``torch.no_grad()``, captured ONCE into a CUDA graph and replayed: the launch-bound
    loops around the hot path (the 250 reverse-diffusion steps of ``p_sample_loop`` over the KANFET stack,
    kan_diffusion_ett.py:565-609, ~40 launches each) cost one graph launch per iteration.

    The hysteresis state of every stateful neuron under ``modules`` is carried from replay to replay exactly as the
    eager sequence carries it: the captured call reads the state tensors that exist after the warm-up calls, and their
    new values are copied back into those tensors inside the graph.  ``reset_state()`` zero-fills them (== the modules'
    own reset: a zero previous input is what a fresh buffer means)."""

    def __init__(self, fn: Callable, inputs, modules=(), warmup: int = 2):
        self.inputs = [t.clone() for t in inputs]
        if not all(t.is_cuda for t in self.inputs):
            raise RuntimeError("GraphedCall: inputs must live on a CUDA device (there is no CPU path)")
        self.fn = fn
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                fn(*self.inputs)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.holders = _ferro_holders(modules)
        start = []
        for m, attr in self.holders:
            t = getattr(m, attr)
            if t is None:
                raise RuntimeError("GraphedCall: a stateful neuron under `modules` was not called by fn during warm-up")
            start.append(t)
        self.state = start
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.out = fn(*self.inputs)
            for (m, attr), r0 in zip(self.holders, start):
                r0.copy_(getattr(m, attr))
        self._rebind()
        self.replays = 0

    def _rebind(self):
        for (m, attr), r0 in zip(self.holders, self.state):
            if attr == "_rows":
                m._set_rows(r0)
            else:
                m._row = r0
                m.prev_x = r0.view(1, -1, 1).expand(1, m.in_dim, m.num_basis)

    def reset_state(self):
        for r0 in self.state:
            r0.zero_()

    def __call__(self, *inputs):
        for dst, src in zip(self.inputs, inputs):
            if src is not dst:
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        return self.out
