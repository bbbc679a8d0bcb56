"""Batch-sharded data parallelism for the KAN-FET path: one process per GPU, full parameter
replica per rank, contiguous row slices, ONE flat fp32 all-reduce of the parameter gradients per optimiser step
(NCCL over NVLink on the GPU box; gloo in the CPU tests).  The reference is single-process; the three semantics
that couple rows are kept explicit so that N ranks reproduce the 1-rank result:

  1. hard-gated LogisticBasis uses the GLOBAL last row as prev_x (train_ecg_kan_fet_nn_ode.py:131-132)
     -> hardgate_prev_sync broadcasts the [in_dim] row from the last rank after every call;
  2. torchdiffeq dopri5 uses one RMS error norm over the whole (B, D) tensor -> norm_sync all-reduces
     (sum of squares, element count) per attempted step (pass options={"norm_sync": norm_sync});
  3. clip_grad_norm_ / loss means are taken after the all-reduce, over the global batch.

Nothing here computes a layer value; it is plumbing around torch.distributed.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_bounds(n_rows: int, rank: Optional[int] = None, world_size: Optional[int] = None) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of n_rows owned by `rank`; the first n_rows % world ranks get one extra row."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    base, extra = divmod(n_rows, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_rows(x: torch.Tensor) -> torch.Tensor:
    lo, hi = shard_bounds(x.shape[0])
    return x[lo:hi]


def _grad_params(params: Iterable[torch.nn.Parameter]) -> List[torch.nn.Parameter]:
    return [p for p in params if p.requires_grad]


def allreduce_gradients(model_or_params, average: bool = True, local_rows: Optional[int] = None,
                        global_rows: Optional[int] = None) -> Optional[torch.Tensor]:
    """Sum every parameter gradient across ranks with a single collective on one flat fp32 buffer and write the
    result back in place.  Parameters that received no gradient on this rank contribute zeros.

    average=True divides by the world size (loss = mean over the LOCAL shard, equal shards).  With unequal shards
    pass local_rows / global_rows: gradients are weighted local_rows / global_rows before the sum, which equals
    the gradient of the mean over the global batch."""
    if getattr(model_or_params, "built", False) and hasattr(model_or_params, "flat_grad"):
        # fetode.optim.FlatAdam after its first step: the gradients already live in one flat buffer
        opt = model_or_params
        opt._sync_grads()
        flat = opt.flat_grad
        rank, w = world()
        if local_rows is not None and global_rows is not None:
            flat.mul_(float(local_rows) / float(global_rows))
        if w > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
            if average and local_rows is None:
                flat.div_(w)
        return flat
    if hasattr(model_or_params, "all_params"):
        model_or_params = model_or_params.all_params
    params = _grad_params(model_or_params.parameters() if hasattr(model_or_params, "parameters") else model_or_params)
    if not params:
        return None
    rank, w = world()
    dev = params[0].device
    flat = torch.zeros(sum(p.numel() for p in params), dtype=torch.float32, device=dev)
    off = 0
    for p in params:
        n = p.numel()
        if p.grad is not None:
            flat[off:off + n].copy_(p.grad.reshape(-1))
        off += n
    if local_rows is not None and global_rows is not None:
        flat.mul_(float(local_rows) / float(global_rows))
    if w > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        if average and local_rows is None:
            flat.div_(w)
    off = 0
    for p in params:
        n = p.numel()
        g = flat[off:off + n].view_as(p)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += n
    return flat


def norm_sync(pair: torch.Tensor) -> None:
    """dopri5 `options["norm_sync"]`: pair = [sum of squared scaled errors, element count] (float64), summed in place."""
    if world()[1] > 1:
        dist.all_reduce(pair, op=dist.ReduceOp.SUM)


def hardgate_prev_sync(module) -> None:
    """Make every rank's hard-gate state the last row of the GLOBAL batch (= the last rank's last row), once, for
    every hard-gated basis under ``module``.  (The per-call form is ``enable_hardgate_sync``.)"""
    rank, w = world()
    if w == 1:
        return
    for m in module.modules():
        row = getattr(m, "_row", None)
        if row is not None and hasattr(m, "branch_breaking_point"):
            dist.broadcast(row, src=w - 1)


def enable_hardgate_sync(module, on: bool = True) -> int:
    """Option 2 of SURVEY section 8e: after EVERY call of a hard-gated ``ecg.LogisticBasis`` under ``module`` the new
    ``prev_x`` row ([in_dim] floats) is broadcast from the last rank, so each rank gates its next call on the last
    row of the GLOBAL batch exactly like the single-process reference (train_ecg_kan_fet_nn_ode.py:129-132).  The
    broadcast is issued on the current stream (capturable).  Returns the number of modules switched."""
    n = 0
    for m in module.modules():
        if hasattr(m, "branch_breaking_point") and hasattr(m, "_dist_sync"):
            m._dist_sync = bool(on)
            n += 1
    return n


def shard_rows_with_sentinel(x: torch.Tensor, rank: Optional[int] = None,
                             world_size: Optional[int] = None) -> Tuple[torch.Tensor, int]:
    """Option 1 of SURVEY section 8e (zero communication): this rank's contiguous rows of the global batch ``x``
    followed -- on every rank but the last -- by the GLOBAL last row.  Rows never interact except through the
    hard-gated basis, whose state is the last row of whatever batch it sees: with the sentinel appended that is the
    global last row on every rank, so every rank computes the reference's function on its rows.  Returns
    (rows [n_real (+1), ...], n_real); the caller drops output row n_real before the loss."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    lo, hi = shard_bounds(x.shape[0], rank, world_size)
    rows = x[lo:hi]
    if world_size > 1 and hi != x.shape[0]:
        rows = torch.cat([rows, x[-1:]], dim=0)
    return rows, hi - lo


def clip_grad_norm_(params, max_norm: float) -> torch.Tensor:
    """clip_grad_norm_ on already all-reduced gradients (train_ecg_kan_fet_nn_ode.py:1110): same on every rank."""
    return torch.nn.utils.clip_grad_norm_(_grad_params(params), max_norm)
