"""Modules of the ETT / Time-MMD "KAN-FET" forecaster that sit on the hot path, same names / constructor arguments /
parameter names as the reference (commit 704a0df4):

  LogisticBasis, LogisticBasisLinear       train_kan_fet_ett.py:741-776
  FullyNonlinearKANCell                    train_kan_fet_ett.py:780-796   (same cell: train_ecg.py:203-217)
  KANRNNEncoder                            train_kan_fet_ett.py:799-818

The 336 sequential cell calls of ``KANRNNEncoder.forward`` (~10 launches each in the reference) are ONE kernel launch
(csrc/kanrnn.cu), forward and backward.  The wide encoder ``KAN([context_len * F, 128]) -> ReLU -> KAN([128, latent])``
(:676-681) is fetode.KAN (tcgen05 path) and needs no module of its own.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import ops
from ._lib import check, f32c, lib, ptr, stream
from .efficientkan import LogisticBasis  # noqa: F401  (same parameters a, b [in_dim, num_basis]; forward -> [B, in, nb])


class _KanRnnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, h0, a_in, b_in, a_h, b_h, H):
        ops._require_cuda(x, "input")
        B, T, F = x.shape
        nb = a_in.shape[1]
        xs = f32c(x)
        ps = [f32c(p.detach()) for p in (a_in, b_in, a_h, b_h)]
        h0c = None if h0 is None else f32c(h0.detach())
        out = torch.empty(B, H, dtype=torch.float32, device=x.device)
        need = any(ctx.needs_input_grad)
        hist = torch.empty(T + 1, B, H, dtype=torch.float32, device=x.device) if (need and H > F * nb) else None
        check(lib().fetode_kanrnn_fwd(ptr(xs), ptr(h0c), *[ptr(p) for p in ps], ptr(out), ptr(hist), C.c_int64(B),
                                      C.c_int32(T), C.c_int32(F), C.c_int32(H), C.c_int32(nb), stream()), "kanrnn_fwd")
        ctx.save_for_backward(xs, *ps, hist if hist is not None else out.new_empty(0))
        ctx.dims = (B, T, F, H, nb, hist is not None, h0 is not None)
        return out

    @staticmethod
    def backward(ctx, g):
        xs, a_in, b_in, a_h, b_h, hist = ctx.saved_tensors
        B, T, F, H, nb, has_hist, has_h0 = ctx.dims
        g = f32c(g)
        gx = torch.zeros_like(xs) if ctx.needs_input_grad[0] else None
        gh0 = torch.empty(B, H, dtype=torch.float32, device=xs.device) if (has_h0 and ctx.needs_input_grad[1]) else None
        gp = [torch.zeros_like(p) if ctx.needs_input_grad[2 + k] else None for k, p in enumerate((a_in, b_in, a_h, b_h))]
        check(lib().fetode_kanrnn_bwd(ptr(xs), ptr(a_in), ptr(b_in), ptr(a_h), ptr(b_h), ptr(hist if has_hist else None),
                                      ptr(g), ptr(gx), ptr(gh0), *[ptr(t) for t in gp], C.c_int64(B), C.c_int32(T),
                                      C.c_int32(F), C.c_int32(H), C.c_int32(nb), stream()), "kanrnn_bwd")
        return (gx, gh0, *gp, None)


class LogisticBasisLinear(nn.Module):
    """train_kan_fet_ett.py:753-776: phi = LogisticBasis(x) flattened, y = phi @ weight (+ bias)."""

    def __init__(self, in_dim: int, out_dim: int, num_basis: int, bias: bool = True):
        super().__init__()
        self.in_dim, self.out_dim, self.num_basis = in_dim, out_dim, num_basis
        self.basis = LogisticBasis(in_dim, num_basis)
        self.weight = nn.Parameter(torch.randn(in_dim * num_basis, out_dim) * 0.02)
        self.bias = nn.Parameter(torch.zeros(out_dim)) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        phi = self.basis(x).reshape(x.shape[0], -1)
        y = ops.LinearFn.apply(ops.f32c(phi), self.weight.t(), self.bias) if phi.is_cuda else phi @ self.weight
        return y


class FullyNonlinearKANCell(nn.Module):
    """train_kan_fet_ett.py:780-796: out = sigmoid([basis_in(x_t) | basis_h(h_prev)])[:, :hidden_size]."""

    def __init__(self, input_size, hidden_size, num_basis):
        super().__init__()
        self.input_basis = LogisticBasis(input_size, num_basis)
        self.hidden_basis = LogisticBasis(hidden_size, num_basis)
        self.activation = nn.Sigmoid()
        self.hidden_size, self.num_basis = hidden_size, num_basis

    def run(self, x_seq: torch.Tensor, h0=None) -> torch.Tensor:
        """h_T after feeding x_seq [B, T, input_size] through the cell T times, one launch."""
        ops.abort_if_tracing()
        return _KanRnnFn.apply(x_seq, h0, self.input_basis.a, self.input_basis.b, self.hidden_basis.a,
                               self.hidden_basis.b, self.hidden_size)

    def forward(self, x_t, h_prev):
        return self.run(x_t.unsqueeze(1), h_prev)


class KANRNNEncoder(nn.Module):
    """train_kan_fet_ett.py:799-818."""

    def __init__(self, num_features: int, hidden_size: int, latent_dim: int, num_basis: int):
        super().__init__()
        self.hidden_size = hidden_size
        self.rnn_cell = FullyNonlinearKANCell(num_features, hidden_size, num_basis)
        self.to_latent = nn.Linear(hidden_size, latent_dim)

    def forward(self, x_ctx: torch.Tensor) -> torch.Tensor:
        h = self.rnn_cell.run(x_ctx)                      # zeros initial state (:811), all T cell steps in one launch
        return ops.linear(h, self.to_latent)
