"""Drop-in for the reference's ``efficient_kan/efficientkan.py``: LogisticBasis, KANLinear, KAN,
plus the ``KANFET`` class the reference's scripts call but the snapshot does not ship
(train_kanfet_node_predprey.py:146, kan_diffusion_ett.py:421-425).

Constructor signatures, attribute / parameter / buffer names and forward semantics follow
efficientkan.py:7-285 (reference checkout, commit 704a0df4) so ``state_dict()`` round-trips and
the training scripts run after an import swap.  ``forward`` never evaluates the layer with torch
ops: it launches the sm_100a kernels in libfetode.so (csrc/kanlinear*.cu, csrc/smallnet*.cu) and
raises if the tensors are not on a CUDA device.

Kept in PyTorch on purpose ("init/update_grid/reg = keep in PyTorch"): parameter
initialisation (least-squares fit of the spline coefficients to noise), ``update_grid`` and
``regularization_loss`` -- none of them is on the right-hand-side path.

KANFET definition (the reference has none -- "parity unpinned"): a KAN in which
every layer's ``logistic_basis`` slot holds a ``TwoDimensionFerroelectricBasis(in, num_basis)``
(ferro_class.py:526-596), whose [B,in,num_basis] output contract equals LogisticBasis's.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from . import ops
from ._lib import BRANCH_FERRO2D
from .ferro_class import TwoDimensionFerroelectricBasis


class LogisticBasis(nn.Module):
    """phi[b,i,j] = 2 / (1 + exp(-a_ij (x_bi - b_ij)))   (efficientkan.py:7-24)."""

    def __init__(self, in_dim: int, num_basis: int):
        super().__init__()
        self.in_dim, self.num_basis = in_dim, num_basis
        self.a = nn.Parameter(torch.randn(in_dim, num_basis))
        self.b = nn.Parameter(torch.randn(in_dim, num_basis))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.dim() == 2 and x.size(1) == self.in_dim
        ops.abort_if_tracing()
        return ops.logistic_basis_apply(self, ops.f32c(x))


def _design_matrix(x: torch.Tensor, knots: torch.Tensor, order: int) -> torch.Tensor:
    """B-spline collocation matrix [B, I, G+order] for *initialisation only* (torch ops, any
    device).  Triangular-table form of the recurrence: column block ``lvl`` is built from the
    previous one with the two knot-difference ratios per entry."""
    t = knots.unsqueeze(0)                                  # [1, I, NK]
    xs = x.unsqueeze(-1)                                    # [B, I, 1]
    table = ((xs >= t[..., :-1]) & (xs < t[..., 1:])).to(x.dtype)
    for lvl in range(1, order + 1):
        n = table.shape[-1] - 1
        w_left = (xs - t[..., :n]) / (t[..., lvl:lvl + n] - t[..., :n])
        w_right = (t[..., lvl + 1:lvl + 1 + n] - xs) / (t[..., lvl + 1:lvl + 1 + n] - t[..., 1:1 + n])
        table = w_left * table[..., :n] + w_right * table[..., 1:]
    return table


class KANLinear(nn.Module):
    """efficientkan.py:27-237."""

    def __init__(self, in_features, out_features, grid_size=5, spline_order=3, scale_noise=0.1,
                 scale_base=1.0, scale_spline=1.0, enable_standalone_scale_spline=True,
                 base_activation=nn.SiLU, grid_eps=0.02, grid_range=[-1, 1],
                 enable_logistic_basis=True, num_basis=10, scale_logistic=1.0,
                 enable_standalone_scale_logistic=True):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.grid_size, self.spline_order = grid_size, spline_order
        step = (grid_range[1] - grid_range[0]) / grid_size
        knots = torch.arange(-spline_order, grid_size + spline_order + 1) * step + grid_range[0]
        self.register_buffer("grid", knots.expand(in_features, -1).contiguous())
        n_coef = grid_size + spline_order
        self.base_weight = nn.Parameter(torch.empty(out_features, in_features))
        self.spline_weight = nn.Parameter(torch.empty(out_features, in_features, n_coef))
        if enable_standalone_scale_spline:
            self.spline_scaler = nn.Parameter(torch.empty(out_features, in_features))
        self.scale_noise, self.scale_base, self.scale_spline = scale_noise, scale_base, scale_spline
        self.enable_standalone_scale_spline = enable_standalone_scale_spline
        if base_activation is not nn.SiLU:
            raise NotImplementedError("fetode KANLinear: the sm_100a kernels implement the SiLU base branch only "
                                      "(every reference call site uses the default)")
        self.base_activation = base_activation()
        self.grid_eps = grid_eps
        self.enable_logistic_basis = enable_logistic_basis
        self.num_basis, self.scale_logistic = num_basis, scale_logistic
        self.enable_standalone_scale_logistic = enable_standalone_scale_logistic
        if enable_logistic_basis:
            self.logistic_basis = LogisticBasis(in_features, num_basis)
            self.logistic_weight = nn.Parameter(torch.empty(out_features, in_features * num_basis))
            if enable_standalone_scale_logistic:
                self.logistic_scaler = nn.Parameter(torch.empty(out_features))
        self._pack = ops.LayerPack(self)
        self.reset_parameters()

    # ---- initialisation (PyTorch; efficientkan.py:92-115) ------------------------------------
    def reset_parameters(self):
        nn.init.kaiming_uniform_(self.base_weight, a=math.sqrt(5) * self.scale_base)
        with torch.no_grad():
            jitter = (torch.rand(self.grid_size + 1, self.in_features, self.out_features) - 0.5) \
                * self.scale_noise / self.grid_size
            pts = self.grid.T[self.spline_order:-self.spline_order]
            gain = 1.0 if self.enable_standalone_scale_spline else self.scale_spline
            self.spline_weight.data.copy_(gain * self.curve2coeff(pts, jitter))
            if self.enable_standalone_scale_spline:
                nn.init.kaiming_uniform_(self.spline_scaler, a=math.sqrt(5) * self.scale_spline)
        if self.enable_logistic_basis:
            nn.init.kaiming_uniform_(self.logistic_weight, a=math.sqrt(5) * self.scale_logistic)
            if self.enable_standalone_scale_logistic:
                nn.init.ones_(self.logistic_scaler)

    def b_splines(self, x: torch.Tensor) -> torch.Tensor:
        """[B, in] -> [B, in, grid_size + spline_order]  (efficientkan.py:117-131).  CUDA inputs
        go through the kernel; host tensors (constructor-time fitting) through the torch helper."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        if x.is_cuda:
            return ops.bspline_basis(self, ops.f32c(x))
        return _design_matrix(x, self.grid, self.spline_order).contiguous()

    def curve2coeff(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Least-squares spline coefficients interpolating y at x (efficientkan.py:133-143)."""
        assert x.dim() == 2 and x.size(1) == self.in_features
        assert y.size() == (x.size(0), self.in_features, self.out_features)
        design = self.b_splines(x).transpose(0, 1)                          # [I, B, S]
        coeff = torch.linalg.lstsq(design, y.transpose(0, 1)).solution      # [I, S, O]
        out = coeff.permute(2, 0, 1)
        assert out.size() == (self.out_features, self.in_features, self.grid_size + self.spline_order)
        return out.contiguous()

    @property
    def scaled_spline_weight(self):
        if self.enable_standalone_scale_spline:
            return self.spline_weight * self.spline_scaler.unsqueeze(-1)
        return self.spline_weight * 1.0

    @property
    def scaled_logistic_weight(self):
        if not self.enable_logistic_basis:
            return None
        w = self.logistic_weight * self.scale_logistic
        if self.enable_standalone_scale_logistic:
            w = w * self.logistic_scaler.unsqueeze(-1)
        return w

    # ---- the hot path ------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.size(-1) == self.in_features
        ops.abort_if_tracing()
        lead = x.shape[:-1]
        x2d = ops.f32c(x.reshape(-1, self.in_features))
        prev = None
        ferro = self.enable_logistic_basis and isinstance(self.logistic_basis, TwoDimensionFerroelectricBasis)
        if ferro:
            prev = self.logistic_basis._prev_rows(x2d.shape[0], x2d.device)
        y = ops.kanlinear_apply(self._pack, x2d, prev)
        if ferro:
            self.logistic_basis._set_rows(x2d.detach().clone())   # own the state (reference: detach().clone())
        return y.reshape(*lead, self.out_features)

    # ---- grid adaptation / regulariser (PyTorch; efficientkan.py:184-237) ----------------------
    @torch.no_grad()
    def update_grid(self, x: torch.Tensor, margin=0.01):
        assert x.dim() == 2 and x.size(1) == self.in_features
        n = x.size(0)
        basis = self.b_splines(x).permute(1, 0, 2)                          # [I, n, S]
        per_input = torch.bmm(basis, self.scaled_spline_weight.permute(1, 2, 0)).permute(1, 0, 2)   # [n, I, O]
        xs = torch.sort(x, dim=0)[0]
        picks = torch.linspace(0, n - 1, self.grid_size + 1, dtype=torch.int64, device=x.device)
        adaptive = xs[picks]
        step = (xs[-1] - xs[0] + 2 * margin) / self.grid_size
        uniform = torch.arange(self.grid_size + 1, dtype=torch.float32, device=x.device).unsqueeze(1) * step \
            + xs[0] - margin
        inner = self.grid_eps * uniform + (1 - self.grid_eps) * adaptive
        k = self.spline_order
        below = inner[:1] - step * torch.arange(k, 0, -1, device=x.device).unsqueeze(1)
        above = inner[-1:] + step * torch.arange(1, k + 1, device=x.device).unsqueeze(1)
        self.grid.copy_(torch.cat([below, inner, above], dim=0).T)
        self.spline_weight.data.copy_(self.curve2coeff(x, per_input))

    def regularization_loss(self, regularize_activation=1.0, regularize_entropy=1.0, regularize_logistic_l1=0.0):
        mag = self.spline_weight.abs().mean(-1)
        total = mag.sum()
        share = mag / (total + 1e-12)
        entropy = -torch.sum(share * (share + 1e-12).log())
        reg = regularize_activation * total + regularize_entropy * entropy
        if self.enable_logistic_basis and regularize_logistic_l1 != 0.0:
            reg = reg + regularize_logistic_l1 * self.logistic_weight.abs().mean()
        return reg


class KAN(nn.Module):
    """efficientkan.py:240-285.  Two-layer stacks whose widths fit the fused kernel
    (e.g. the Lotka-Volterra [2,10,2]) evaluate all layers in ONE launch."""

    def __init__(self, layers_hidden, grid_size=5, spline_order=3, scale_noise=0.1, scale_base=1.0,
                 scale_spline=1.0, base_activation=nn.SiLU, grid_eps=0.02, grid_range=[-1, 1]):
        super().__init__()
        self.grid_size, self.spline_order = grid_size, spline_order
        self.layers = nn.ModuleList(
            KANLinear(i, o, grid_size=grid_size, spline_order=spline_order, scale_noise=scale_noise,
                      scale_base=scale_base, scale_spline=scale_spline, base_activation=base_activation,
                      grid_eps=grid_eps, grid_range=grid_range)
            for i, o in zip(layers_hidden, layers_hidden[1:]))
        self._plan: Optional[ops.SmallNetPlan] = None
        self._plan_checked = False
        self._fused_state: Optional[torch.Tensor] = None

    # ---- fused small-width path --------------------------------------------------------------
    def small_plan(self) -> Optional[ops.SmallNetPlan]:
        """SmallNetPlan if the fused kernel has an instantiation for this stack, else None."""
        if not self._plan_checked:
            self._plan_checked = True
            if ops.SmallNetPlan.supported(list(self.layers)):
                self._plan = ops.SmallNetPlan(list(self.layers))
        return self._plan

    def _is_ferro(self) -> bool:
        return isinstance(getattr(self.layers[0], "logistic_basis", None), TwoDimensionFerroelectricBasis)

    def hysteresis_state(self, B: int, device) -> Optional[torch.Tensor]:
        """[B, sum(in_features)] previous inputs of every layer for the fused kernels, or None when
        every layer is in its reset state (previous input == 0)."""
        if not self._is_ferro():
            return None
        rows = [l.logistic_basis._rows for l in self.layers]
        if all(r is None for r in rows):
            return None
        st = self._fused_state
        if st is not None and st.shape[0] == B:
            off, same = 0, True
            for l, r in zip(self.layers, rows):
                w = l.in_features
                same &= r is not None and r.data_ptr() == st[:, off:off + w].data_ptr() and r.shape[0] == B
                off += w
            if same:
                return st
        cols = [l.logistic_basis._prev_rows(B, device) if r is not None
                else torch.zeros(B, l.in_features, device=device) for l, r in zip(self.layers, rows)]
        st = torch.cat(cols, dim=1).contiguous()
        self.adopt_state(st)
        return st

    def adopt_state(self, st: torch.Tensor) -> None:
        """Make the per-layer ``prev_x`` buffers views of the fused state tensor."""
        self._fused_state = st
        off = 0
        for l in self.layers:
            w = l.in_features
            l.logistic_basis._set_rows(st[:, off:off + w])
            off += w

    def forward(self, x: torch.Tensor, update_grid=False):
        plan = None if update_grid else self.small_plan()
        tr = ops.active_trace()
        if tr is not None:
            if plan is None or x.dim() != 2 or x.shape[1] != plan.D or plan.net.widths[2] != plan.D:
                raise ops.TraceAbort()
            sentinel = torch.empty(0, device=x.device)
            tr.calls.append((self, x, sentinel))
            return sentinel
        if plan is not None and x.dim() == 2 and x.shape[1] == plan.D and plan.net.widths[2] == plan.D \
                and not (torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters()))):
            x2 = ops.f32c(x)
            if self._is_ferro():
                st = self.hysteresis_state(x2.shape[0], x2.device)
                fresh = st is None
                if fresh:
                    st = torch.empty(x2.shape[0], plan.state_floats, dtype=torch.float32, device=x2.device)
                y = plan.rhs(x2, st, fresh)
                self.adopt_state(st)
                return y
            return plan.rhs(x2, None, True)
        for layer in self.layers:
            if update_grid:
                layer.update_grid(x)
            x = layer(x)
        return x

    def regularization_loss(self, regularize_activation=1.0, regularize_entropy=1.0):
        return sum(l.regularization_loss(regularize_activation, regularize_entropy) for l in self.layers)


class KANFET(KAN):
    """KAN with the ferroelectric hysteresis neuron in every layer's branch slot (see module
    docstring).  Call sites: train_kanfet_node_predprey.py:146 ``KANFET(layers_hidden=[2,10,2],
    grid_size=5)``, kan_diffusion_ett.py:421-425 ``KANFET([in,256])``."""

    def __init__(self, layers_hidden, grid_size=5, spline_order=3, **kw):
        super().__init__(layers_hidden, grid_size=grid_size, spline_order=spline_order, **kw)
        for layer in self.layers:
            layer.logistic_basis = TwoDimensionFerroelectricBasis(layer.in_features, layer.num_basis)

    def reset_state(self):
        self._fused_state = None
        for layer in self.layers:
            layer.logistic_basis.reset_state()
