"""Drop-in for the model classes of the reference's ``mnist_kuramoto_kan.py`` (commit 704a0df4):
``LogisticBasis``, ``KANLinear`` (the copy with ``use_logistic_basis`` / ``logistic_bias`` and no
logistic scaler, :25-142), ``Kuramoto2D`` (:145-204) and ``KuramotoKANClassifier`` (:207-225).

Same constructor arguments, parameter / buffer names and forward semantics; the arithmetic runs in
libfetode.so: csrc/kuramoto.cu (the whole Euler loop of the phase lattice in one launch, forward and
reverse sweep) and csrc/kanlinear*.cu (the 1568 -> 10 head).  No CPU path.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from . import ops
from .efficientkan import LogisticBasis, _design_matrix  # noqa: F401  (same class as efficientkan.py:7-24)


class KANLinear(nn.Module):
    """mnist_kuramoto_kan.py:25-142."""

    def __init__(self, in_features, out_features, grid_size=5, spline_order=3, scale_noise=0.1, scale_base=1.0,
                 scale_spline=1.0, enable_standalone_scale_spline=True, base_activation=nn.SiLU, grid_eps=0.02,
                 grid_range=(-1.0, 1.0), use_logistic_basis=True, num_basis=10, scale_logistic=1.0):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.grid_size, self.spline_order = grid_size, spline_order
        step = (grid_range[1] - grid_range[0]) / grid_size
        knots = torch.arange(-spline_order, grid_size + spline_order + 1) * step + grid_range[0]
        self.register_buffer("grid", knots.expand(in_features, -1).contiguous())
        self.base_weight = nn.Parameter(torch.empty(out_features, in_features))
        self.spline_weight = nn.Parameter(torch.empty(out_features, in_features, grid_size + spline_order))
        self.enable_standalone_scale_spline = enable_standalone_scale_spline
        if enable_standalone_scale_spline:
            self.spline_scaler = nn.Parameter(torch.empty(out_features, in_features))
        self.scale_noise, self.scale_base, self.scale_spline = scale_noise, scale_base, scale_spline
        if base_activation is not nn.SiLU:
            raise NotImplementedError("fetode KANLinear: SiLU base branch only")
        self.base_activation = base_activation()
        self.grid_eps = grid_eps
        self.use_logistic_basis = use_logistic_basis
        self.num_basis, self.scale_logistic = num_basis, scale_logistic
        if use_logistic_basis:
            self.logistic_basis = LogisticBasis(in_features, num_basis)
            self.logistic_weight = nn.Parameter(torch.empty(out_features, in_features * num_basis))
            self.logistic_bias = nn.Parameter(torch.zeros(out_features))
        self._pack = ops.LayerPack(self)
        self.reset_parameters()

    def reset_parameters(self):                                      # :82-99
        nn.init.kaiming_uniform_(self.base_weight, a=math.sqrt(5) * self.scale_base)
        with torch.no_grad():
            jitter = (torch.rand(self.grid_size + 1, self.in_features, self.out_features) - 0.5) \
                * self.scale_noise / self.grid_size
            gain = 1.0 if self.enable_standalone_scale_spline else self.scale_spline
            pts = self.grid.T[self.spline_order:-self.spline_order]
            self.spline_weight.data.copy_(gain * self.curve2coeff(pts, jitter))
            if self.enable_standalone_scale_spline:
                nn.init.kaiming_uniform_(self.spline_scaler, a=math.sqrt(5) * self.scale_spline)
        if self.use_logistic_basis:
            nn.init.kaiming_uniform_(self.logistic_weight, a=math.sqrt(5) * self.scale_logistic)
            nn.init.zeros_(self.logistic_bias)

    def b_splines(self, x):                                          # :101-110
        assert x.dim() == 2 and x.size(1) == self.in_features
        if x.is_cuda:
            return ops.bspline_basis(self, ops.f32c(x))
        return _design_matrix(x, self.grid, self.spline_order).contiguous()

    def curve2coeff(self, x, y):                                     # :112-118
        assert x.dim() == 2 and x.size(1) == self.in_features
        assert y.size() == (x.size(0), self.in_features, self.out_features)
        design = self.b_splines(x).transpose(0, 1)
        coeff = torch.linalg.lstsq(design, y.transpose(0, 1)).solution
        return coeff.permute(2, 0, 1).contiguous()

    @property
    def scaled_spline_weight(self):                                  # :120-124
        if self.enable_standalone_scale_spline:
            return self.spline_weight * self.spline_scaler.unsqueeze(-1)
        return self.spline_weight

    def forward(self, x):                                            # :126-142
        assert x.size(-1) == self.in_features
        lead = x.shape[:-1]
        x2 = ops.f32c(x.reshape(-1, self.in_features))
        y = ops.kanlinear_apply(self._pack, x2, None)
        return y.reshape(*lead, self.out_features)


class Kuramoto2D(nn.Module):
    """mnist_kuramoto_kan.py:145-204: theta <- theta + dt (omega + K sum_4nbr sin(theta_n - theta)), ``steps``
    times, features [cos theta, sin theta] of shape (B, 2 H W)."""

    def __init__(self, H=28, W=28, steps=10, dt=0.15, learn_K=True, learn_omega=True):
        super().__init__()
        self.H, self.W, self.steps, self.dt = H, W, steps, dt
        if learn_K:
            self.K = nn.Parameter(torch.tensor(0.5))
        else:
            self.register_buffer("K", torch.tensor(0.5), persistent=False)
        if learn_omega:
            self.omega = nn.Parameter(torch.zeros(1, 1, H, W))
        else:
            self.register_buffer("omega", torch.zeros(1, 1, H, W))
        k = torch.zeros(1, 1, 3, 3)
        k[0, 0, 0, 1] = k[0, 0, 2, 1] = k[0, 0, 1, 0] = k[0, 0, 1, 2] = 1.0
        self.register_buffer("neighbor_kernel", k)       # kept for state_dict parity; the stencil is in the kernel

    def forward(self, x_img):
        B, C, H, W = x_img.shape
        assert C == 1 and H == self.H and W == self.W
        return ops.kuramoto_apply(x_img, self.K, self.omega, H, W, self.steps, self.dt)


class KuramotoKANClassifier(nn.Module):
    """mnist_kuramoto_kan.py:207-225."""

    def __init__(self, H=28, W=28, num_classes=10, kuramoto_steps=10, num_basis=8):
        super().__init__()
        self.osc = Kuramoto2D(H=H, W=W, steps=kuramoto_steps, dt=0.15, learn_K=True, learn_omega=True)
        self.head = KANLinear(2 * H * W, num_classes, grid_size=5, spline_order=3, base_activation=nn.SiLU,
                              use_logistic_basis=True, num_basis=num_basis)

    def forward(self, x_img):
        return self.head(self.osc(x_img))
