"""Right-hand-side modules of the ECG scripts, same names / constructor arguments / parameter names as the
reference (commit 704a0df4):

  LogisticBasis (hysteretic, hard-gated)   train_ecg_kan_fet_nn_ode.py:54-134   (same math smooth_test_KAN_ferro.py:58-80)
  KANFeatureMixer                          train_ecg_kan_fet_nn_ode.py:408-420
  MLPKANODEFunc / No_MLP_KANODEFunc        train_ecg_kan_fet_nn_ode.py:423-508
  KANFetODEFunc                            train_ecg.py:986-1013

The neuron / KAN arithmetic runs in libfetode.so (csrc/ferro.cu, csrc/kanlinear*.cu).  The hard gate
``sigmoid(5 (x - prev_x)) > 0.5`` is evaluated in the kernel with IEEE exp / add / divide so the branch
taken is bit-identical to torch's CPU result; ``branch_state`` is the uint8 gate the kernel reports.
The plain ``kan.KAN`` of the missing ``kan_diffusion`` submodule is substituted by fetode.efficientkan.KAN
("parity unpinned" for that submodule).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .efficientkan import KAN
from .ferro_class import FerroelectricBasis


class LogisticBasis(nn.Module):
    """Hysteretic two-branch logistic basis, forward(x:(B,in)) -> (B,in,num_basis)."""

    def __init__(self, in_dim: int, num_basis: int, gate_slope: float = 5.0, init_prev: float = 0.0,
                 eps: float = 1e-6, branch_breaking_point=0.5, use_noise=False, noise_std=0.05):
        super().__init__()
        self.in_dim, self.num_basis = in_dim, num_basis
        self.gate_slope, self.eps = gate_slope, eps
        self.use_noise, self.noise_std = use_noise, noise_std
        self.branch_breaking_point = branch_breaking_point
        self.k = nn.Parameter(torch.rand(in_dim, num_basis) * 2 + 0.5)
        self.Ec = nn.Parameter(torch.rand(in_dim, num_basis) * 2 + 0.5)
        self.Ps = nn.Parameter(torch.rand(in_dim, num_basis) * 1.5 + 0.5)
        self.bias = nn.Parameter(torch.randn(in_dim, num_basis) * 0.1)
        self.coef = nn.Parameter(torch.randn(in_dim, num_basis))          # unused by forward, as in the reference
        self.register_buffer("prev_x", torch.zeros(1, in_dim, num_basis))
        self.register_buffer("branch_state", torch.ones(1, in_dim, num_basis))
        self._row = None                                                   # compact [in_dim] previous last row
        self._dist_sync = False                                            # fetode.dist.enable_hardgate_sync

    def reset_state(self, value: float = 0.0):
        self._row = None
        self.prev_x = torch.zeros(1, self.in_dim, self.num_basis, device=self.prev_x.device)
        self.branch_state = torch.ones(1, self.in_dim, self.num_basis, device=self.prev_x.device)

    def _load_from_state_dict(self, state_dict, prefix, *args):
        """prev_x is a stride-0 view of the compact row after any call: receive the checkpoint's buffers in real tensors
        of the incoming shape (train_kan_fet_ett.py:353-358 restores the best state_dict), then re-read the row lazily."""
        for name in ("prev_x", "branch_state"):
            t = state_dict.get(prefix + name)
            cur = self._buffers.get(name)
            if t is not None and cur is not None:
                self._buffers[name] = torch.empty(tuple(t.shape), dtype=t.dtype if name == "branch_state" else cur.dtype,
                                                  device=cur.device)
        super()._load_from_state_dict(state_dict, prefix, *args)
        self._row = None                                                   # _prev_row() picks the loaded buffer up

    def _prev_row(self, device):
        if self._row is None:
            # honour a buffer written from outside (load_state_dict): first basis column carries the row
            if bool((self.prev_x != 0).any()):
                self._row = self.prev_x[0, :, 0].to(device).contiguous()
        return self._row

    def _call(self, x, apply_sigmoid):
        """apply_sigmoid: False/0 = basis only, True/1 = sigmoid fused in (KANFeatureMixer), 2 = sigmoid followed by
        nan_to_num(nan=0, posinf=1e3, neginf=-1e3) fused in (MLPKANODEFunc.forward :464-465)."""
        if x.dim() != 2 or x.size(1) != self.in_dim:
            raise ValueError(f"x must be (B,{self.in_dim}), got {tuple(x.shape)}")
        if self.branch_breaking_point != 0.5:
            raise NotImplementedError("fetode: the bit-exact gate kernel implements branch_breaking_point=0.5")
        ops.abort_if_tracing()
        x2 = ops.f32c(x)
        prev = self._prev_row(x2.device)
        new_row = torch.empty(self.in_dim, dtype=torch.float32, device=x2.device)
        noisy = self.use_noise and self.noise_std != 0.0
        out, side = ops.ferro_apply(self, ops.FERRO_HARDGATE, x2, prev, want_gate=True,
                                    apply_sigmoid=0 if noisy else int(apply_sigmoid), prev_out=new_row,
                                    gate_slope=self.gate_slope, alpha=0.5)
        if noisy:
            out = out + torch.randn_like(out) * self.noise_std
            if apply_sigmoid:
                out = torch.sigmoid(out)
            if int(apply_sigmoid) == 2:
                out = torch.nan_to_num(out, nan=0.0, posinf=1e3, neginf=-1e3)
        if self._dist_sync:
            # batch-sharded run: the reference's prev_x is the last row of the GLOBAL batch (:131-132) = the last
            # rank's last row
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                dist.broadcast(new_row, src=dist.get_world_size() - 1)
        self._row = new_row
        self.prev_x = new_row.view(1, -1, 1).expand(1, self.in_dim, self.num_basis)
        self.branch_state = side["gate"]                                   # uint8 [B,in,K]: 1 = up branch
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._call(x, False)


class KANFeatureMixer(nn.Module):
    """x:(B,D) -> sigmoid(basis(x)) flattened to (B, D*num_basis); the sigmoid is fused into the kernel."""

    def __init__(self, dim, num_basis, act=None):
        super().__init__()
        self.basis = LogisticBasis(dim, num_basis)
        self.act = nn.Sigmoid() if act is None else act

    def forward(self, x, nan_to_num: bool = False):
        """nan_to_num=True additionally applies torch.nan_to_num(phi, nan=0, posinf=1e3, neginf=-1e3) -- in the same
        kernel when the activation is the default Sigmoid."""
        if isinstance(self.act, nn.Sigmoid):
            phi = self.basis._call(x, 2 if nan_to_num else 1)
        else:
            phi = self.act(self.basis(x))
            if nan_to_num:
                phi = torch.nan_to_num(phi, nan=0.0, posinf=1e3, neginf=-1e3)
        return phi.reshape(x.size(0), -1)


class MLPKANODEFunc(nn.Module):
    """train_ecg_kan_fet_nn_ode.py:423-481."""

    def __init__(self, latent_dim=64, num_basis=10, hidden=128, h_bound=1.0, dh_clip=10.0, init_out_std=1e-3):
        super().__init__()
        self.latent_dim, self.num_basis = latent_dim, num_basis
        self.h_bound, self.dh_clip = h_bound, dh_clip
        self.ln = nn.LayerNorm(latent_dim)
        self.h_feat = KANFeatureMixer(latent_dim, num_basis, act=nn.Sigmoid())
        self.kan1 = KAN([latent_dim * num_basis, hidden])
        self.act = nn.SiLU()
        self.kan2 = KAN([hidden, hidden])
        self.out = nn.Linear(hidden, latent_dim, bias=True)
        nn.init.zeros_(self.out.bias)
        nn.init.normal_(self.out.weight, mean=0.0, std=init_out_std)
        self.log_alpha = nn.Parameter(torch.tensor(-3.0))
        self.scale = nn.Parameter(torch.tensor(1.0))
        self._gain_key, self._gain_val = None, None

    def _gain(self):
        """scale * softplus(log_alpha) (:480), one autograd node shared by every right-hand-side call between two
        parameter updates instead of three tiny launches (and their backward) per call."""
        key = (self.scale._version, self.log_alpha._version, self.scale.data_ptr(), self.log_alpha.data_ptr(),
               torch.is_grad_enabled())
        if self._gain_key != key or (self._gain_val.requires_grad and self._gain_val.grad_fn is None):
            # drop the stale node first: while it lives, its graph pins the parameters' AccumulateGrad nodes (and the
            # stream they were created on), which a later CUDA-graph capture on another stream cannot use
            self._gain_val = None
            self._gain_val = self.scale * F.softplus(self.log_alpha)
            self._gain_key = key
            if self._gain_val.requires_grad:
                # once this node has been backpropagated its saved tensors are gone: a second forward + backward
                # before any optimiser step (gradient accumulation) must build a new one
                self._gain_val.register_hook(self._gain_used)
        return self._gain_val

    def _gain_used(self, grad):
        self._gain_key = None
        return grad

    def forward(self, t, h):
        h = ops.ln_tanh(h, self.ln, self.h_bound)           # LayerNorm + h_bound * tanh(h / h_bound) :459-461, one kernel
        phi = self.h_feat(h, nan_to_num=True)                # sigmoid + nan_to_num fused into the neuron kernel :464-465
        z = self.act(self.kan1(phi))
        z = self.kan2(z)
        # act -> nn.Linear(hidden, latent) -> scale * softplus(log_alpha) (:471-480) as one product: SiLU is its
        # operand prologue, the gain its epilogue
        return ops.head(z, self.act, self.out, self._gain())


class No_MLP_KANODEFunc(nn.Module):
    """train_ecg_kan_fet_nn_ode.py:483-508."""

    def __init__(self, latent_dim=64, num_basis=10, hidden=128):
        super().__init__()
        self.latent_dim, self.num_basis = latent_dim, num_basis
        self.feat = KANFeatureMixer(latent_dim, num_basis, act=nn.Sigmoid())
        self.proj = nn.Linear(latent_dim * num_basis, latent_dim)
        nn.init.zeros_(self.proj.bias)
        nn.init.normal_(self.proj.weight, mean=0.0, std=0.01)

    def forward(self, t, h):
        dh = ops.linear(self.feat(h), self.proj)
        assert dh.shape == h.shape, (dh.shape, h.shape)
        return dh


class KANFetODEFunc(nn.Module):
    """train_ecg.py:986-1013: tanh-bound -> FerroelectricBasis -> tanh -> FerroelectricBasis -> nan_to_num -> clamp."""

    def __init__(self, latent_dim: int, hidden_dim: int, num_basis: int, h_bound: float = 1.0):
        super().__init__()
        self.h_bound = h_bound
        self.fc1 = FerroelectricBasis(latent_dim, hidden_dim, num_basis)
        self.act = nn.Tanh()
        self.fc2 = FerroelectricBasis(hidden_dim, latent_dim, num_basis)

    def _fusable(self) -> bool:
        return all(type(f) is FerroelectricBasis and not (f.use_noise and f.noise_std != 0.0) for f in (self.fc1, self.fc2)) \
            and isinstance(self.act, nn.Tanh) and self.h_bound > 0

    @staticmethod
    def _mode(f, B):
        """Hysteresis source of a FerroelectricBasis call: (prev rows or None, prev_mode) -- the reference's rule
        (ferro_class.py:373-378): fresh [1,...] zeros match only B == 1, any other shape mismatch re-creates prev = x."""
        r = f._rows
        if r is None:
            return None, (1 if B == 1 else 2)
        if r.shape[0] != B:
            return None, 2
        return (r if r.is_contiguous() else r.contiguous()), 0

    def forward(self, t, h):
        if h.dim() == 1:
            h = h.unsqueeze(0)
        if self._fusable() and h.is_cuda and h.dim() == 2:
            # two launches: [tanh bound -> fc1 -> tanh] and [fc2 -> nan_to_num -> clamp]; the glue of :1000-1011 runs
            # inside the neuron kernels (csrc/ferro.cu, pre_bound / post_op)
            ops.abort_if_tracing()
            x = ops.f32c(h)
            B = x.shape[0]
            p1, m1 = self._mode(self.fc1, B)
            hb = torch.empty_like(x)                                   # the bounded state = fc1's next prev_x
            z, _ = ops.ferro_apply(self.fc1, ops.FERRO_FULL, x, p1, prev_out=hb, pre_bound=self.h_bound, post_op=1,
                                   prev_mode=m1)
            self.fc1._set_rows(hb)
            p2, m2 = self._mode(self.fc2, B)
            dh, _ = ops.ferro_apply(self.fc2, ops.FERRO_FULL, z, p2, post_op=2, prev_mode=m2)
            self.fc2._set_rows(z.detach().clone())                     # own the state (z is saved for backward)
            return dh
        h = self.h_bound * torch.tanh(h / self.h_bound)
        z = self.act(self.fc1(h))
        dh = self.fc2(z)
        dh = torch.nan_to_num(dh, nan=0.0, posinf=1e3, neginf=-1e3)
        return torch.clamp(dh, -50.0, 50.0)


class KanFet_MLP_NODE(nn.Module):
    """train_ecg_kan_fet_nn_ode.py:985-1048: Linear(T -> latent) -> odeint(MLPKANODEFunc, linspace(0,1,9)) ->
    Dropout -> KANFeatureMixer -> Linear(latent*num_basis -> classes).  The reference hard-codes method="dopri5";
    ``ode_method`` (default "dopri5") additionally accepts "rk4" / "rk4_classic" / "midpoint" / "euler"."""

    def __init__(self, T: int, num_classes: int, latent_dim: int = 64, num_basis: int = 10, ode_hidden: int = 128,
                 dropout: float = 0.1, solver: str = "dopri5", rtol: float = 1e-3, atol: float = 1e-4):
        super().__init__()
        self.T, self.num_classes, self.latent_dim = T, num_classes, latent_dim
        self.solver, self.rtol, self.atol = solver, rtol, atol
        self.ode_method = "dopri5"
        self.encoder = nn.Linear(T, latent_dim)
        self.odefunc = MLPKANODEFunc(latent_dim=latent_dim, num_basis=num_basis, hidden=ode_hidden)
        self.dropout = nn.Dropout(dropout)
        self.cls_feat = KANFeatureMixer(latent_dim, num_basis, act=nn.Sigmoid())
        self.cls = nn.Linear(latent_dim * num_basis, num_classes)
        self._t_eval = torch.linspace(0.0, 1.0, 9, dtype=torch.float64)

    def forward(self, x):
        from .odeint import odeint
        h0 = self.encoder(x)
        # the grid stays on the host (the reference builds it on x.device, :1031): odeint only reads its values, and a
        # device tensor would cost a kernel plus a synchronising read-back per forward
        t_eval = self._t_eval
        h_traj = odeint(self.odefunc, h0, t_eval, method=self.ode_method, rtol=self.rtol, atol=self.atol)
        hT = self.dropout(h_traj[-1])
        return self.cls(self.cls_feat(hT))
