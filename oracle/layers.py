"""Oracle: layer math of the KAN-FET right-hand side (torch CPU, functional form).

TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.  Every function cites the reference
lines (relative to the reference checkout, commit 704a0df4) whose arithmetic it restates.
All functions are dtype-generic: run them on float32 tensors to reproduce the reference
bit for bit, or on float64 tensors to obtain a high-precision ground truth for gradient and
error-budget checks.

Parameter containers are plain dicts of tensors keyed by the reference's own parameter
and buffer names (so a reference ``state_dict()`` can be fed in directly).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
Params = Dict[str, Tensor]


# --------------------------------------------------------------------------------------
# B-spline basis  (efficient_kan/efficientkan.py:117-131, copy mnist_kuramoto_kan.py:101-110)
# --------------------------------------------------------------------------------------
def bspline_basis(x: Tensor, knots: Tensor, order: int = 3) -> Tensor:
    """Cox-de Boor recursion exactly as the reference evaluates it.

    x: [B, I]; knots: [I, G+2*order+1]  ->  [B, I, G+order].
    Level 0 is the half-open indicator ``t_j <= x < t_{j+1}`` (efficientkan.py:122);
    level p combines neighbours with two true divisions per term (:124-128).
    """
    xe = x.unsqueeze(-1)
    val = ((xe >= knots[:, :-1]) & (xe < knots[:, 1:])).to(x.dtype)
    for p in range(1, order + 1):
        lo = knots[:, : -(p + 1)]
        left = (xe - lo) / (knots[:, p:-1] - lo) * val[:, :, :-1]
        hi = knots[:, p + 1:]
        right = (hi - xe) / (hi - knots[:, 1:(-p)]) * val[:, :, 1:]
        val = left + right
    return val.contiguous()


def uniform_knots(in_features: int, grid_size: int = 5, order: int = 3,
                  grid_range=(-1.0, 1.0), dtype=torch.float32) -> Tensor:
    """Default knot table (efficientkan.py:55-61): int64 arange * python-float h + lo,
    evaluated in float32 by torch's type promotion."""
    h = (grid_range[1] - grid_range[0]) / grid_size
    row = torch.arange(-order, grid_size + order + 1) * h + grid_range[0]
    return row.to(dtype).expand(in_features, -1).contiguous()


def curve2coeff(x: Tensor, y: Tensor, knots: Tensor, order: int = 3) -> Tensor:
    """Least-squares spline coefficients (efficientkan.py:133-143): x [B, I], y [B, I, O] -> [O, I, G+order]."""
    A = bspline_basis(x, knots, order).transpose(0, 1)          # [I, B, S]
    Bm = y.transpose(0, 1)                                       # [I, B, O]
    return torch.linalg.lstsq(A, Bm).solution.permute(2, 0, 1).contiguous()


def update_grid(x: Tensor, p: Params, *, grid_size: int = 5, order: int = 3, grid_eps: float = 0.02,
                margin: float = 0.01):
    """KANLinear.update_grid (efficientkan.py:184-221): returns (new knots [I, G+2k+1], new spline_weight [O, I, S]).
    Sample-quantile knots blended with a uniform grid, `order` extrapolated knots either side, coefficients refitted so
    that the spline part reproduces its old outputs on the batch."""
    batch = x.size(0)
    splines = bspline_basis(x, p["grid"], order).permute(1, 0, 2)                       # [I, B, S]
    sw = p["spline_weight"] * p["spline_scaler"].unsqueeze(-1) if "spline_scaler" in p else p["spline_weight"]
    unreduced = torch.bmm(splines, sw.permute(1, 2, 0)).permute(1, 0, 2)                # [B, I, O]
    xs = torch.sort(x, dim=0)[0]
    adaptive = xs[torch.linspace(0, batch - 1, grid_size + 1, dtype=torch.int64)]
    step = (xs[-1] - xs[0] + 2 * margin) / grid_size
    uniform = torch.arange(grid_size + 1, dtype=torch.float32).unsqueeze(1) * step + xs[0] - margin
    grid = grid_eps * uniform + (1 - grid_eps) * adaptive
    grid = torch.concatenate([grid[:1] - step * torch.arange(order, 0, -1).unsqueeze(1), grid,
                              grid[-1:] + step * torch.arange(1, order + 1).unsqueeze(1)], dim=0)
    knots = grid.T.contiguous()
    return knots, curve2coeff(x, unreduced, knots, order)


def regularization_loss(p: Params, regularize_activation=1.0, regularize_entropy=1.0, regularize_logistic_l1=0.0):
    """efficientkan.py:223-237."""
    l1 = p["spline_weight"].abs().mean(-1)
    act = l1.sum()
    q = l1 / (act + 1e-12)
    reg = regularize_activation * act + regularize_entropy * (-torch.sum(q * (q + 1e-12).log()))
    if regularize_logistic_l1 != 0.0 and "logistic_weight" in p:
        reg = reg + regularize_logistic_l1 * p["logistic_weight"].abs().mean()
    return reg


# --------------------------------------------------------------------------------------
# Logistic basis  (efficientkan.py:20-24)
# --------------------------------------------------------------------------------------
def logistic_basis(x: Tensor, a: Tensor, b: Tensor) -> Tensor:
    """phi[b,i,j] = 2 / (1 + exp(-a[i,j] * (x[b,i] - b[i,j])))."""
    return 2.0 / (1.0 + torch.exp(-a * (x.unsqueeze(-1) - b)))


# --------------------------------------------------------------------------------------
# Ferroelectric neurons  (ferro_class.py)
# --------------------------------------------------------------------------------------
def ferro_core(x_exp: Tensor, prev: Tensor, br: Tensor, p: Params,
               gate_slope: float = 10.0, alpha: float = 0.8) -> Tuple[Tensor, Tensor]:
    """Shared soft-branch arithmetic (ferro_class.py:384-408 == :556-578 == :482-506).

    x_exp/prev/br broadcast against the parameter shape.  Returns (basis, target_sign);
    ``prev`` and ``br`` must already be detached snapshots (ferro_class.py:381-382).
    """
    dx = x_exp - prev
    up = torch.sigmoid(gate_slope * dx)
    cpos = torch.sigmoid(gate_slope * (x_exp - p["Ec"]))
    cneg = torch.sigmoid(gate_slope * (-x_exp - p["Ec"]))
    to_upper = up * cpos
    to_lower = (1 - up) * cneg
    tgt = to_upper * 1.0 + to_lower * (-1.0) + (1 - to_upper - to_lower) * br
    mom = alpha * br + (1.0 - alpha) * tgt
    shifted = x_exp + p["Ec"] * mom
    basis = p["Ps"] * torch.tanh(p["k"] * shifted) + p["bias"]
    return basis, tgt


class FerroFullState:
    """prev_x / branch_sign life-cycle of FerroelectricBasis (ferro_class.py:365-378,409)
    and NoisyFerroelectricBasis (:463-476,512-514)."""

    def __init__(self, p: Params):
        shape = (1,) + tuple(p["k"].shape)
        self.prev = torch.zeros(shape, dtype=p["k"].dtype)
        self.br = torch.ones(shape, dtype=p["k"].dtype)

    def reset(self):
        self.prev.zero_()
        self.br.fill_(1.0)


def ferro_full_forward(x: Tensor, p: Params, st: FerroFullState, *, gate_slope=10.0,
                       alpha=0.8, update_branch=False, return_basis=False):
    """FerroelectricBasis.forward (ferro_class.py:368-420); ``update_branch=True`` gives the
    Noisy variant's state update with noise_std=0 (:466-519).

    x: [B, I]; params [I, O, K]  ->  out [B, O].
    First-call rule (:373-378): if the stored buffers do not have shape [B,I,O,K] they are
    re-created as prev=x (so dx=0) and br=1.  With B==1 the initial [1,I,O,K] zeros match.
    """
    if x.dim() > 2:
        x = x.reshape(x.shape[0], -1)
    I, O, K = p["k"].shape
    x_exp = x.unsqueeze(2).unsqueeze(3).expand(-1, -1, O, K)
    if st.prev.shape != x_exp.shape or st.prev.dtype != x_exp.dtype:
        st.prev = x_exp.detach().clone()
    if st.br.shape != x_exp.shape or st.br.dtype != x_exp.dtype:
        st.br = torch.ones_like(x_exp)
    basis, tgt = ferro_core(x_exp, st.prev.clone(), st.br.clone(), p, gate_slope, alpha)
    st.prev = x_exp.detach().clone()
    if update_branch:
        st.br = tgt.detach().clone()
    out = (basis * p["coef"]).sum(dim=(1, 3))
    if return_basis:
        return out, basis.detach()
    return out


class FerroSeqState:
    """prev_x / branch_sign [1,I,O,K] of OriginalBatchedFerroelectricBasis (ferro_class.py:44-46)."""

    def __init__(self, p: Params):
        shape = (1,) + tuple(p["k"].shape)
        self.prev = torch.zeros(shape, dtype=p["k"].dtype)
        self.br = torch.ones(shape, dtype=p["k"].dtype)

    def reset(self):
        self.prev = torch.zeros_like(self.prev)
        self.br = torch.ones_like(self.br)


def ferro_seq_forward(x: Tensor, p: Params, st: FerroSeqState, *, gate_slope=10.0, alpha=0.8, return_basis=False):
    """OriginalBatchedFerroelectricBasis.forward (ferro_class.py:49-99): a Python loop over the batch in which sample b
    sees prev_x = x[b-1] and the branch momentum of sample b-1, which is also the value used in shifted_x (:73-77) and
    is NOT detached: the gradient flows through the recurrence (within one call; the state is detached between calls
    here, the reference keeps the stale graph in the buffer)."""
    if x.dim() > 2:
        x = x.reshape(x.shape[0], -1)
    I, O, K = p["k"].shape
    x_exp = x.unsqueeze(2).unsqueeze(3).expand(-1, -1, O, K)
    br, prev = st.br, st.prev
    all_basis = []
    for b in range(x.shape[0]):
        xs = x_exp[b:b + 1]
        dx = xs - prev
        up = torch.sigmoid(gate_slope * dx)
        cpos = torch.sigmoid(gate_slope * (xs - p["Ec"]))
        cneg = torch.sigmoid(gate_slope * (-xs - p["Ec"]))
        to_upper = up * cpos
        to_lower = (1 - up) * cneg
        tgt = to_upper * 1.0 + to_lower * (-1.0) + (1 - to_upper - to_lower) * br
        br = alpha * br + (1.0 - alpha) * tgt
        all_basis.append(p["Ps"] * torch.tanh(p["k"] * (xs + p["Ec"] * br)) + p["bias"])
        prev = xs.detach().clone()
    basis_all = torch.cat(all_basis, dim=0)
    st.prev, st.br = prev, br.detach().clone()
    out = (basis_all * p["coef"]).sum(dim=(1, 3))
    if return_basis:
        return out, basis_all.detach()
    return out


class FerroConvState:
    """_prev_x / _branch_sign / _state_ready of FerroelectricBasisConv2d (ferro_class.py:666-687) and of
    MemEfficient_FerroelectricBasisConv2d (:826-838): shape-checked, re-created as prev = x_exp, br = 1."""

    def __init__(self):
        self.prev = None
        self.br = None

    def reset(self):
        self.prev = self.br = None


def ferro_conv2d_forward(x: Tensor, p: Params, st: Optional[FerroConvState] = None, *, stride=1, padding=0,
                         gate_slope=10.0, alpha=0.8, stateful=False, mem_efficient=False, return_basis=False):
    """FerroelectricBasisConv2d.forward (ferro_class.py:689-779) / MemEfficient_...forward (:840-944).

    x: [B, Cin, H, W]; params k, Ec, Ps, bias, coef: [Cout, Cin, K, kH, kW]; out_bias [Cout] -> [B, Cout, Hout, Wout].
    Stateless (default, :724-727): prev = x_exp (dx = 0), branch = +1.  Stateful: prev <- x_exp after the call; the plain
    class also stores branch <- branch_momentum (:763-767) over the full (B,Cout,Cin,K,kH,kW,L) shape, the
    memory-efficient class keeps its Cout-free branch buffer at +1 (:927-931).  The plain class's stateful mode
    only runs for Cout == 1 (its branch buffer has no Cout axis; see the raise below)."""
    B, Cin, H, W = x.shape
    Cout, _, K, kH, kW = p["k"].shape
    xu = F.unfold(x, kernel_size=(kH, kW), padding=padding, stride=stride)
    L = xu.shape[-1]
    x_exp = xu.view(B, Cin, kH, kW, L).unsqueeze(2).expand(-1, -1, K, -1, -1, -1).unsqueeze(1)   # (B,1,Cin,K,kH,kW,L)
    q = {n: p[n].unsqueeze(0).unsqueeze(-1) for n in ("k", "Ec", "Ps", "bias", "coef")}
    if stateful:
        assert st is not None
        if st.prev is None or st.prev.shape != x_exp.shape or st.prev.dtype != x_exp.dtype:
            st.prev = x_exp.detach().clone()
            st.br = torch.ones_like(x_exp).detach()           # (B,1,...): no Cout axis in either class (:686, :837)
        prev, br = st.prev.detach().clone(), st.br.detach().clone()
    else:
        prev, br = x_exp.detach().clone(), torch.ones_like(x_exp).detach()
    dx = x_exp - prev
    up = torch.sigmoid(gate_slope * dx)
    cpos = torch.sigmoid(gate_slope * (x_exp - q["Ec"]))
    cneg = torch.sigmoid(gate_slope * (-x_exp - q["Ec"]))
    to_upper = up * cpos
    to_lower = (1.0 - up) * cneg
    tgt = to_upper * 1.0 + to_lower * (-1.0) + (1.0 - to_upper - to_lower) * br
    mom = alpha * br + (1.0 - alpha) * tgt
    basis = q["Ps"] * torch.tanh(q["k"] * (x_exp + q["Ec"] * mom)) + q["bias"]
    out = (basis * q["coef"]).sum(dim=(2, 3, 4, 5))
    if stateful:
        st.prev = x_exp.detach().clone()
        if not mem_efficient:
            # the reference does _branch_sign.copy_(branch_momentum) (:767) into the Cout-free buffer: this only
            # broadcasts for Cout == 1 and raises RuntimeError otherwise -- restated as is
            if mom.shape != st.br.shape:
                raise RuntimeError(f"output with shape {list(st.br.shape)} doesn't match the broadcast shape {list(mom.shape)}")
            st.br = mom.detach().clone()
    Hout = (H + 2 * padding - kH) // stride + 1
    Wout = (W + 2 * padding - kW) // stride + 1
    out = out.view(B, Cout, Hout, Wout) + p["out_bias"].view(1, -1, 1, 1)
    if return_basis:
        return out, basis.detach()
    return out


class Ferro2DState:
    """TwoDimensionFerroelectricBasis buffers (ferro_class.py:544-545): prev [1,I,K] zeros
    which *broadcast* on the first call (prev=0) and are re-created as [B,I,K] afterwards
    (:584-590).  branch_sign is never written, so it stays 1."""

    def __init__(self, p: Params):
        shape = (1,) + tuple(p["k"].shape)
        self.prev = torch.zeros(shape, dtype=p["k"].dtype)
        self.br = torch.ones(shape, dtype=p["k"].dtype)

    def reset(self):
        self.prev.zero_()
        self.br.fill_(1.0)


def ferro2d_forward(x: Tensor, p: Params, st: Ferro2DState, *, gate_slope=10.0, alpha=0.8) -> Tensor:
    """TwoDimensionFerroelectricBasis.forward (ferro_class.py:547-592): returns
    ``basis * coef`` of shape [B, I, K] (the LogisticBasis output contract)."""
    if x.dim() > 2:
        x = x.reshape(x.shape[0], -1)
    K = p["k"].shape[1]
    x_exp = x.unsqueeze(-1).expand(-1, -1, K)
    basis, _ = ferro_core(x_exp, st.prev.clone(), st.br.clone(), p, gate_slope, alpha)
    st.prev = x_exp.detach().clone()
    return basis * p["coef"]


class HardGateState:
    """prev_x buffer of the hysteretic LogisticBasis (train_ecg_kan_fet_nn_ode.py:93):
    shape [1,I,K]; after every call it holds the LAST batch row (:131-132)."""

    def __init__(self, p: Params):
        self.prev = torch.zeros((1,) + tuple(p["k"].shape), dtype=p["k"].dtype)
        self.gate = None

    def reset(self):
        self.prev.zero_()


def hardgate_forward(x: Tensor, p: Params, st: HardGateState, *, gate_slope=5.0,
                     breaking_point=0.5) -> Tensor:
    """Hysteretic two-branch logistic (train_ecg_kan_fet_nn_ode.py:102-134).
    Returns basis [B,I,K]; the 0/1 gate actually used is left in ``st.gate``."""
    if x.dim() != 2 or x.size(1) != p["k"].shape[0]:
        raise ValueError(f"x must be (B,{p['k'].shape[0]}), got {tuple(x.shape)}")
    K = p["k"].shape[1]
    x_exp = x.unsqueeze(-1).expand(-1, -1, K)
    up = p["Ps"] * (1 / (1 + torch.exp(-p["k"] * (x_exp - p["Ec"])))) * 2 - p["Ps"]
    down = p["Ps"] * (1 / (1 + torch.exp(-p["k"] * (x_exp + p["Ec"])))) * 2 - p["Ps"]
    g = torch.sigmoid(gate_slope * (x_exp - st.prev))
    gate = (g > breaking_point).to(x.dtype)
    st.gate = gate
    basis = gate * up + (1.0 - gate) * down + p["bias"]
    st.prev = x_exp[-1:, :, :].detach().clone()
    return basis


# --------------------------------------------------------------------------------------
# KANLinear / KAN / KANFET  (efficientkan.py:145-182, 274-279)
# --------------------------------------------------------------------------------------
def kanlinear_forward(x: Tensor, p: Params, *, order: int = 3, scale_logistic: float = 1.0,
                      branch: str = "logistic", branch_state=None,
                      gate_slope: float = 10.0, alpha: float = 0.8) -> Tensor:
    """KANLinear.forward (efficientkan.py:160-182).

    p keys: grid, base_weight, spline_weight, [spline_scaler], and for the branch
    [logistic_weight, logistic_scaler | logistic_bias] plus the branch-basis parameters
    under ``logistic_basis.a/.b`` (logistic) or ``logistic_basis.k/.Ec/.Ps/.bias/.coef`` (ferro2d).
    branch: "none" | "logistic" | "ferro2d".  ``logistic_bias`` (MNIST copy,
    mnist_kuramoto_kan.py:138-140) is added when present.
    """
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    O = p["base_weight"].shape[0]
    out = F.linear(F.silu(x2), p["base_weight"])
    sw = p["spline_weight"]
    if "spline_scaler" in p:
        sw = sw * p["spline_scaler"].unsqueeze(-1)           # efficientkan.py:145-147
    out = out + F.linear(bspline_basis(x2, p["grid"], order).view(x2.size(0), -1), sw.view(O, -1))
    if branch != "none":
        if branch == "logistic":
            phi = logistic_basis(x2, p["logistic_basis.a"], p["logistic_basis.b"])
        elif branch == "ferro2d":
            sub = {k.split(".", 1)[1]: v for k, v in p.items() if k.startswith("logistic_basis.")}
            phi = ferro2d_forward(x2, sub, branch_state, gate_slope=gate_slope, alpha=alpha)
        else:
            raise ValueError(branch)
        w = p["logistic_weight"] * scale_logistic            # efficientkan.py:153-157
        if "logistic_scaler" in p:
            w = w * p["logistic_scaler"].unsqueeze(-1)
        out = out + F.linear(phi.reshape(x2.size(0), -1), w, p.get("logistic_bias"))
    return out.reshape(*shape[:-1], O)


def kan_forward(x: Tensor, layers: List[Params], *, branch="logistic", states=None, **kw) -> Tensor:
    """KAN.forward (efficientkan.py:274-279): layers applied back to back, nothing between."""
    for li, p in enumerate(layers):
        x = kanlinear_forward(x, p, branch=branch,
                              branch_state=None if states is None else states[li], **kw)
    return x


def kanfet_states(layers: List[Params]) -> List[Ferro2DState]:
    return [Ferro2DState({k.split(".", 1)[1]: v for k, v in p.items() if k.startswith("logistic_basis.")})
            for p in layers]


# --------------------------------------------------------------------------------------
# Right-hand-side glue  (train_ecg_kan_fet_nn_ode.py:408-508, train_ecg.py:997-1013)
# --------------------------------------------------------------------------------------
def feature_mixer(x: Tensor, p: Params, st: HardGateState) -> Tensor:
    """KANFeatureMixer (train_ecg_kan_fet_nn_ode.py:417-420): sigmoid(hard-gate basis), flattened (i,j)."""
    return torch.sigmoid(hardgate_forward(x, p, st)).reshape(x.size(0), -1)


def mlp_kan_rhs(h: Tensor, p: Dict, st: HardGateState, *, h_bound=1.0, kan_branch="logistic") -> Tensor:
    """MLPKANODEFunc.forward (train_ecg_kan_fet_nn_ode.py:458-481).
    p: ln.weight/bias, basis (hard-gate params), kan1/kan2 (lists of layer dicts),
    out.weight/bias, log_alpha, scale."""
    h = F.layer_norm(h, (h.shape[-1],), p["ln.weight"], p["ln.bias"], 1e-5)
    h = h_bound * torch.tanh(h / h_bound)
    phi = feature_mixer(h, p["basis"], st)
    phi = torch.nan_to_num(phi, nan=0.0, posinf=1e3, neginf=-1e3)
    z = F.silu(kan_forward(phi, p["kan1"], branch=kan_branch))
    z = F.silu(kan_forward(z, p["kan2"], branch=kan_branch))
    dh = F.linear(z, p["out.weight"], p["out.bias"])
    return p["scale"] * F.softplus(p["log_alpha"]) * dh


def nomlp_kan_rhs(h: Tensor, p: Dict, st: HardGateState) -> Tensor:
    """No_MLP_KANODEFunc.forward (train_ecg_kan_fet_nn_ode.py:503-508)."""
    return F.linear(feature_mixer(h, p["basis"], st), p["proj.weight"], p["proj.bias"])


def kanfet_ode_rhs(h: Tensor, p1: Params, p2: Params, s1: FerroFullState, s2: FerroFullState,
                   *, h_bound=1.0) -> Tensor:
    """KANFetODEFunc.forward (train_ecg.py:997-1013)."""
    if h.dim() == 1:
        h = h.unsqueeze(0)
    h = h_bound * torch.tanh(h / h_bound)
    z = torch.tanh(ferro_full_forward(h, p1, s1))
    dh = ferro_full_forward(z, p2, s2)
    dh = torch.nan_to_num(dh, nan=0.0, posinf=1e3, neginf=-1e3)
    return torch.clamp(dh, -50.0, 50.0)


# --------------------------------------------------------------------------------------
# Logistic-KAN recurrent encoder  (train_kan_fet_ett.py:741-818)
# --------------------------------------------------------------------------------------
def kan_rnn_cell(x_t: Tensor, h_prev: Tensor, p: Params) -> Tensor:
    """FullyNonlinearKANCell.forward (train_kan_fet_ett.py:791-796).  p: input_basis.a/.b [F,nb], hidden_basis.a/.b [H,nb]."""
    H = p["hidden_basis.a"].shape[0]
    x_phi = logistic_basis(x_t, p["input_basis.a"], p["input_basis.b"]).view(x_t.size(0), -1)
    h_phi = logistic_basis(h_prev, p["hidden_basis.a"], p["hidden_basis.b"]).view(h_prev.size(0), -1)
    return torch.sigmoid(torch.cat([x_phi, h_phi], dim=1))[:, :H]


def kan_rnn_encoder(x_ctx: Tensor, p: Params) -> Tensor:
    """KANRNNEncoder.forward (train_kan_fet_ett.py:808-818).  p: rnn_cell.* (as kan_rnn_cell), to_latent.weight/.bias."""
    cell = {k[len("rnn_cell."):]: v for k, v in p.items() if k.startswith("rnn_cell.")}
    B, T, _ = x_ctx.shape
    h = torch.zeros(B, cell["hidden_basis.a"].shape[0], dtype=x_ctx.dtype)
    for t in range(T):
        h = kan_rnn_cell(x_ctx[:, t, :], h, cell)
    return F.linear(h, p["to_latent.weight"], p["to_latent.bias"])


# --------------------------------------------------------------------------------------
# Kuramoto lattice  (mnist_kuramoto_kan.py:175-204)
# --------------------------------------------------------------------------------------
def kuramoto_forward(x_img: Tensor, K: Tensor, omega: Tensor, steps: int = 10, dt: float = 0.15) -> Tensor:
    """[B,1,H,W] in [0,1] -> [B, 2*H*W] = [cos(theta), sin(theta)] after ``steps`` Euler steps
    of the 4-neighbour Kuramoto coupling (zero padding)."""
    B = x_img.shape[0]
    ker = torch.zeros(1, 1, 3, 3, dtype=x_img.dtype)
    ker[0, 0, 0, 1] = ker[0, 0, 2, 1] = ker[0, 0, 1, 0] = ker[0, 0, 1, 2] = 1.0
    theta = math.pi * (2.0 * x_img - 1.0)
    om = omega.expand(B, -1, -1, -1)
    for _ in range(steps):
        s, c = torch.sin(theta), torch.cos(theta)
        sn = F.conv2d(s, ker, padding=1)
        cn = F.conv2d(c, ker, padding=1)
        theta = theta + dt * (om + K * (c * sn - s * cn))
    return torch.cat([torch.cos(theta), torch.sin(theta)], dim=1).view(B, -1)


# --------------------------------------------------------------------------------------
# Seeded parameter factories (same distributions as the reference constructors; used to
# build identical weights for the oracle and the CUDA path on the GPU box).
# --------------------------------------------------------------------------------------
def make_kanlinear_params(I: int, O: int, gen: torch.Generator, *, grid_size=5, order=3, nb=10,
                          branch="logistic", mnist_style=False, dtype=torch.float32) -> Params:
    """Random KANLinear parameters with reference-like magnitudes.  NOT the reference's init
    procedure (that is ``fetode.kan.KANLinear.reset_parameters``); parity tests only need
    both sides to share the same numbers."""
    S = grid_size + order
    bound = 1.0 / math.sqrt(I)
    u = lambda *s: (torch.rand(*s, generator=gen, dtype=torch.float64) * 2 - 1)
    p = {
        "grid": uniform_knots(I, grid_size, order).to(dtype),
        "base_weight": (u(O, I) * bound).to(dtype),
        "spline_weight": (u(O, I, S) * 0.1).to(dtype),
    }
    p["spline_scaler"] = (u(O, I) * bound).to(dtype)
    if branch != "none":
        p["logistic_weight"] = (u(O, I * nb) * (1.0 / math.sqrt(I * nb))).to(dtype)
        if mnist_style:
            p["logistic_bias"] = (u(O) * 0.1).to(dtype)
        else:
            p["logistic_scaler"] = (1.0 + 0.1 * u(O)).to(dtype)
        if branch == "logistic":
            p["logistic_basis.a"] = torch.randn(I, nb, generator=gen, dtype=torch.float64).to(dtype)
            p["logistic_basis.b"] = torch.randn(I, nb, generator=gen, dtype=torch.float64).to(dtype)
        else:
            for k, v in make_ferro_params((I, nb), gen, dtype).items():
                p["logistic_basis." + k] = v
    return p


def make_ferro_params(shape, gen: torch.Generator, dtype=torch.float32) -> Params:
    """k,Ec in [0.5,2.5], Ps in [0.5,2], bias ~ 0.1 N(0,1), coef ~ N(0,1)
    (ferro_class.py:358-362 / :537-541 / train_ecg_kan_fet_nn_ode.py:87-91)."""
    r = lambda: torch.rand(*shape, generator=gen, dtype=torch.float64)
    n = lambda: torch.randn(*shape, generator=gen, dtype=torch.float64)
    return {
        "k": (r() * 2 + 0.5).to(dtype),
        "Ec": (r() * 2 + 0.5).to(dtype),
        "Ps": (r() * 1.5 + 0.5).to(dtype),
        "bias": (n() * 0.1).to(dtype),
        "coef": n().to(dtype),
    }


def to_dtype(obj, dtype):
    """Deep-convert a nest of dict/list/tensor to ``dtype`` (for float64 ground truth)."""
    if isinstance(obj, torch.Tensor):
        return obj.to(dtype) if obj.is_floating_point() else obj
    if isinstance(obj, dict):
        return {k: to_dtype(v, dtype) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(to_dtype(v, dtype) for v in obj)
    return obj
