"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the two training configurations of
BASELINE.json as whole models, built from oracle.layers / oracle.solvers.

  * ``kanfet_mlp_node_forward``  -- KanFet_MLP_NODE.forward, train_ecg_kan_fet_nn_ode.py:1026-1048:
        Linear(T -> latent) :1028 -> ODE solve of MLPKANODEFunc :458-481 over linspace(0, 1, 9) -> Dropout (p = 0 here)
        -> KANFeatureMixer :417-420 -> Linear(latent * nb -> classes) :1047.
    The reference solves with torchdiffeq dopri5 (absent, unpinned); this restatement uses the reference's OWN classic
    RK4 loop (integrate_rk4, :693-705; pinned by tests/golden/rk_loops_mlp) with 8 steps, the same choice the GPU
    benchmark makes (``ode_method = "rk4_classic"``).  The plain ``kan.KAN`` of the missing kan_diffusion submodule is
    substituted by efficientkan.KAN (parity unpinned for that submodule).
  * ``lv_train_loss``            -- one iteration of train_kanfet_node_predprey.py:246-257 with the fixed-step solver:
        pred = solve(calDeriv, X0, t_learn); loss = mean((pred - target)^2).

dtype-generic: fp32 reproduces the reference's arithmetic, fp64 is the ground truth for gradients.
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

from . import layers as OL
from . import solvers as OS

_STATE = ("prev_x", "branch_sign", "branch_state")


def _sub(sd: Dict[str, torch.Tensor], prefix: str) -> Dict[str, torch.Tensor]:
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix) and k.split(".")[-1] not in _STATE}


def ecg_params(state_dict: Dict[str, torch.Tensor], dtype=torch.float32, requires_grad=False) -> Dict:
    """state_dict of KanFet_MLP_NODE (reference names) -> nested oracle parameter dict."""
    def conv(t):
        t = t.detach().cpu().to(dtype).clone()
        return t
    sd = {k: conv(v) for k, v in state_dict.items() if k.split(".")[-1] not in _STATE}
    if requires_grad:
        for k, v in sd.items():
            if not k.endswith("grid"):
                v.requires_grad_(True)
    rhs = {"ln.weight": sd["odefunc.ln.weight"], "ln.bias": sd["odefunc.ln.bias"],
           "basis": _sub(sd, "odefunc.h_feat.basis."),
           "kan1": [_sub(sd, "odefunc.kan1.layers.0.")], "kan2": [_sub(sd, "odefunc.kan2.layers.0.")],
           "out.weight": sd["odefunc.out.weight"], "out.bias": sd["odefunc.out.bias"],
           "log_alpha": sd["odefunc.log_alpha"], "scale": sd["odefunc.scale"]}
    return {"flat": sd, "rhs": rhs, "encoder.weight": sd["encoder.weight"], "encoder.bias": sd["encoder.bias"],
            "cls_basis": _sub(sd, "cls_feat.basis."), "cls.weight": sd["cls.weight"], "cls.bias": sd["cls.bias"]}


class EcgStates:
    """The two hysteretic bases of the model (odefunc.h_feat.basis, cls_feat.basis): each carries the last batch row
    from call to call (train_ecg_kan_fet_nn_ode.py:131-132), also ACROSS forward passes."""

    def __init__(self, P: Dict):
        self.rhs = OL.HardGateState(P["rhs"]["basis"])
        self.cls = OL.HardGateState(P["cls_basis"])


def kanfet_mlp_node_forward(x: torch.Tensor, P: Dict, st: EcgStates, *, n_steps: int = 8, h_bound: float = 1.0,
                            kan_branch: str = "logistic") -> torch.Tensor:
    h0 = F.linear(x, P["encoder.weight"], P["encoder.bias"])                        # :1028
    f = lambda t, h: OL.mlp_kan_rhs(h, P["rhs"], st.rhs, h_bound=h_bound, kan_branch=kan_branch)
    hT = OS.rk4_classic(f, h0, 0.0, 1.0, n_steps)                                   # :693-705 (see module docstring)
    z = OL.feature_mixer(hT, P["cls_basis"], st.cls)                                # :1046
    return F.linear(z, P["cls.weight"], P["cls.bias"])                              # :1047


def lv_train_loss(layers, x0: torch.Tensor, target: torch.Tensor, t1: float, n_steps: int,
                  branch: str = "ferro2d") -> torch.Tensor:
    """train_kanfet_node_predprey.py:250-255 with classic RK4: fresh hysteresis state, solve, MSE over all samples."""
    states = OL.kanfet_states(layers) if branch == "ferro2d" else None
    f = lambda t, h: OL.kan_forward(h, layers, branch=branch, states=states)
    pred = OS.rk4_classic(f, x0, 0.0, t1, n_steps, trajectory=True)
    return ((pred - target) ** 2).mean()


def p_sample_loop(model, past, schedule_tensors, pred_len, d_in, noise):
    """kan_diffusion_ett.py:565-609 restated with the noise made explicit: noise[0] is the initial y (:576), noise[1 + k]
    the z drawn at the k-th iteration (ti = T-1-k, :602-603)."""
    betas, alphas, abar, _sa, _s1 = schedule_tensors
    T = betas.shape[0]
    B = past.shape[0]
    y = noise[0].clone()
    for k, ti in enumerate(reversed(range(T))):
        t = torch.full((B,), ti, dtype=torch.long)
        eps_hat = model(y, past, t)
        a_bar = abar[ti]
        sqrt_a_bar = torch.sqrt(a_bar)
        sqrt_1m_a_bar = torch.sqrt(1.0 - a_bar)
        y0_hat = (y - sqrt_1m_a_bar * eps_hat) / (sqrt_a_bar + 1e-8)
        beta_t, alpha_t = betas[ti], alphas[ti]
        a_bar_prev = abar[ti - 1] if ti > 0 else torch.tensor(1.0)
        beta_tilde = beta_t * (1.0 - a_bar_prev) / (1.0 - a_bar + 1e-8)
        coef1 = (torch.sqrt(a_bar_prev) * beta_t) / (1.0 - a_bar + 1e-8)
        coef2 = (torch.sqrt(alpha_t) * (1.0 - a_bar_prev)) / (1.0 - a_bar + 1e-8)
        mean = coef1 * y0_hat + coef2 * y
        y = mean + torch.sqrt(beta_tilde + 1e-8) * noise[1 + k] if ti > 0 else mean
    return y
