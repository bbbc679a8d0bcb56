"""GPU parity of the differentiable fused Runge-Kutta path (csrc/smallnet_bwd.cu) and of the adaptive dopri5
driver (fetode/dopri5.py + csrc/rkstage.cu) against the oracle."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from helpers import fetode_kan_random, lv_inputs
from oracle import layers as OL
from oracle import solvers as OS

pytestmark = pytest.mark.gpu
GRAD_TOL = 5e-4
TRAJ_TOL = 1e-4


def _oracle64(layers, branch):
    l64 = [{k: v.double().requires_grad_(k != "grid") for k, v in l.items()} for l in layers]
    states = OL.kanfet_states(l64) if branch == "ferro2d" else None
    return l64, (lambda t, h: OL.kan_forward(h, l64, branch=branch, states=states))


@pytest.mark.parametrize("branch", ["none", "logistic", "ferro2d"])
@pytest.mark.parametrize("tableau", ["rk4_classic", "rk4", "midpoint", "euler"])
@pytest.mark.parametrize("B", [1, 70])
def test_lv_training_gradients_match_oracle_autograd(cuda, branch, tableau, B):
    """loss = mean((pred[:, :, :] - target)^2) over the whole trajectory, as train_kanfet_node_predprey.py:252-256."""
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=21, device=cuda)
    l64, f64 = _oracle64(layers, branch)
    n, dt = 12, 0.05
    t = torch.arange(n + 1, dtype=torch.float64) * dt
    x0 = lv_inputs(B, 3)
    x0[: max(1, B // 3)] = torch.rand(max(1, B // 3), 2) * 3 - 1.5       # some rows inside the knot span
    target = torch.randn(n + 1, B, 2, dtype=torch.float64)
    x64 = x0.double().requires_grad_(True)
    if tableau == "rk4_classic":
        pred64 = OS.rk4_classic(f64, x64, 0.0, n * dt, n, trajectory=True)
    else:
        pred64 = OS.odeint(f64, x64, t, method=tableau)
    loss64 = ((pred64 - target) ** 2).mean()
    loss64.backward()

    xg = x0.to(cuda).requires_grad_(True)
    pred = fetode.odeint(lambda tt, X: m(X), xg, t, method=tableau)
    assert pred.requires_grad
    assert max_rel(pred, pred64) < TRAJ_TOL
    loss = ((pred - target.float().to(cuda)) ** 2).mean()
    loss.backward()
    assert max_rel(xg.grad, x64.grad) < GRAD_TOL
    for li, layer in enumerate(m.layers):
        for name, tparam in layer.named_parameters():
            assert tparam.grad is not None, name
            assert max_rel(tparam.grad, l64[li][name].grad) < GRAD_TOL, (li, name)


def test_integrate_rk4_final_state_gradient(cuda):
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=5, device=cuda)
    l64, f64 = _oracle64(layers, "ferro2d")
    x0 = lv_inputs(130, 8)
    x64 = x0.double().requires_grad_(True)
    want = OS.rk4_classic(f64, x64, 0.0, 1.0, 8)
    want.square().sum().backward()
    xg = x0.to(cuda).requires_grad_(True)
    got = fetode.integrate_rk4(lambda t, h: m(h), xg, 0.0, 1.0, 8)
    got.square().sum().backward()
    assert max_rel(got, want) < TRAJ_TOL
    assert max_rel(xg.grad, x64.grad) < GRAD_TOL
    for li, layer in enumerate(m.layers):
        for name, tparam in layer.named_parameters():
            assert max_rel(tparam.grad, l64[li][name].grad) < GRAD_TOL, (li, name)


def test_fused_gradients_equal_staged_gradients(cuda):
    """One-launch reverse sweep == per-stage autograd through the any-width kernels."""
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=9, device=cuda)
    x0 = lv_inputs(300, 4, cuda)
    t = torch.arange(9, dtype=torch.float64) * 0.03
    res = []
    for fused in (True, False):
        m.reset_state()
        m.zero_grad()
        xg = x0.clone().requires_grad_(True)
        f = (lambda tt, X: m(X)) if fused else (lambda tt, X: m(X) * 1.0)
        fetode.odeint(f, xg, t, method="rk4").square().mean().backward()
        res.append((xg.grad.clone(), {n: p.grad.clone() for n, p in m.named_parameters()}))
    assert max_rel(res[0][0], res[1][0]) < GRAD_TOL
    for n in res[0][1]:
        assert max_rel(res[0][1][n], res[1][1][n]) < GRAD_TOL, n


@pytest.mark.parametrize("rtol,atol", [(1e-5, 1e-7), (1e-2, 1e-3)])
def test_dopri5_matches_oracle_restatement(cuda, rtol, atol):
    p, d = load_golden("rk_loops_mlp")
    import torch.nn.functional as F
    f_cpu = lambda t, z: F.linear(torch.tanh(F.linear(z, p["0.weight"], p["0.bias"])), p["2.weight"], p["2.bias"])
    pc = {k: v.to(cuda) for k, v in p.items()}
    f_gpu = lambda t, z: F.linear(torch.tanh(F.linear(z, pc["0.weight"], pc["0.bias"])), pc["2.weight"], pc["2.bias"])
    t = torch.linspace(0, 1, 9, dtype=torch.float64)
    s_ref, s_got = {}, {}
    with torch.no_grad():
        want = OS.odeint(f_cpu, d["z0"], t, rtol=rtol, atol=atol, stats=s_ref)
        got = fetode.odeint(f_gpu, d["z0"].to(cuda), t, rtol=rtol, atol=atol, options={"stats": s_got})
    assert got.shape == want.shape
    if rtol >= 1e-3:
        assert s_got == s_ref                                   # same accept / reject sequence
    else:                                                       # fp32 error estimates near the noise floor
        assert abs(s_got["accepted"] - s_ref["accepted"]) <= 0.02 * s_ref["accepted"] + 2
    assert max_rel(got, want) < TRAJ_TOL


def test_dopri5_on_kanfet_rhs_and_backward(cuda):
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=13, device=cuda)
    states = OL.kanfet_states(layers)
    f = lambda t, h: OL.kan_forward(h, layers, branch="ferro2d", states=states)
    x0 = torch.tensor([[1.0, 1.0]])
    t = torch.linspace(0, 0.5, 6, dtype=torch.float64)
    with torch.no_grad():
        want = OS.odeint(f, x0, t, rtol=1e-5, atol=1e-7)
    xg = x0.to(cuda).requires_grad_(True)
    got = fetode.odeint(lambda tt, X: m(X), xg, t, rtol=1e-5, atol=1e-7)
    assert max_rel(got, want) < TRAJ_TOL
    got[:, 0, :].square().mean().backward()
    assert xg.grad is not None and torch.isfinite(xg.grad).all()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())


@pytest.mark.parametrize("B", [1, 5, 300])
@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
def test_both_training_kernel_pairs_agree_with_oracle_autograd(cuda, branch, B, monkeypatch):
    """The hidden-units-on-lanes pair (csrc/smallnet_wide.cu, small batches) and the lane-per-trajectory pair
    (smallnet_v2.cu + smallnet_bwd.cu, large batches) are selected by batch size; force each in turn on the same
    problem: trajectory, input gradient and every parameter gradient against fp64 oracle autograd."""
    from fetode import training
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=31 + B, device=cuda)
    l64 = [{k: v.double().requires_grad_(k != "grid") for k, v in lay.items()} for lay in layers]
    x0 = lv_inputs(B, 3) if B > 1 else torch.ones(1, 2)
    n, T = 12, 0.6
    gen = torch.Generator().manual_seed(5)
    w = torch.randn(n + 1, B, 2, generator=gen, dtype=torch.float64)
    x64 = x0.double().requires_grad_(True)
    st = OL.kanfet_states(l64) if branch == "ferro2d" else None
    f = lambda t, h: OL.kan_forward(h, l64, branch=branch, states=st)
    want = OS.rk4_classic(f, x64, 0.0, T, n, trajectory=True)
    (want * w).sum().backward()
    t = torch.linspace(0, T, n + 1, dtype=torch.float64)
    for mode in ("1", "0"):
        monkeypatch.setattr(training, "_WIDE_ENV", mode)
        m.zero_grad()
        if branch == "ferro2d":
            m.reset_state()
        xg = x0.to(cuda).requires_grad_(True)
        got = fetode.odeint(lambda tt, X: m(X), xg, t, method="rk4_classic")
        assert max_rel(got, want) < 1e-4, mode
        (got * w.float().to(cuda)).sum().backward()
        assert max_rel(xg.grad, x64.grad) < 5e-4, mode
        for li, layer in enumerate(m.layers):
            for name, p in layer.named_parameters():
                assert max_rel(p.grad, l64[li][name].grad) < 5e-4, (mode, li, name)
