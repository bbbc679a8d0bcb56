"""The fused adaptive solve (fetode/dopri5.py ``_FusedSolve`` + fetode_smallnet_attempt_fwd / _bwd, csrc/smallnet_wide.cu):
``odeint(lambda t, X: model(X), X0, t)`` with torchdiffeq's default method, the exact call of
train_kanfet_node_predprey.py:252.  One launch per attempted step; checked against the oracle's dopri5 restatement
(oracle/solvers.py, PARITY UNPINNED like every torchdiffeq restatement here), against the staged path of the same
driver, and -- gradients -- against fp64 oracle autograd."""
import pytest
import torch

import fetode
from conftest import max_rel
from helpers import fetode_kan_random, lv_inputs
from oracle import layers as OL
from oracle import solvers as OS

pytestmark = pytest.mark.gpu
TRAJ_TOL = 1e-4
GRAD_TOL = 2e-3          # fp32 step sequence vs fp64 step sequence: both within rtol of the ODE solution
GRAD_TOL_HYST = 2e-2     # ferro2d: the gate sees the previous STAGE input, so the solution depends on the step sequence at
                         # O(dt): fp32 and fp64 controllers pick different sequences.  The tight check for the hysteretic
                         # model is test_fused_dopri5_gradient_equals_staged_path_gradient (same controller, same steps).


def _rhs_cpu(layers, branch, dtype=torch.float32, grad=False):
    ls = [{k: (v.to(dtype).requires_grad_(k != "grid") if grad else v.to(dtype)) for k, v in l.items()} for l in layers]
    st = OL.kanfet_states(ls) if branch == "ferro2d" else None
    return ls, (lambda t, h: OL.kan_forward(h, ls, branch=branch, states=st))


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
@pytest.mark.parametrize("B", [1, 7, 100])
def test_fused_dopri5_matches_oracle_and_staged_path(cuda, branch, B, monkeypatch):
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=3 + B, device=cuda)
    _, f = _rhs_cpu(layers, branch)
    x0 = lv_inputs(B, 5) if B > 1 else torch.tensor([[1.0, 1.0]])
    t = torch.linspace(0, 0.75, 16, dtype=torch.float64)
    s_ref, s_fused, s_staged = {}, {}, {}
    with torch.no_grad():
        want = OS.odeint(f, x0, t, rtol=1e-5, atol=1e-7, stats=s_ref)
        got = fetode.odeint(lambda tt, X: m(X), x0.to(cuda), t, rtol=1e-5, atol=1e-7, options={"stats": s_fused})
        state_fused = m.hysteresis_state(B, cuda).clone() if branch == "ferro2d" else None
        monkeypatch.setenv("FETODE_DOPRI5_FUSED", "0")
        if branch == "ferro2d":
            m.reset_state()
        staged = fetode.odeint(lambda tt, X: m(X), x0.to(cuda), t, rtol=1e-5, atol=1e-7, options={"stats": s_staged})
    assert s_fused.get("fused") is True and "fused" not in s_staged
    assert got.shape == want.shape == (16, B, 2)
    assert torch.equal(got[0].cpu(), x0)
    assert max_rel(got, want) < TRAJ_TOL
    assert max_rel(got, staged) < TRAJ_TOL
    for s in (s_fused, s_staged):
        assert abs(s["accepted"] - s_ref["accepted"]) <= 0.05 * s_ref["accepted"] + 2
    if branch == "ferro2d" and {k: v for k, v in s_fused.items() if k != "fused"} == s_staged:
        # the hysteresis snapshot left behind is the input of the LAST right-hand side evaluated (accepted or not);
        # comparable when both paths took the same accept / reject sequence
        st_staged = m.hysteresis_state(B, cuda)
        assert max_rel(state_fused, st_staged) < 1e-3


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
@pytest.mark.parametrize("B", [1, 12])
def test_fused_dopri5_gradients_match_oracle_autograd(cuda, branch, B):
    """loss over the whole trajectory as train_kanfet_node_predprey.py:252-256; the adjoints of the dense-output
    combinations, of the FSAL slope and of the first right-hand side all enter."""
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=17 + B, device=cuda)
    l64, f64 = _rhs_cpu(layers, branch, torch.float64, grad=True)
    x0 = lv_inputs(B, 9) if B > 1 else torch.tensor([[1.0, 1.0]])
    t = torch.linspace(0, 0.4, 5, dtype=torch.float64)
    gen = torch.Generator().manual_seed(2)
    w = torch.randn(5, B, 2, generator=gen, dtype=torch.float64)
    tol = GRAD_TOL_HYST if branch == "ferro2d" else GRAD_TOL
    x64 = x0.double().requires_grad_(True)
    want = OS.odeint(f64, x64, t, rtol=1e-6, atol=1e-8)
    (want * w).sum().backward()

    xg = x0.to(cuda).requires_grad_(True)
    st = {}
    got = fetode.odeint(lambda tt, X: m(X), xg, t, rtol=1e-6, atol=1e-8, options={"stats": st})
    assert st.get("fused") is True and got.requires_grad
    assert max_rel(got, want) < TRAJ_TOL
    (got * w.float().to(cuda)).sum().backward()
    assert max_rel(xg.grad, x64.grad) < tol
    for li, layer in enumerate(m.layers):
        for name, p in layer.named_parameters():
            assert p.grad is not None, name
            assert max_rel(p.grad, l64[li][name].grad) < tol, (li, name)


def test_fused_dopri5_gradient_equals_staged_path_gradient(cuda, monkeypatch):
    """Same driver, same controller: the fused reverse sweep against autograd through the staged path's kernels."""
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=23, device=cuda)
    x0 = lv_inputs(9, 4)
    t = torch.linspace(0, 0.8, 9, dtype=torch.float64)
    res = []
    for fused in ("1", "0"):
        monkeypatch.setenv("FETODE_DOPRI5_FUSED", fused)
        m.zero_grad()
        m.reset_state()
        xg = x0.to(cuda).requires_grad_(True)
        out = fetode.odeint(lambda tt, X: m(X), xg, t, rtol=1e-5, atol=1e-7)
        out[1:].square().mean().backward()
        res.append((out.detach(), xg.grad.clone(), {n: p.grad.clone() for n, p in m.named_parameters()}))
    assert max_rel(res[0][0], res[1][0]) < TRAJ_TOL
    assert max_rel(res[0][1], res[1][1]) < GRAD_TOL
    for n in res[0][2]:
        assert max_rel(res[0][2][n], res[1][2][n]) < GRAD_TOL, n


def test_fused_dopri5_single_time_and_first_step_option(cuda):
    _, m = fetode_kan_random([2, 10, 2], "logistic", seed=1, device=cuda)
    x0 = torch.ones(3, 2, device=cuda)
    with torch.no_grad():
        one = fetode.odeint(lambda tt, X: m(X), x0, torch.tensor([0.0]))
        assert one.shape == (1, 3, 2) and torch.equal(one[0], x0)
        st = {}
        two = fetode.odeint(lambda tt, X: m(X), x0, torch.tensor([0.0, 0.3]), rtol=1e-4, atol=1e-6,
                            options={"first_step": 0.05, "stats": st})
    assert two.shape == (2, 3, 2) and st["accepted"] >= 1 and torch.isfinite(two).all()


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
def test_device_controller_equals_host_controller(cuda, branch, monkeypatch):
    """B <= 12: the accept / reject loop itself runs in the kernel (fetode_smallnet_dp5_solve_fwd / _bwd).  Same
    arithmetic per attempt as the host-controlled launches, so values, step counts and gradients agree closely."""
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=41, device=cuda)
    x0 = lv_inputs(5, 6)
    t = torch.linspace(0, 0.9, 10, dtype=torch.float64)
    res = []
    for dev_ctl in ("1", "0"):
        monkeypatch.setenv("FETODE_DOPRI5_DEVICE", dev_ctl)
        m.zero_grad()
        if branch == "ferro2d":
            m.reset_state()
        st = {}
        xg = x0.to(cuda).requires_grad_(True)
        out = fetode.odeint(lambda tt, X: m(X), xg, t, rtol=1e-5, atol=1e-7, options={"stats": st})
        assert st.get("fused") is True and bool(st.get("device_controller")) == (dev_ctl == "1")
        out[1:].square().mean().backward()
        state = m.hysteresis_state(5, cuda).clone() if branch == "ferro2d" else None
        res.append((out.detach(), xg.grad.clone(), {n: p.grad.clone() for n, p in m.named_parameters()}, st, state))
    a, b = res
    assert abs(a[3]["accepted"] - b[3]["accepted"]) <= 2 and abs(a[3]["rejected"] - b[3]["rejected"]) <= 2
    assert max_rel(a[0], b[0]) < 2e-5
    if a[3]["accepted"] == b[3]["accepted"] and a[3]["rejected"] == b[3]["rejected"]:
        assert max_rel(a[1], b[1]) < 1e-4
        for n in a[2]:
            assert max_rel(a[2][n], b[2][n]) < 1e-4, n
        if state is not None:
            assert max_rel(a[4], b[4]) < 1e-4


def test_device_controller_resumes_across_launches(cuda, monkeypatch):
    """Checkpoint chunks of 3 accepted steps and 2 attempts per launch: the solve is re-launched many times and must
    give what one launch gives (the controller state travels in ctrl[8], the adjoint carries in carry_y / carry_k)."""
    from fetode import dopri5
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=43, device=cuda)
    x0 = lv_inputs(12, 2)
    t = torch.linspace(0, 1.0, 6, dtype=torch.float64)
    res = []
    for chunk, attempts in ((4096, 1 << 16), (3, 2)):
        monkeypatch.setattr(dopri5._FusedSolve, "CHUNK", chunk)
        monkeypatch.setattr(dopri5._FusedSolve, "ATTEMPTS", attempts)
        m.zero_grad()
        m.reset_state()
        st = {}
        xg = x0.to(cuda).requires_grad_(True)
        out = fetode.odeint(lambda tt, X: m(X), xg, t, rtol=1e-7, atol=1e-9, options={"stats": st})
        assert st.get("device_controller") is True
        (out * torch.arange(1, 7, device=cuda).view(6, 1, 1)).sum().backward()
        res.append((out.detach(), xg.grad.clone(), {n: p.grad.clone() for n, p in m.named_parameters()}, st))
    a, b = res
    assert a[3] == b[3] and a[3]["accepted"] >= 10, a[3]
    assert torch.equal(a[0], b[0])
    assert max_rel(a[1], b[1]) < 1e-5
    for n in a[2]:
        assert max_rel(a[2][n], b[2][n]) < 1e-5, n
