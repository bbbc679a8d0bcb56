"""GPU parity of the fused small-width KAN / KANFET kernels (csrc/smallnet.cu) through the C ABI.

Bars (BASELINE.json north_star): per-RHS outputs within 1e-5 relative (max-norm, fp32), integrated
trajectories within 1e-4 relative after the horizon."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from helpers import fetode_kan_from, fetode_kan_random, lv_inputs, split_layers
from oracle import layers as OL
from oracle import solvers as OS

pytestmark = pytest.mark.gpu

RHS_TOL = 1e-5
TRAJ_TOL = 1e-4


def _oracle_rhs(layers, branch):
    states = OL.kanfet_states(layers) if branch == "ferro2d" else None
    return (lambda t, h: OL.kan_forward(h, layers, branch=branch, states=states)), states


@pytest.mark.parametrize("branch", ["none", "logistic", "ferro2d"])
@pytest.mark.parametrize("B", [1, 33, 4096])
def test_rhs_matches_oracle(cuda, branch, B):
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=3, device=cuda)
    assert m.small_plan() is not None
    f, _ = _oracle_rhs(layers, branch)
    x = lv_inputs(B, 5)
    x[: max(1, B // 4)] = torch.rand(max(1, B // 4), 2) * 4.4 - 2.2        # inside the knot span
    with torch.no_grad():
        for call in range(3):                                               # hysteresis state advances
            xi = x + 0.07 * call
            want = f(None, xi)
            got = m(xi.to(cuda))
            assert max_rel(got, want) < RHS_TOL, (branch, B, call)


@pytest.mark.parametrize("name,ferro", [("kan_2_10_2", False), ("kanfet_2_10_2", True)])
def test_rhs_and_rk4_match_reference_golden(cuda, name, ferro):
    params, d = load_golden(name)
    m = fetode_kan_from(params, [2, 10, 2], ferro, cuda)
    with torch.no_grad():
        if ferro:
            for i in range(3):
                got = m(d[f"x{i}"].to(cuda))
                assert max_rel(got, d[f"y{i}"]) < RHS_TOL, i
            m.reset_state()
            x0 = d["x0"]
        else:
            assert max_rel(m(d["x"].to(cuda)), d["y"]) < RHS_TOL
            x0 = d["x"]
        t = torch.linspace(0.0, 0.14, 11, dtype=torch.float64)
        traj = fetode.odeint(lambda t, X: m(X), x0.to(cuda), t, method="rk4_classic")
    assert traj.shape == d["traj"].shape
    assert max_rel(traj, d["traj"]) < TRAJ_TOL


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
@pytest.mark.parametrize("tableau", ["euler", "midpoint", "rk4_classic", "rk4"])
def test_fixed_step_trajectories_match_oracle(cuda, branch, tableau):
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=11, device=cuda)
    f, states = _oracle_rhs(layers, branch)
    x0 = lv_inputs(257, 2)
    n, dt = 50, 0.014
    t = torch.arange(n + 1, dtype=torch.float64) * dt
    with torch.no_grad():
        if tableau == "rk4_classic":
            want = OS.rk4_classic(f, x0, 0.0, n * dt, n, trajectory=True)
        else:
            want = OS.odeint(f, x0, t, method=tableau)
        got = fetode.odeint(lambda tt, X: m(X), x0.to(cuda), t, method=tableau)
    assert got.shape == want.shape
    assert max_rel(got[-1], want[-1]) < TRAJ_TOL
    assert max_rel(got, want) < TRAJ_TOL


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
@pytest.mark.parametrize("tableau", ["euler", "midpoint", "rk4_classic", "rk4"])
def test_lane_per_trajectory_kernel_all_tableaus_above_wide_threshold(cuda, branch, tableau):
    """13 001 rows (odd: a half-filled pair and a ragged last CTA) is above the hidden-units-on-lanes threshold, so the
    split-lane kernel of csrc/smallnet_v2.cu runs (two lanes per trajectory pair, the classic and the 3/8 register
    layouts); every row against the oracle, plus the hysteresis state left behind."""
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=12, device=cuda)
    f, states = _oracle_rhs(layers, branch)
    B = 13001
    x0 = lv_inputs(B, 3)
    x0[:4000] = torch.rand(4000, 2) * 3 - 1.5                     # rows inside the knot span
    n, dt = 12, 0.014
    t = torch.arange(n + 1, dtype=torch.float64) * dt
    with torch.no_grad():
        if tableau == "rk4_classic":
            want = OS.rk4_classic(f, x0, 0.0, n * dt, n, trajectory=True)
        else:
            want = OS.odeint(f, x0, t, method=tableau)
        got = fetode.odeint(lambda tt, X: m(X), x0.to(cuda), t, method=tableau)
    assert got.shape == want.shape
    assert max_rel(got, want) < TRAJ_TOL
    assert torch.isfinite(got).all()


def test_fused_path_equals_staged_path(cuda):
    """The one-launch fused integration and the stage-by-stage driver agree."""
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=4, device=cuda)
    x0 = lv_inputs(1000, 9, cuda)
    t = torch.arange(21, dtype=torch.float64) * 0.014
    with torch.no_grad():
        fused = fetode.odeint(lambda tt, X: m(X), x0, t, method="rk4")
        m.reset_state()
        staged = fetode.odeint(lambda tt, X: m(X) * 1.0, x0, t, method="rk4")   # '* 1.0' defeats recognition
    assert max_rel(fused, staged) < 1e-5


def test_ragged_and_empty_batches(cuda):
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=6, device=cuda)
    f, _ = _oracle_rhs(layers, "ferro2d")
    with torch.no_grad():
        for B in (0, 1, 127, 128, 129):
            m.reset_state()
            x = lv_inputs(max(B, 1), B)[:B]
            got = m(x.to(cuda))
            assert got.shape == (B, 2)
            if B:
                f2, _ = _oracle_rhs(layers, "ferro2d")
                assert max_rel(got, f2(None, x)) < RHS_TOL


def test_long_horizon_properties_full_size(cuda):
    """At the benchmark's own size (2^20 trajectories) the oracle is too slow; check size-independent
    properties instead: (i) trajectories are independent -- the first 2^18 rows of the big run equal a 2^18-row run
    bit for bit (same launch shape), and a 512-row run (a different launch shape: two lanes per trajectory pair, other
    summation order) to the trajectory tolerance; (ii) integrating 2 x 50 steps with the state carried equals 1 x 100."""
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=8, device=cuda)
    plan = m.small_plan()
    from fetode._lib import TABLEAU
    B = 1 << 20
    x0 = lv_inputs(B, 1, cuda)
    st = torch.empty(B, plan.state_floats, device=cuda)
    big = plan.rk_fwd(TABLEAU["rk4_classic"], x0, 0.014, 100, state=st, state_is_zero=True)
    n_mid = 1 << 18
    st_m = torch.empty(n_mid, plan.state_floats, device=cuda)
    mid = plan.rk_fwd(TABLEAU["rk4_classic"], x0[:n_mid].contiguous(), 0.014, 100, state=st_m, state_is_zero=True)
    assert torch.equal(big[:n_mid], mid)
    st_s = torch.empty(512, plan.state_floats, device=cuda)
    small = plan.rk_fwd(TABLEAU["rk4_classic"], x0[:512].contiguous(), 0.014, 100, state=st_s, state_is_zero=True)
    assert max_rel(small, big[:512]) < TRAJ_TOL
    st2 = torch.empty(B, plan.state_floats, device=cuda)
    half = plan.rk_fwd(TABLEAU["rk4_classic"], x0, 0.014, 50, state=st2, state_is_zero=True)
    full = plan.rk_fwd(TABLEAU["rk4_classic"], half, 0.014, 50, state=st2, state_is_zero=False)
    assert torch.equal(full, big)
    assert torch.equal(st2, st)
    assert torch.isfinite(big).all()


@pytest.mark.parametrize("branch", ["logistic", "ferro2d"])
def test_rhs_on_adapted_non_uniform_grid(cuda, branch):
    """update_grid (efficientkan.py:184-221) leaves per-feature NON-uniform knots: the fused kernel's interval search
    and its per-interval cubic tables must follow the stored knots, not a closed-form spacing."""
    layers, m = fetode_kan_random([2, 10, 2], branch, seed=17, device=cuda)
    gen = torch.Generator().manual_seed(3)
    for li, lay in enumerate(layers):
        g = lay["grid"]
        step = (torch.rand(g.shape, generator=gen) * 0.9 + 0.1) * 0.4            # spacings in [0.04, 0.4]
        newg = (torch.cumsum(step, dim=1) - 2.3).float().contiguous()
        lay["grid"] = newg
        m.layers[li].grid.copy_(newg.to(cuda))
        torch.autograd.graph.increment_version(m.layers[li].base_weight)          # repack
    f, _ = _oracle_rhs(layers, branch)
    x = torch.rand(777, 2, generator=gen) * 5.0 - 2.5
    x[:12] = layers[0]["grid"][0, :12, None].expand(12, 2)                        # knot-exact inputs
    with torch.no_grad():
        for call in range(2):
            xi = x + 0.05 * call
            assert max_rel(m(xi.to(cuda)), f(None, xi)) < RHS_TOL, (branch, call)


def test_benchmark_horizon_1000_steps_vs_oracle(cuda):
    """BASELINE configs[1] horizon: 1000 classic-RK4 steps, dt = 0.014 -- trajectories within 1e-4 (max-norm relative)
    of the fp32 oracle (the reference's arithmetic) for 256 rows, every dumped sample, and all states finite."""
    layers, m = fetode_kan_random([2, 10, 2], "ferro2d", seed=8, device=cuda)
    f, _ = _oracle_rhs(layers, "ferro2d")
    x0 = lv_inputs(256, 1)
    n, dt = 1000, 0.014
    with torch.no_grad():
        want = OS.rk4_classic(f, x0, 0.0, n * dt, n, trajectory=True)
        t = torch.arange(n + 1, dtype=torch.float64) * dt
        got = fetode.odeint(lambda tt, X: m(X), x0.to(cuda), t, method="rk4_classic")
    assert torch.isfinite(got).all()
    err_T = max_rel(got[-1], want[-1])
    err_all = max_rel(got, want)
    print(f"1000-step horizon: final-state rel err {err_T:.3e}, whole-trajectory rel err {err_all:.3e}")
    assert err_T < TRAJ_TOL and err_all < TRAJ_TOL
