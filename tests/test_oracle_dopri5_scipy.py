"""Partial pin of the adaptive-solver restatement (a12, torchdiffeq dopri5: absent third-party dependency, PARITY UNPINNED
for the controller details).  What CAN be anchored on an independent implementation in this image:

  * the Dormand-Prince tableau (nodes, stage matrix, 5th-order weights) of oracle/solvers.py AND of the product's host
    controller fetode/dopri5.py equals scipy.integrate's RK45 tables (scipy 1.x, `scipy/integrate/_ivp/rk.py`);
  * the embedded error weights and the dense-output mid-point weights -- torchdiffeq uses Shampine's variant, which is
    NOT scipy's, so these are pinned by the order conditions they must satisfy (a 4th-order companion: E = b - b* kills
    polynomials up to degree 3 and not degree 4; y_mid is the solution at t + h/2 to 4th order);
  * the solver as a whole: same smooth problem, oracle dopri5 vs scipy RK45 at tight tolerances, both against the
    closed-form solution.
The CUDA tables (csrc/smallnet_wide.cu snw_dp5_tableau) are checked against the host controller on the GPU
(tests/test_dopri5_fused_gpu.py)."""
import math

import numpy as np
import pytest
import torch
from scipy.integrate import solve_ivp
from scipy.integrate._ivp.rk import RK45

from oracle import solvers as OS


def _tables(kind):
    if kind == "oracle":
        return OS._DP_ALPHA, OS._DP_BETA, OS._DP_CERR, OS._DP_CMID
    from fetode import dopri5 as D                       # host-side tables only: nothing here touches the GPU library
    return D._ALPHA, D._BETA, D._C_ERR, D._C_MID


@pytest.mark.parametrize("kind", ["oracle", "product"])
def test_tableau_equals_scipy_rk45(kind):
    alpha, beta, _cerr, _cmid = _tables(kind)
    assert np.allclose(np.array(alpha[:5], dtype=np.float64), RK45.C[1:], rtol=0, atol=1e-16)
    assert alpha[5] == 1.0
    for s in range(5):                                   # stage rows 1..5 of scipy's A
        assert np.allclose(np.array(beta[s], dtype=np.float64), RK45.A[s + 1, :s + 1], rtol=0, atol=1e-15), s
        assert math.isclose(sum(beta[s]), alpha[s], rel_tol=0, abs_tol=1e-15)      # row-sum condition
    assert np.allclose(np.array(beta[5], dtype=np.float64), RK45.B, rtol=0, atol=1e-16)   # FSAL: last row == weights


@pytest.mark.parametrize("kind", ["oracle", "product"])
def test_error_and_midpoint_weights_satisfy_their_order_conditions(kind):
    alpha, beta, cerr, cmid = _tables(kind)
    c = np.array([0.0] + list(alpha), dtype=np.float64)                  # nodes of the 7 stages
    A = np.zeros((7, 7))
    for s in range(6):
        A[s + 1, :s + 1] = beta[s]
    E = np.array(cerr, dtype=np.float64)
    b = np.array(list(beta[5]) + [0.0], dtype=np.float64)
    bstar = b - E                                                        # the embedded 4th-order weights
    # quadrature conditions up to order 4, tree conditions of order 3 and 4
    for w, order in ((b, 5), (bstar, 4)):
        for q in range(order):
            assert abs(w @ c ** q - 1.0 / (q + 1)) < 1e-14, (order, q)
        assert abs(w @ (A @ c) - 1.0 / 6) < 1e-14
        assert abs(w @ (A @ c ** 2) - 1.0 / 12) < 1e-14
        assert abs((w * c) @ (A @ c) - 1.0 / 8) < 1e-14
        assert abs(w @ (A @ (A @ c)) - 1.0 / 24) < 1e-14
    assert abs(bstar @ c ** 4 - 1.0 / 5) > 1e-4                          # ... and the companion is NOT 5th order
    assert abs(E.sum()) < 1e-15
    # Shampine's companion as torchdiffeq writes it
    assert np.allclose(bstar, [1951 / 21600, 0, 22642 / 50085, 451 / 720, -12231 / 42400, 649 / 6300, 1 / 60], atol=1e-15)
    # dense output: y_mid = y + h sum_i cmid_i k_i is the solution at t + h/2 to 4th order
    m = np.array(cmid, dtype=np.float64)
    for q in range(4):
        assert abs(m @ c ** q - 0.5 ** (q + 1) / (q + 1)) < 1e-12, q
    assert abs(m @ (A @ c) - 0.5 ** 3 / 6) < 1e-12


def test_oracle_dopri5_and_scipy_rk45_agree_on_a_smooth_problem():
    """y' = [[-0.5, 2], [-2, -0.5]] y: a damped rotation with a closed form.  Not a step-for-step comparison (the two
    controllers differ in their embedded pair and rejection rule) -- both must sit within their tolerance of the truth,
    and take a comparable number of steps."""
    M = torch.tensor([[-0.5, 2.0], [-2.0, -0.5]], dtype=torch.float64)
    f = lambda t, y: y @ M.T
    y0 = torch.tensor([[1.0, 0.5]], dtype=torch.float64)
    t = torch.linspace(0, 3.0, 13, dtype=torch.float64)
    stats = {}
    got = OS.odeint(f, y0, t, rtol=1e-8, atol=1e-10, stats=stats)[:, 0, :].numpy()
    ref = solve_ivp(lambda tt, y: M.numpy() @ y, (0.0, 3.0), y0[0].numpy(), method="RK45", t_eval=t.numpy(), rtol=1e-8,
                    atol=1e-10)
    tt = t.numpy()
    rot = np.stack([np.cos(2 * tt) * 1.0 + np.sin(2 * tt) * 0.5, -np.sin(2 * tt) * 1.0 + np.cos(2 * tt) * 0.5], 1)
    truth = np.exp(-0.5 * tt)[:, None] * rot
    assert np.abs(got - truth).max() < 5e-8
    assert np.abs(ref.y.T - truth).max() < 5e-7
    assert np.abs(got - ref.y.T).max() < 5e-7
    n_scipy = ref.nfev // 6
    assert 0.5 * n_scipy <= stats["accepted"] + stats["rejected"] <= 2.0 * n_scipy


@pytest.mark.parametrize("method,order", [("euler", 1), ("midpoint", 2), ("rk4", 4), ("rk4_classic", 4)])
def test_fixed_grid_restatements_converge_at_their_order(method, order):
    """torchdiffeq's fixed-grid solvers are restated too ("rk4" = the 3/8 rule); an independent property each must have:
    halving the step divides the global error by 2^order on a smooth problem with a closed form."""
    M = torch.tensor([[-0.5, 2.0], [-2.0, -0.5]], dtype=torch.float64)
    f = lambda t, y: y @ M.T
    y0 = torch.tensor([[1.0, 0.5]], dtype=torch.float64)
    T = 1.0
    truth = math.exp(-0.5 * T) * np.array([math.cos(2 * T) + 0.5 * math.sin(2 * T), -math.sin(2 * T) + 0.5 * math.cos(2 * T)])
    errs = []
    for n in (16, 32):
        if method == "rk4_classic":
            y = OS.rk4_classic(f, y0, 0.0, T, n)
        else:
            y = OS.odeint(f, y0, torch.linspace(0, T, n + 1, dtype=torch.float64), method=method)[-1]
        errs.append(np.abs(y[0].numpy() - truth).max())
    rate = math.log2(errs[0] / errs[1])
    assert abs(rate - order) < 0.25, (errs, rate)


@pytest.mark.parametrize("rtol,atol", [(1e-7, 1e-9), (1e-3, 1e-6)])
def test_initial_step_equals_scipy_select_initial_step(rtol, atol):
    """Hairer-Norsett-Wanner II.4 starting step, as torchdiffeq (_select_initial_step, order 4 -> exponent 1/5) and scipy
    (select_initial_step with the error-estimator order 4) both implement it: same number on the same problem."""
    from scipy.integrate._ivp.common import select_initial_step
    M = torch.tensor([[-0.5, 2.0], [-2.0, -0.5]], dtype=torch.float64)
    y0 = torch.tensor([[1.0, 0.5]], dtype=torch.float64)
    st = {}
    OS.odeint(lambda t, y: torch.tanh(y) @ M.T, y0, torch.tensor([0.0, 0.05], dtype=torch.float64), rtol=rtol, atol=atol,
              stats=st)
    fun = lambda t, y: M.numpy() @ np.tanh(y)
    y0n = y0[0].numpy()
    want = select_initial_step(fun, 0.0, y0n, np.inf, np.inf, fun(0.0, y0n), 1.0, 4, rtol, atol)
    assert math.isclose(st["first_step"], want, rel_tol=1e-9), (st["first_step"], want)


def test_product_dense_output_weights_equal_the_quartic_fit():
    """fetode/dopri5.py evaluates the dense output as ONE linear combination of (ya, yb, k_0..k_6) (host-side weights
    `_dense_weights`, also computed on the device by the solve kernel).  Against the definition it was derived from
    (torchdiffeq _interp_fit / _interp_evaluate: quartic through ya, y_mid, yb with end slopes), on random data."""
    from fetode import dopri5 as D
    rng = np.random.default_rng(0)
    for _ in range(20):
        ya, yb = rng.normal(size=3), rng.normal(size=3)
        ks = rng.normal(size=(7, 3))
        ta = rng.uniform(-1, 1)
        h = rng.uniform(0.01, 0.5)
        tb = ta + h
        tq = ta + rng.uniform(0, 1) * h
        y_mid = ya + h * (np.array(D._C_MID) @ ks)
        f0, f1 = ks[0], ks[6]
        a = 2 * h * (f1 - f0) - 8 * (yb + ya) + 16 * y_mid
        b = h * (5 * f0 - 3 * f1) + 18 * ya + 14 * yb - 32 * y_mid
        c = h * (f1 - 4 * f0) - 11 * ya - 5 * yb + 16 * y_mid
        x = (tq - ta) / (tb - ta)
        want = ya + x * h * f0 + x ** 2 * c + x ** 3 * b + x ** 4 * a
        w_ya, w_yb, w_k = D._dense_weights(ta, tb, tq, h)
        got = w_ya * ya + w_yb * yb + np.array(w_k) @ ks
        assert np.allclose(got, want, rtol=1e-12, atol=1e-12)
    assert D._dense_weights(0.0, 1.0, 1.0, 1.0) == (0.0, 1.0, [0.0] * 7)
    assert D._dense_weights(0.0, 1.0, 0.0, 1.0) == (1.0, 0.0, [0.0] * 7)
