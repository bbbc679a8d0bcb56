"""The oracle reproduces every golden vector under tests/golden/ (outputs of the live reference,
see oracle/make_golden.py).  This is what pins the oracle on the GPU box, where the reference
checkout does not exist.  Tolerance 2e-6 max-norm: libm / vector-width differences between hosts."""
import torch

from conftest import load_golden, max_rel
from oracle import layers as OL
from oracle import solvers as OS

TOL = 2e-6
FERRO_KEYS = ("k", "Ec", "Ps", "bias", "coef")


def _layers(params, n):
    return [{k[len(f"layers.{i}."):]: v for k, v in params.items() if k.startswith(f"layers.{i}.")} for i in range(n)]


def test_kanlinear_golden():
    for name in ("kanlinear_2x10", "kanlinear_10x2", "kanlinear_48x24"):
        p, d = load_golden(name)
        assert max_rel(OL.kanlinear_forward(d["x"], p), d["y"]) < TOL
        assert max_rel(OL.bspline_basis(d["x"], p["grid"]), d["bsplines"]) < TOL
    p, d = load_golden("kanlinear_nobranch_adapted")
    assert max_rel(OL.kanlinear_forward(d["x"], p, branch="none"), d["y"]) < TOL


def test_bspline_known_answers():
    """SURVEY appendix A.2 spot values on the default grid."""
    g = OL.uniform_knots(1)
    v = OL.bspline_basis(torch.tensor([[-1.0], [0.0], [1.0], [2.3], [-2.3]]), g)[:, 0]
    assert torch.allclose(v[0], torch.tensor([1 / 6, 2 / 3, 1 / 6, 0, 0, 0, 0, 0]), atol=1e-6)
    assert torch.allclose(v[1], torch.tensor([0, 0, 1 / 48, 23 / 48, 23 / 48, 1 / 48, 0, 0]), atol=1e-6)
    assert torch.allclose(v[2], torch.tensor([0, 0, 0, 0, 0, 1 / 6, 2 / 3, 1 / 6]), atol=1e-6)
    assert torch.count_nonzero(v[3]) == 0 and torch.count_nonzero(v[4]) == 0
    inside = OL.bspline_basis(torch.linspace(-1, 0.999, 64).unsqueeze(1), g).sum(-1)
    assert torch.allclose(inside, torch.ones_like(inside), atol=1e-6)          # partition of unity on [-1,1)


def test_kan_and_kanfet_golden():
    p, d = load_golden("kan_2_10_2")
    layers = _layers(p, 2)
    assert max_rel(OL.kan_forward(d["x"], layers), d["y"]) < TOL
    traj = OS.rk4_classic(lambda t, h: OL.kan_forward(h, layers), d["x"], 0.0, 0.14, 10, trajectory=True)
    assert max_rel(traj, d["traj"]) < 10 * TOL
    p, d = load_golden("kanfet_2_10_2")
    layers = _layers(p, 2)
    st = OL.kanfet_states(layers)
    for i in range(3):
        assert max_rel(OL.kan_forward(d[f"x{i}"], layers, branch="ferro2d", states=st), d[f"y{i}"]) < TOL
    st = OL.kanfet_states(layers)
    traj = OS.rk4_classic(lambda t, h: OL.kan_forward(h, layers, branch="ferro2d", states=st), d["x0"], 0.0, 0.14, 10,
                          trajectory=True)
    assert max_rel(traj, d["traj"]) < 10 * TOL


def test_ferro_golden():
    p, d = load_golden("ferro2d_5x7")
    st = OL.Ferro2DState(p)
    assert max_rel(OL.ferro2d_forward(d["xa"], p, st), d["oa"]) < TOL
    assert max_rel(OL.ferro2d_forward(d["xb"], p, st), d["ob"]) < TOL
    p, d = load_golden("ferrofull_6x4x5")
    st = OL.FerroFullState(p)
    assert max_rel(OL.ferro_full_forward(d["x1"], p, st), d["o1"]) < TOL
    assert max_rel(OL.ferro_full_forward(d["x1b"], p, st), d["o1b"]) < TOL
    st.reset()
    assert max_rel(OL.ferro_full_forward(d["xa"], p, st), d["oa"]) < TOL
    assert max_rel(OL.ferro_full_forward(d["xb"], p, st), d["ob"]) < TOL
    p, d = load_golden("ferronoisy_6x4x5")
    st = OL.FerroFullState(p)
    for k in "abc":
        assert max_rel(OL.ferro_full_forward(d["x" + k], p, st, update_branch=True), d["o" + k]) < TOL


def test_hardgate_golden_gate_bits():
    p, d = load_golden("hardgate_8x6")
    st = OL.HardGateState(p)
    assert torch.equal(OL.hardgate_forward(d["xa"], p, st), d["oa"]) or max_rel(OL.hardgate_forward(d["xa"], p, OL.HardGateState(p)), d["oa"]) < TOL
    assert torch.equal(st.prev, d["prev_after_a"])
    ob = OL.hardgate_forward(d["xb"], p, st)
    assert max_rel(ob, d["ob"]) < TOL


def test_kuramoto_and_rk_loops_golden():
    p, d = load_golden("kuramoto_clf")
    feat = OL.kuramoto_forward(d["img"], p["osc.K"], p["osc.omega"])
    assert max_rel(feat, d["feat"]) < 5 * TOL
    head = {k[len("head."):]: v for k, v in p.items() if k.startswith("head.")}
    assert max_rel(OL.kanlinear_forward(feat, head), d["logits"]) < 5 * TOL
    p, d = load_golden("rk_loops_mlp")
    import torch.nn.functional as F
    f = lambda t, z: F.linear(torch.tanh(F.linear(z, p["0.weight"], p["0.bias"])), p["2.weight"], p["2.bias"])
    assert max_rel(OS.euler_fixed(f, d["z0"]), d["euler"]) < TOL
    assert max_rel(OS.rk2_midpoint(f, d["z0"]), d["rk2"]) < TOL
    assert max_rel(OS.rk4_classic(f, d["z0"]), d["rk4"]) < TOL
    assert max_rel(OS.rk4_substeps(f, d["z0"], torch.linspace(0, 1, 4), 3), d["rk4_sub"]) < TOL


def test_update_grid_and_regularization_loss_against_reference_vectors():
    """oracle.layers.update_grid / regularization_loss (efficientkan.py:184-237) against the vectors the live reference
    produced (oracle/make_golden_grid.py).  The knots are exact fp32 arithmetic on sorted data; the refitted
    coefficients come out of LAPACK's least squares, whose last bits depend on the thread count: 1e-5."""
    from conftest import load_golden
    from oracle import layers as OL
    p, g = load_golden("update_grid_6x4")
    assert torch.allclose(OL.regularization_loss(p), g["reg_default"], rtol=1e-6)
    assert torch.allclose(OL.regularization_loss(p, 0.5, 2.0, 0.1), g["reg_custom"], rtol=1e-6)
    knots, coef = OL.update_grid(g["x"], p, margin=0.01)
    assert torch.equal(knots, g["grid_after"])
    assert torch.allclose(coef, g["spline_weight_after"], rtol=1e-4, atol=1e-6)


def test_ferro_conv2d_against_reference_vectors():
    """oracle.layers.ferro_conv2d_forward (ferro_class.py:689-779, :840-944) against the vectors the live reference
    classes produced (oracle/make_golden_conv.py): stateless calls, true-hysteresis sequences, and the plain class's
    stateful failure for Cout > 1."""
    import pytest
    from conftest import load_golden, max_rel
    from oracle import layers as OL
    for tag, kw, me in (("plain_s1p1", dict(stride=1, padding=1), False),
                        ("plain_s2p0_c1", dict(stride=2, padding=0), False),
                        ("memeff_s1p1", dict(stride=1, padding=1), True)):
        p, d = load_golden("ferroconv_" + tag)
        for i in (0, 1):
            assert max_rel(OL.ferro_conv2d_forward(d[f"x{i}"], p, mem_efficient=me, **kw), d[f"stateless{i}"]) < 1e-6
        st = OL.FerroConvState()
        if "stateful0" in d:
            for i in range(3):
                got = OL.ferro_conv2d_forward(d[f"x{i}"], p, st, stateful=True, mem_efficient=me, **kw)
                assert max_rel(got, d[f"stateful{i}"]) < 1e-6, (tag, i)
        else:
            with pytest.raises(RuntimeError):
                OL.ferro_conv2d_forward(d["x0"], p, st, stateful=True, **kw)


def test_kan_rnn_encoder_against_reference_vectors():
    """oracle.layers.kan_rnn_cell / kan_rnn_encoder (train_kan_fet_ett.py:780-818) against the live reference's
    outputs (oracle/make_golden_rnn.py), including the recurrent regime H > F*nb."""
    from conftest import load_golden, max_rel
    from oracle import layers as OL
    for tag in ("script_7_10_64", "recurrent_3_4_20"):
        p, d = load_golden("kanrnn_" + tag)
        assert max_rel(OL.kan_rnn_encoder(d["x"], p), d["z0"]) < 1e-6
        cell = {k[len("rnn_cell."):]: v for k, v in p.items() if k.startswith("rnn_cell.")}
        assert max_rel(OL.kan_rnn_cell(d["x"][:, 0, :], d["h_prev"], cell), d["cell"]) < 1e-6
