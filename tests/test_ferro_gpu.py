"""GPU parity of the stand-alone ferroelectric neurons (csrc/ferro.cu) and of the ECG right-hand sides built
from them, through the C ABI: golden vectors generated from the live reference (oracle/make_golden.py), the
oracle on random inputs, gradients against oracle autograd, and bit-identical hard-gate branch selection."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from fetode import ecg
from oracle import layers as OL

pytestmark = pytest.mark.gpu
RHS_TOL = 1e-5
GRAD_TOL = 2e-4


def _load(mod, params, device):
    missing, unexpected = mod.load_state_dict(params, strict=False)
    assert not unexpected, unexpected
    assert all(k.split(".")[-1] in ("prev_x", "branch_sign", "branch_state", "neighbor_kernel") for k in missing), missing
    return mod.to(device)


def _no_state(params):
    return {k: v for k, v in params.items() if k.split(".")[-1] not in ("prev_x", "branch_sign", "branch_state")}


def test_two_dimension_golden(cuda):
    p, d = load_golden("ferro2d_5x7")
    m = _load(fetode.TwoDimensionFerroelectricBasis(5, 7), _no_state(p), cuda)
    with torch.no_grad():
        assert max_rel(m(d["xa"].to(cuda)), d["oa"]) < RHS_TOL
        assert max_rel(m(d["xb"].to(cuda)), d["ob"]) < RHS_TOL
        assert m.prev_x.shape == (9, 5, 7)
        assert torch.equal(m.prev_x[:, :, 0].cpu(), d["xb"])


def test_full_golden_first_call_rules(cuda):
    p, d = load_golden("ferrofull_6x4x5")
    m = _load(fetode.FerroelectricBasis(6, 4, 5), _no_state(p), cuda)
    with torch.no_grad():
        assert max_rel(m(d["x1"].to(cuda)), d["o1"]) < RHS_TOL          # B == 1: prev = 0
        assert max_rel(m(d["x1b"].to(cuda)), d["o1b"]) < RHS_TOL
        m.reset_state()
        assert max_rel(m(d["xa"].to(cuda)), d["oa"]) < RHS_TOL          # B != stored rows: prev = x
        assert max_rel(m(d["xb"].to(cuda)), d["ob"]) < RHS_TOL
        assert m.prev_x.shape == (7, 6, 4, 5)


def test_noisy_golden_branch_state(cuda):
    p, d = load_golden("ferronoisy_6x4x5")
    m = _load(fetode.NoisyFerroelectricBasis(6, 4, 5, noise_std=0.0), _no_state(p), cuda)
    with torch.no_grad():
        for k in "abc":
            assert max_rel(m(d["x" + k].to(cuda)), d["o" + k]) < RHS_TOL, k
    assert m.branch_sign.shape == (7, 6, 4, 5)


def test_return_activations_and_batched_variant(cuda):
    gen = torch.Generator().manual_seed(5)
    p = OL.make_ferro_params((6, 4, 5), gen)
    m = _load(fetode.FerroelectricBasis(6, 4, 5), p, cuda)
    x = torch.randn(1, 6, generator=gen)
    st = OL.FerroFullState(p)
    want, wb = OL.ferro_full_forward(x, p, st, return_basis=True)
    out, basis, coef = m(x.to(cuda), return_activations=True)
    assert max_rel(out, want) < RHS_TOL and max_rel(basis, wb) < RHS_TOL and torch.equal(coef.cpu(), p["coef"])
    # per-sample loop variant: row b sees row b-1 as its previous input (ferro_class.py:107-214)
    mb = _load(fetode.BatchedFerroelectricBasis(6, 4, 5), p, cuda)
    xs = torch.randn(9, 6, generator=gen)
    st = OL.FerroFullState(p)
    want = torch.cat([OL.ferro_full_forward(xs[b:b + 1], p, st) for b in range(9)], 0)
    with torch.no_grad():
        assert max_rel(mb(xs.to(cuda)), want) < RHS_TOL
        want2 = torch.cat([OL.ferro_full_forward(xs[b:b + 1] * 0.5, p, st) for b in range(9)], 0)
        assert max_rel(mb((xs * 0.5).to(cuda)), want2) < RHS_TOL


def test_hardgate_golden_and_gate_bits(cuda):
    p, d = load_golden("hardgate_8x6")
    m = _load(ecg.LogisticBasis(8, 6), _no_state(p), cuda)
    st = OL.HardGateState(_no_state(p))
    with torch.no_grad():
        oa = m(d["xa"].to(cuda))
        OL.hardgate_forward(d["xa"], _no_state(p), st)
        assert max_rel(oa, d["oa"]) < RHS_TOL
        assert torch.equal(m.branch_state.cpu().float(), st.gate)
        assert torch.equal(m.prev_x.cpu(), d["prev_after_a"])
        ob = m(d["xb"].to(cuda))                                         # dx sweep around the rounding threshold
        OL.hardgate_forward(d["xb"], _no_state(p), st)
        assert torch.equal(m.branch_state.cpu().float(), st.gate)
        assert max_rel(ob, d["ob"]) < RHS_TOL


def test_hardgate_bits_random_sweep(cuda):
    """Branch selection is bit-identical to torch CPU for dx over 12 decades, both signs, and dx == 0."""
    gen = torch.Generator().manual_seed(11)
    I, K, B = 64, 12, 4096
    p = OL.make_ferro_params((I, K), gen)
    m = _load(ecg.LogisticBasis(I, K), p, cuda)
    st = OL.HardGateState(p)
    base = torch.randn(1, I, generator=gen)
    with torch.no_grad():
        m(base.to(cuda))
        OL.hardgate_forward(base, p, st)
        mag = 10.0 ** (torch.rand(B, I, generator=gen) * 12 - 12)
        sign = torch.where(torch.rand(B, I, generator=gen) < 0.5, -1.0, 1.0)
        x = base + mag * sign
        x[::7] = base                                                     # dx == 0 rows
        got = m(x.to(cuda))
        want = OL.hardgate_forward(x, p, st)
    assert torch.equal(m.branch_state.cpu().float(), st.gate)
    assert max_rel(got, want) < RHS_TOL
    assert torch.equal(m.prev_x[0, :, 0].cpu(), x[-1])


@pytest.mark.parametrize("variant", ["full", "full_br", "two_d", "hardgate", "logistic"])
def test_gradients_match_oracle_autograd(cuda, variant):
    gen = torch.Generator().manual_seed(3)
    B, I, O, K = 37, 6, 5, 4
    xa = torch.randn(B, I, generator=gen) * 1.5
    xb = xa + 0.3 * torch.randn(B, I, generator=gen)
    if variant in ("full", "full_br"):
        p = OL.make_ferro_params((I, O, K), gen)
        cls = fetode.NoisyFerroelectricBasis if variant == "full_br" else fetode.FerroelectricBasis
        m = _load(cls(I, O, K, noise_std=0.0), p, cuda)
    elif variant == "two_d":
        p = OL.make_ferro_params((I, K), gen)
        m = _load(fetode.TwoDimensionFerroelectricBasis(I, K), p, cuda)
    elif variant == "hardgate":
        p = OL.make_ferro_params((I, K), gen)
        m = _load(ecg.LogisticBasis(I, K), p, cuda)
    else:
        p = {"a": torch.randn(I, K, generator=gen), "b": torch.randn(I, K, generator=gen)}
        m = _load(fetode.LogisticBasis(I, K), p, cuda)
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}

    def oracle_calls(xs):
        outs = []
        if variant in ("full", "full_br"):
            st = OL.FerroFullState(p64)
            for x in xs:
                outs.append(OL.ferro_full_forward(x, p64, st, update_branch=variant == "full_br"))
        elif variant == "two_d":
            st = OL.Ferro2DState(p64)
            for x in xs:
                outs.append(OL.ferro2d_forward(x, p64, st))
        elif variant == "hardgate":
            st = OL.HardGateState(p64)
            for x in xs:
                outs.append(OL.hardgate_forward(x, p64, st))
        else:
            for x in xs:
                outs.append(OL.logistic_basis(x, p64["a"], p64["b"]))
        return outs

    xs64 = [xa.double().requires_grad_(True), xb.double().requires_grad_(True)]
    outs64 = oracle_calls(xs64)
    w = [torch.randn(o.shape, generator=gen, dtype=torch.float64) for o in outs64]
    sum((o * wi).sum() for o, wi in zip(outs64, w)).backward()

    xs = [xa.to(cuda).requires_grad_(True), xb.to(cuda).requires_grad_(True)]
    outs = [m(x) for x in xs]
    for o, o64 in zip(outs, outs64):
        assert max_rel(o, o64) < RHS_TOL
    sum((o * wi.float().to(cuda)).sum() for o, wi in zip(outs, w)).backward()
    for x, x64 in zip(xs, xs64):
        assert max_rel(x.grad, x64.grad) < GRAD_TOL
    for name, t in m.named_parameters():
        if p64[name].grad is None:                       # e.g. hard-gate coef is unused
            assert t.grad is None or float(t.grad.abs().max()) == 0.0
            continue
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name


def test_kanfet_ode_func_golden(cuda):
    p, d = load_golden("kanfetodefunc_6_9_4")
    m = _load(ecg.KANFetODEFunc(6, 9, 4), _no_state(p), cuda)
    with torch.no_grad():
        h = d["h"].to(cuda)
        d1 = m(torch.tensor(0.0), h)
        assert max_rel(d1, d["d1"]) < RHS_TOL
        d2 = m(torch.tensor(0.1), h + 0.1 * d1)
        assert max_rel(d2, d["d2"]) < RHS_TOL


def test_mlp_kan_rhs_golden(cuda):
    p, d = load_golden("mlpkanrhs_16_4_24")
    m = _load(ecg.MLPKANODEFunc(16, 4, 24), _no_state(p), cuda)
    with torch.no_grad():
        h = d["h"].to(cuda)
        d1 = m(torch.tensor(0.0), h)
        assert max_rel(d1, d["d1"]) < RHS_TOL
        d2 = m(torch.tensor(0.1), h + 0.1 * d1)
        assert max_rel(d2, d["d2"]) < RHS_TOL


def test_empty_and_shape_errors(cuda):
    m = fetode.FerroelectricBasis(6, 4, 5).to(cuda)
    assert m(torch.empty(0, 6, device=cuda)).shape == (0, 4)
    hg = ecg.LogisticBasis(8, 6).to(cuda)
    with pytest.raises(ValueError):
        hg(torch.randn(3, 7, device=cuda))


@pytest.mark.parametrize("B", [1, 37])
def test_kanfet_ode_func_fused_glue_values_and_gradients(cuda, B):
    """KANFetODEFunc (train_ecg.py:986-1013) with its glue fused into the two neuron launches: two chained calls
    (the second sees the hysteresis state of the first) against the oracle in fp32 (values) and fp64 autograd
    (input and every parameter gradient); an out-of-range output exercises the clamp's zero gradient."""
    gen = torch.Generator().manual_seed(100 + B)
    p1 = OL.make_ferro_params((6, 9, 4), gen)
    p2 = OL.make_ferro_params((9, 6, 4), gen)
    p2["coef"] = p2["coef"] * 25.0                                      # push some outputs beyond +-50
    m = ecg.KANFetODEFunc(6, 9, 4, h_bound=1.5)
    m.fc1.load_state_dict(p1, strict=False)
    m.fc2.load_state_dict(p2, strict=False)
    m = m.to(cuda)
    assert m._fusable()
    hs = [torch.randn(B, 6, generator=gen) * 2.0, torch.randn(B, 6, generator=gen) * 2.0]
    ws = [torch.randn(B, 6, generator=gen, dtype=torch.float64) for _ in hs]

    def run_oracle(dtype):
        q1 = {k: v.to(dtype).clone().requires_grad_(dtype == torch.float64) for k, v in p1.items()}
        q2 = {k: v.to(dtype).clone().requires_grad_(dtype == torch.float64) for k, v in p2.items()}
        s1, s2 = OL.FerroFullState(q1), OL.FerroFullState(q2)
        xs = [h.to(dtype).clone().requires_grad_(dtype == torch.float64) for h in hs]
        outs = [OL.kanfet_ode_rhs(x, q1, q2, s1, s2, h_bound=1.5) for x in xs]
        return q1, q2, xs, outs
    with torch.no_grad():
        _, _, _, want32 = run_oracle(torch.float32)
    q1, q2, x64, out64 = run_oracle(torch.float64)
    sum((o * w).sum() for o, w in zip(out64, ws)).backward()
    xg = [h.to(cuda).requires_grad_(True) for h in hs]
    outs = [m(None, x) for x in xg]
    for o, w32 in zip(outs, want32):
        assert max_rel(o, w32) < RHS_TOL
    assert float(torch.cat([o.detach().abs().flatten() for o in outs]).max()) == 50.0     # the clamp is active
    sum((o * w.float().to(cuda)).sum() for o, w in zip(outs, ws)).backward()
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < GRAD_TOL
    for mod, q in ((m.fc1, q1), (m.fc2, q2)):
        for name, t in mod.named_parameters():
            assert max_rel(t.grad, q[name].grad) < GRAD_TOL, name
    # state left behind: fc1 remembers the BOUNDED input, fc2 remembers tanh(fc1 out)
    assert max_rel(m.fc1._rows, 1.5 * torch.tanh(hs[1] / 1.5)) < 1e-6


@pytest.mark.parametrize("B,I,O,K", [(1, 5, 3, 4), (9, 5, 3, 4), (64, 64, 16, 12)])
def test_original_batched_scan_values_and_recurrent_gradients(cuda, B, I, O, K):
    """OriginalBatchedFerroelectricBasis (ferro_class.py:9-104) as one in-kernel scan over the batch: two consecutive
    calls (row and branch state carried) against the oracle, and the gradients -- which flow through the branch
    momentum recurrence across samples -- against fp64 oracle autograd."""
    gen = torch.Generator().manual_seed(B * 7 + I)
    p = OL.make_ferro_params((I, O, K), gen)
    m = fetode.OriginalBatchedFerroelectricBasis(I, O, K)
    m.load_state_dict(p, strict=False)
    m = m.to(cuda)
    xs = [torch.randn(B, I, generator=gen) * 1.5, torch.randn(B, I, generator=gen) * 1.5]
    st32 = OL.FerroSeqState(p)
    with torch.no_grad():
        want = [OL.ferro_seq_forward(x, p, st32) for x in xs]
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    st64 = OL.FerroSeqState(p64)
    x64 = [x.double().requires_grad_(True) for x in xs]
    ws = [torch.randn(B, O, generator=gen, dtype=torch.float64) for _ in xs]
    sum((OL.ferro_seq_forward(x, p64, st64) * w).sum() for x, w in zip(x64, ws)).backward()
    xg = [x.to(cuda).requires_grad_(True) for x in xs]
    outs = [m(x) for x in xg]
    for o, w32 in zip(outs, want):
        assert max_rel(o, w32) < RHS_TOL
    sum((o * w.float().to(cuda)).sum() for o, w in zip(outs, ws)).backward()
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < GRAD_TOL
    for name, t in m.named_parameters():
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name
    assert m.prev_x.shape == (1, I, O, K) and m.branch_sign.shape == (1, I, O, K)
    assert max_rel(m.branch_sign, st32.br) < 1e-5
    m.reset_state()
    assert float(m.prev_x.abs().max()) == 0.0 and float((m.branch_sign - 1).abs().max()) == 0.0
