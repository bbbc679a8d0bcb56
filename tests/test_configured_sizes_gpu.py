"""Parity at the sizes BASELINE.json names, against the ORACLE (not against another kernel of this repo):
  * tcgen05 wide KANLinear forward + backward at B = 4096, 768 -> 128 (config 3's kan1 shape, split-K and every row tile);
  * the whole config-3 model KanFet_MLP_NODE(T=96, latent 64, nb 12, hidden 128), classic RK4 x 8: loss, EVERY parameter
    gradient and the hysteresis rows the step leaves behind, against fp64 oracle autograd -- this exercises the one fold
    node per layer accumulating over the 32 right-hand-side calls of a step (ops.FoldFn);
  * the KANFET([928,256]) -> ([256,256]) -> ([256,672]) stack of kan_diffusion_ett.py:421-425 (config 4), forward and
    gradients.
Tolerances: 1e-5 max-norm relative per evaluation (north star), 2e-4 / 5e-4 for gradients (fp32 accumulation over
4096 rows / 32 chained calls against fp64)."""
import pytest
import torch
import torch.nn.functional as F

import fetode
from conftest import max_rel
from fetode import ecg
from oracle import layers as OL
from oracle import models as OM

pytestmark = pytest.mark.gpu


def test_tc_forward_backward_full_ecg_size_vs_oracle(cuda):
    I, O, nb, B = 768, 128, 10, 4096
    gen = torch.Generator().manual_seed(41)
    p = OL.make_kanlinear_params(I, O, gen, nb=nb, branch="logistic")
    layer = fetode.KANLinear(I, O, num_basis=nb)
    layer.load_state_dict(p, strict=False)
    layer = layer.to(cuda)
    x = torch.randn(B, I, generator=gen) * 0.9
    w = torch.randn(B, O, generator=gen, dtype=torch.float64)
    with torch.no_grad():
        want32 = OL.kanlinear_forward(x, p, branch="logistic")               # the reference's own fp32 arithmetic
    p64 = {k: v.double().requires_grad_(k != "grid") for k, v in p.items()}
    x64 = x.double().requires_grad_(True)
    y64 = OL.kanlinear_forward(x64, p64, branch="logistic")
    (y64 * w).sum().backward()
    xg = x.to(cuda).requires_grad_(True)
    y = layer(xg)
    assert max_rel(y, want32) < 1e-5 and max_rel(y, y64) < 1e-5
    (y * w.float().to(cuda)).sum().backward()
    assert max_rel(xg.grad, x64.grad) < 2e-4
    for name, t in layer.named_parameters():
        assert max_rel(t.grad, p64[name].grad) < 2e-4, name


@pytest.mark.parametrize("B", [64, 257])
def test_config3_model_loss_and_all_gradients_vs_fp64_oracle(cuda, B):
    torch.manual_seed(0)
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    # a larger output gain than the script's 1e-3 initialisation, so that the ODE actually moves the state and the
    # gradients through the 32 right-hand-side calls are not vanishingly small
    with torch.no_grad():
        model.odefunc.out.weight.mul_(200.0)
        model.odefunc.log_alpha.fill_(0.5)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    P64 = OM.ecg_params(sd, torch.float64, requires_grad=True)
    P32 = OM.ecg_params(sd, torch.float32)
    gen = torch.Generator().manual_seed(B)
    x = torch.randn(B, 96, generator=gen)
    x = (x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)               # per-row z-score :44
    y = torch.randint(0, 2, (B,), generator=gen)
    st64, st32 = OM.EcgStates(P64), OM.EcgStates(P32)
    loss64 = F.cross_entropy(OM.kanfet_mlp_node_forward(x.double(), P64, st64), y)
    loss64.backward()
    with torch.no_grad():
        logits32 = OM.kanfet_mlp_node_forward(x, P32, st32)
    model = model.to(cuda)
    logits = model(x.to(cuda))
    loss = F.cross_entropy(logits, y.to(cuda))
    loss.backward()
    assert max_rel(logits, logits32) < 1e-4                                  # 8 RK4 steps: the trajectory bar
    assert abs(float(loss) - float(loss64)) < 1e-5 * max(1.0, abs(float(loss64)))
    flat64 = P64["flat"]
    checked = 0
    for name, t in model.named_parameters():
        g64 = flat64[name].grad
        if g64 is None:                                                       # hard-gate coef: unused by forward
            assert t.grad is None or float(t.grad.abs().max()) == 0.0, name
            continue
        assert t.grad is not None, name
        assert max_rel(t.grad, g64) < 5e-4, name
        checked += 1
    assert checked >= 25
    # hysteresis rows left behind by the step == the oracle's (last batch row of the last call of each basis)
    assert max_rel(model.odefunc.h_feat.basis._row, st32.rhs.prev[0, :, 0]) < 1e-4
    assert max_rel(model.cls_feat.basis._row, st32.cls.prev[0, :, 0]) < 1e-4


def test_config4_kanfet_stack_forward_and_gradients_vs_oracle(cuda):
    """kan_diffusion_ett.py:421-425: KANFET([928,256]) -> KANFET([256,256]) -> KANFET([256,672]), applied twice so the
    second call sees the hysteresis state of the first."""
    widths = [(928, 256), (256, 256), (256, 672)]
    B = 96
    gen = torch.Generator().manual_seed(7)
    ps = [OL.make_kanlinear_params(i, o, gen, nb=10, branch="ferro2d") for i, o in widths]
    nets = []
    for (i, o), p in zip(widths, ps):
        n = fetode.KANFET([i, o])
        missing, unexpected = n.load_state_dict({f"layers.0.{k}": v for k, v in p.items()}, strict=False)
        assert not unexpected
        nets.append(n.to(cuda))
    p64 = [{k: v.double().requires_grad_(k != "grid") for k, v in p.items()} for p in ps]
    st64 = [OL.kanfet_states([p])[0] for p in p64]
    st32 = [OL.kanfet_states([p])[0] for p in ps]
    xs = [torch.randn(B, 928, generator=gen) * 0.8, torch.randn(B, 928, generator=gen) * 0.8]
    ws = [torch.randn(B, 672, generator=gen, dtype=torch.float64) for _ in xs]

    def oracle(x, pp, sts):
        for p, s in zip(pp, sts):
            x = OL.kanlinear_forward(x, p, branch="ferro2d", branch_state=s)
        return x
    x64 = [x.double().requires_grad_(True) for x in xs]
    tot = 0.0
    for x, w in zip(x64, ws):
        tot = tot + (oracle(x, p64, st64) * w).sum()
    tot.backward()
    with torch.no_grad():
        want32 = [oracle(x, ps, st32) for x in xs]
    xg = [x.to(cuda).requires_grad_(True) for x in xs]
    outs = []
    for x in xg:
        h = x
        for n in nets:
            h = n(h)
        outs.append(h)
    for o, w32 in zip(outs, want32):
        assert torch.isfinite(o).all()
        assert max_rel(o, w32) < 3e-5                                        # three chained layers
    sum((o * w.float().to(cuda)).sum() for o, w in zip(outs, ws)).backward()
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < 5e-4
    for n, p in zip(nets, p64):
        for name, t in n.layers[0].named_parameters():
            assert t.grad is not None, name
            assert max_rel(t.grad, p[name].grad) < 5e-4, name
