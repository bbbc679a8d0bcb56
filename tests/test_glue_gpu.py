"""GPU parity of the fused LayerNorm + tanh-bound prologue (csrc/glue.cu) against the two torch ops the reference
uses (train_ecg_kan_fet_nn_ode.py:459-461), forward and gradients, fp32.  Tolerance: 2e-6 of the max-norm forward
(fp32 rounding of the statistics), 2e-5 for gradients (row reductions in a different order)."""
import copy

import pytest
import torch

from fetode import ops
from conftest import max_rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,D,hb", [(4096, 64, 1.0), (1, 64, 1.0), (37, 16, 2.5), (300, 200, 1.0), (5, 256, 0.7),
                                    (129, 3, 1.0)])
def test_ln_tanh_matches_torch(cuda, B, D, hb):
    gen = torch.Generator().manual_seed(B * 7 + D)
    x = (torch.randn(B, D, generator=gen) * 2.0 + 0.3).to(cuda)
    ln = torch.nn.LayerNorm(D).to(cuda)
    with torch.no_grad():
        ln.weight.copy_(torch.rand(D, generator=gen) + 0.5)
        ln.bias.copy_(torch.randn(D, generator=gen) * 0.2)
    w = torch.randn(B, D, generator=gen).to(cuda)

    def run(fused):
        ln.zero_grad()
        xi = x.clone().requires_grad_(True)
        if fused:
            y = ops.ln_tanh(xi, ln, hb)
        else:
            y = hb * torch.tanh(ln(xi) / hb)
        (y * w).sum().backward()
        return y.detach(), xi.grad, ln.weight.grad.clone(), ln.bias.grad.clone()

    y0, gx0, gg0, gb0 = run(False)
    y1, gx1, gg1, gb1 = run(True)
    assert max_rel(y1, y0) < 2e-6
    assert max_rel(gx1, gx0) < 2e-5
    assert max_rel(gg1, gg0) < 2e-5
    assert max_rel(gb1, gb0) < 2e-5


def test_ln_tanh_falls_back_to_torch_ops_outside_the_kernel_cases(cuda):
    x = torch.randn(8, 300, device=cuda)                     # D > 256: the reference's two ops
    ln = torch.nn.LayerNorm(300).to(cuda)
    assert torch.equal(ops.ln_tanh(x, ln, 1.0), torch.tanh(ln(x)))


@pytest.mark.parametrize("B,K,O,bias", [(4096, 128, 64, True), (1, 128, 64, True), (77, 768, 64, True), (300, 10, 3, False),
                                        (5000, 33, 130, True)])
def test_linear_matches_torch_fp32(cuda, B, K, O, bias):
    """csrc/linear.cu against nn.Linear in fp32 (TF32 off, the torch default): forward 2e-6 of the max-norm,
    gradients 2e-5 (batch reductions in a different order)."""
    gen = torch.Generator().manual_seed(B + K + O)
    lin = torch.nn.Linear(K, O, bias=bias).to(cuda)
    x = torch.randn(B, K, generator=gen).to(cuda)
    w = torch.randn(B, O, generator=gen).to(cuda)

    def run(fused):
        lin.zero_grad()
        xi = x.clone().requires_grad_(True)
        y = ops.linear(xi, lin) if fused else lin(xi)
        (y * w).sum().backward()
        return y.detach(), xi.grad, lin.weight.grad.clone(), None if not bias else lin.bias.grad.clone()

    y0, gx0, gw0, gb0 = run(False)
    y1, gx1, gw1, gb1 = run(True)
    assert max_rel(y1, y0) < 2e-6
    assert max_rel(gx1, gx0) < 2e-6
    assert max_rel(gw1, gw0) < 2e-5
    if bias:
        assert max_rel(gb1, gb0) < 2e-5


@pytest.mark.parametrize("B", [3, 26, 33, 300])
def test_shared_gradient_accumulators_over_repeated_calls(cuda, B):
    """A module applied several times in one graph (once per Runge-Kutta stage) routes its parameter gradients through
    ONE accumulator per backward pass (ops.ParamGroup): every call must ADD -- also for batches so small that the
    weight-gradient GEMM has a single K split (regression: a 26-row shard overwrote the accumulator)."""
    from fetode import ops
    torch.manual_seed(B)
    lin = torch.nn.Linear(24, 10).to(cuda)
    ln = torch.nn.LayerNorm(24).to(cuda)
    ref_lin, ref_ln = copy.deepcopy(lin), copy.deepcopy(ln)
    xs = [torch.randn(B, 24, device=cuda, requires_grad=True) for _ in range(3)]
    xr = [x.detach().clone().requires_grad_(True) for x in xs]
    out = sum(ops.linear(ops.ln_tanh(x, ln, 1.5), lin).square().sum() for x in xs)
    out.backward()
    want = sum(ref_lin(1.5 * torch.tanh(ref_ln(x) / 1.5)).square().sum() for x in xr)
    want.backward()
    for a, b in zip(xs, xr):
        assert max_rel(a.grad, b.grad) < 2e-4
    for m, r in ((lin, ref_lin), (ln, ref_ln)):
        for (n, p), q in zip(m.named_parameters(), r.parameters()):
            assert max_rel(p.grad, q.grad) < 2e-4, n
    # a second pass without an optimiser step accumulates on top, like torch
    out2 = sum(ops.linear(ops.ln_tanh(x, ln, 1.5), lin).square().sum() for x in xs)
    out2.backward()
    want2 = sum(ref_lin(1.5 * torch.tanh(ref_ln(x) / 1.5)).square().sum() for x in xr)
    want2.backward()
    for (n, p), q in zip(lin.named_parameters(), ref_lin.parameters()):
        assert max_rel(p.grad, q.grad) < 2e-4, n


@pytest.mark.parametrize("B,K,O", [(1, 128, 64), (26, 128, 64), (300, 40, 7), (4096, 128, 64)])
def test_fused_head_matches_the_three_torch_steps(cuda, B, K, O):
    """ops.head = gain * lin(silu(z)) (train_ecg_kan_fet_nn_ode.py:471-480) against torch in fp64, called THREE times in
    one graph with a shared gain node (as once per Runge-Kutta stage): value, input gradient and the accumulated
    gradients of weight, bias and of the two scalars behind the gain."""
    import torch.nn.functional as F
    gen = torch.Generator().manual_seed(B + K)
    lin = torch.nn.Linear(K, O).to(cuda)
    scale = torch.nn.Parameter(torch.tensor(1.3, device=cuda))
    log_alpha = torch.nn.Parameter(torch.tensor(-1.0, device=cuda))
    zs = [(torch.randn(B, K, generator=gen) * 2).to(cuda) for _ in range(3)]
    ws = [torch.randn(B, O, generator=gen).to(cuda) for _ in range(3)]
    act = torch.nn.SiLU()

    zi = [z.clone().requires_grad_(True) for z in zs]
    gain = scale * F.softplus(log_alpha)
    outs = [ops.head(z, act, lin, gain) for z in zi]
    sum((o * w).sum() for o, w in zip(outs, ws)).backward()
    got = ([o.detach() for o in outs], [z.grad for z in zi], lin.weight.grad.clone(), lin.bias.grad.clone(),
           scale.grad.clone(), log_alpha.grad.clone())

    W, b = lin.weight.detach().double().requires_grad_(True), lin.bias.detach().double().requires_grad_(True)
    s64, a64 = scale.detach().double().requires_grad_(True), log_alpha.detach().double().requires_grad_(True)
    z64 = [z.double().requires_grad_(True) for z in zs]
    g64 = s64 * F.softplus(a64)
    o64 = [g64 * F.linear(F.silu(z), W, b) for z in z64]
    sum((o * w.double()).sum() for o, w in zip(o64, ws)).backward()
    for i in range(3):
        assert max_rel(got[0][i], o64[i]) < 2e-6, i
        assert max_rel(got[1][i], z64[i].grad) < 5e-6, i
    assert max_rel(got[2], W.grad) < 2e-5
    assert max_rel(got[3], b.grad) < 2e-5
    assert max_rel(got[4], s64.grad) < 2e-5
    assert max_rel(got[5], a64.grad) < 2e-5
    with torch.no_grad():                                    # inference path: no saved buffers, same values
        assert max_rel(ops.head(zs[0], act, lin, scale * F.softplus(log_alpha)), o64[0]) < 2e-6
