"""GPU parity of the any-width KANLinear backward (csrc/kanlinear_bwd.cu) against oracle autograd (fp64)."""
import pytest
import torch

import fetode
from conftest import max_rel
from oracle import layers as OL

pytestmark = pytest.mark.gpu
GRAD_TOL = 2e-4


def _make(I, O, nb, branch, cuda, seed):
    gen = torch.Generator().manual_seed(seed)
    p = OL.make_kanlinear_params(I, O, gen, nb=nb, branch=branch)
    kw = dict(num_basis=nb) if branch != "none" else dict(enable_logistic_basis=False)
    layer = fetode.KANLinear(I, O, **kw)
    if branch == "ferro2d":
        layer.logistic_basis = fetode.TwoDimensionFerroelectricBasis(I, nb)
    layer.load_state_dict(p, strict=False)
    return p, layer.to(cuda), gen


@pytest.mark.parametrize("I,O,nb,branch", [(2, 10, 10, "logistic"), (10, 2, 10, "ferro2d"), (48, 24, 10, "logistic"),
                                           (64, 130, 12, "logistic"), (96, 64, 10, "ferro2d"), (128, 128, 0, "none"),
                                           (7, 3, 5, "ferro2d"), (100, 10, 8, "logistic")])
@pytest.mark.parametrize("B", [1, 200])
def test_gradients_match_oracle_autograd(cuda, I, O, nb, branch, B):
    p, layer, gen = _make(I, O, nb, branch, cuda, I * 13 + O)
    p64 = {k: v.double().requires_grad_(k != "grid") for k, v in p.items()}
    xs = [torch.randn(B, I, generator=gen) * 1.1, torch.randn(B, I, generator=gen) * 0.8]
    ws = [torch.randn(B, O, generator=gen, dtype=torch.float64) for _ in xs]
    st = OL.Ferro2DState({k.split(".", 1)[1]: v for k, v in p64.items() if k.startswith("logistic_basis.")}) \
        if branch == "ferro2d" else None
    x64 = [x.double().requires_grad_(True) for x in xs]
    loss64 = sum((OL.kanlinear_forward(x, p64, branch=branch, branch_state=st) * w).sum() for x, w in zip(x64, ws))
    loss64.backward()
    xg = [x.to(cuda).requires_grad_(True) for x in xs]
    loss = sum((layer(x) * w.float().to(cuda)).sum() for x, w in zip(xg, ws))
    loss.backward()
    assert abs(float(loss) - float(loss64)) < 1e-4 * max(1.0, abs(float(loss64)))
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < GRAD_TOL
    for name, t in layer.named_parameters():
        assert t.grad is not None, name
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name


def test_gradient_at_full_ecg_size_is_linear_in_upstream(cuda):
    """768->128 at B=4096 (the ECG kan1 shape): too slow for the oracle, so check linearity of the backward
    in the upstream gradient and batch-split additivity of the weight gradients."""
    p, layer, gen = _make(768, 128, 10, "logistic", cuda, 1)
    x = (torch.randn(4096, 768, generator=gen) * 0.9).to(cuda)
    g = torch.randn(4096, 128, generator=gen).to(cuda)

    def grads(xs, gs, scale):
        layer.zero_grad()
        xs = xs.clone().requires_grad_(True)
        (layer(xs) * gs * scale).sum().backward()
        return xs.grad, {n: t.grad.clone() for n, t in layer.named_parameters()}

    gx1, gp1 = grads(x, g, 1.0)
    gx2, gp2 = grads(x, g, 2.0)
    assert max_rel(gx2, 2 * gx1) < 1e-6
    for n in gp1:
        assert max_rel(gp2[n], 2 * gp1[n]) < 1e-5, n
    _, ga = grads(x[:2048], g[:2048], 1.0)
    _, gb = grads(x[2048:], g[2048:], 1.0)
    for n in gp1:
        assert max_rel(ga[n] + gb[n], gp1[n]) < 1e-4, n


def test_shared_fold_node_survives_partial_and_repeated_backward(cuda):
    """One fold node serves every call of a layer between two parameter updates and the folded gradient is unfolded
    once (ops.FoldFn).  A partial pass (autograd.grad w.r.t. the input only, which never reaches the fold node) must
    leave nothing behind, and a second pass over fresh graphs must not see the first pass's accumulator."""
    p, layer, gen = _make(48, 24, 10, "logistic", cuda, 5)
    xs = [(torch.randn(64, 48, generator=gen) * 0.9).to(cuda) for _ in range(3)]
    ws = [torch.randn(64, 24, generator=gen).to(cuda) for _ in range(3)]

    def loss_of(inputs):
        return sum((layer(x) * w).sum() for x, w in zip(inputs, ws))

    layer.zero_grad()
    loss_of(xs).backward()
    ref = {n: t.grad.clone() for n, t in layer.named_parameters()}
    # partial pass first: gradient w.r.t. the inputs only
    xg = [x.clone().requires_grad_(True) for x in xs]
    gx = torch.autograd.grad(loss_of(xg), xg)
    layer.zero_grad()
    xg2 = [x.clone().requires_grad_(True) for x in xs]
    loss_of(xg2).backward()
    for a, b in zip(gx, xg2):
        assert max_rel(a, b.grad) < 1e-6
    for n, t in layer.named_parameters():
        assert max_rel(t.grad, ref[n]) < 1e-5, n
    # per-call gradients add up to the shared-node gradient
    tot = {n: torch.zeros_like(t) for n, t in layer.named_parameters()}
    for x, w in zip(xs, ws):
        layer.zero_grad()
        (layer(x) * w).sum().backward()
        for n, t in layer.named_parameters():
            tot[n] += t.grad
    for n in tot:
        assert max_rel(tot[n], ref[n]) < 1e-5, n
    # frozen parameters: only the input gradient is produced
    for t in layer.parameters():
        t.requires_grad_(False)
    layer.zero_grad()
    xf = xs[0].clone().requires_grad_(True)
    (layer(xf) * ws[0]).sum().backward()
    assert max_rel(xf.grad, xg2[0].grad) < 1e-6
    assert all(t.grad is None for t in layer.parameters())


def test_deepcopy_after_training_step(cuda):
    """copy.deepcopy(model) (best-model snapshots in the training scripts) must work after forwards / backwards have
    populated the packed-image caches, and the copy must compute on its own parameters."""
    import copy
    p, layer, gen = _make(48, 24, 10, "logistic", cuda, 7)
    x = (torch.randn(32, 48, generator=gen) * 0.9).to(cuda)
    layer(x).sum().backward()
    twin = copy.deepcopy(layer)
    y0 = layer(x).detach()
    assert torch.equal(twin(x).detach(), y0)
    with torch.no_grad():
        twin.base_weight.mul_(2.0)
    assert not torch.equal(twin(x).detach(), y0)
    assert torch.equal(layer(x).detach(), y0)
    net = fetode.KANFET([2, 10, 2], grid_size=5).to(cuda)
    z = torch.rand(8, 2, device=cuda) + 0.5
    out = fetode.integrate_rk4(lambda t, h: net(h), z, 0.0, 0.1, 4)
    net2 = copy.deepcopy(net)
    net.reset_state(); net2.reset_state()
    a = fetode.integrate_rk4(lambda t, h: net(h), z, 0.0, 0.1, 4)
    b = fetode.integrate_rk4(lambda t, h: net2(h), z, 0.0, 0.1, 4)
    assert torch.equal(a, b)
