"""GPU parity of the tcgen05 backward kernels of the wide KANLinear (csrc/kanlinear_tc_bwd.cu) against fp64 oracle
autograd, and against the CUDA-core backward at the full ECG size."""
import os

import pytest
import torch

import fetode
from conftest import max_rel
from oracle import layers as OL

pytestmark = pytest.mark.gpu
GRAD_TOL = 2e-4


def _layer(I, O, nb, branch, cuda, seed):
    gen = torch.Generator().manual_seed(seed)
    p = OL.make_kanlinear_params(I, O, gen, nb=nb, branch=branch)
    kw = dict(num_basis=nb) if branch != "none" else dict(enable_logistic_basis=False)
    layer = fetode.KANLinear(I, O, **kw)
    if branch == "ferro2d":
        layer.logistic_basis = fetode.TwoDimensionFerroelectricBasis(I, nb)
    layer.load_state_dict(p, strict=False)
    return p, layer.to(cuda), gen


def _tc_mode(mode):
    from fetode._lib import check, lib
    check(lib().fetode_set_tc_mode(int(mode)), "set_tc_mode")


@pytest.fixture
def force_tc():
    _tc_mode(1)
    yield
    _tc_mode(-1)


@pytest.mark.parametrize("I,O,nb,branch,B", [
    (768, 128, 10, "logistic", 300),
    (128, 128, 0, "none", 257),
    (100, 130, 10, "logistic", 77),       # ragged 8-input blocks, two output tiles, ragged row chunk
    (96, 64, 10, "ferro2d", 200),
    (256, 10, 8, "logistic", 130),
    (64, 256, 10, "logistic", 140),       # two 128-output slices: the input gradient accumulates across CTAs
    (2352, 128, 10, "logistic", 64),      # ETT encoder width at the script's batch size
])
def test_tc_gradients_match_oracle_autograd(cuda, force_tc, I, O, nb, branch, B):
    p, layer, gen = _layer(I, O, nb, branch, cuda, I * 3 + O)
    p64 = {k: v.double().requires_grad_(k != "grid") for k, v in p.items()}
    xs = [torch.randn(B, I, generator=gen) * 1.1, torch.randn(B, I, generator=gen) * 0.8]
    ws = [torch.randn(B, O, generator=gen, dtype=torch.float64) for _ in xs]
    st = OL.Ferro2DState({k.split(".", 1)[1]: v for k, v in p64.items() if k.startswith("logistic_basis.")}) \
        if branch == "ferro2d" else None
    x64 = [x.double().requires_grad_(True) for x in xs]
    sum((OL.kanlinear_forward(x, p64, branch=branch, branch_state=st) * w).sum() for x, w in zip(x64, ws)).backward()
    xg = [x.to(cuda).requires_grad_(True) for x in xs]
    sum((layer(x) * w.float().to(cuda)).sum() for x, w in zip(xg, ws)).backward()
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < GRAD_TOL
    for name, t in layer.named_parameters():
        assert t.grad is not None, name
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name


def test_tc_backward_equals_cuda_core_backward_at_ecg_size(cuda):
    p, layer, gen = _layer(768, 128, 10, "logistic", cuda, 5)
    x = (torch.randn(4096, 768, generator=gen) * 0.9).to(cuda)
    g = torch.randn(4096, 128, generator=gen).to(cuda)
    res = {}
    try:
        for mode in ("0", "1"):
            _tc_mode(int(mode))
            layer.zero_grad()
            xs = x.clone().requires_grad_(True)
            (layer(xs) * g).sum().backward()
            res[mode] = (xs.grad.clone(), {n: t.grad.clone() for n, t in layer.named_parameters()})
    finally:
        _tc_mode(-1)
    assert max_rel(res["1"][0], res["0"][0]) < 5e-5
    for n in res["0"][1]:
        assert max_rel(res["1"][1][n], res["0"][1][n]) < 5e-5, n
