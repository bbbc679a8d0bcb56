"""GPU parity of the any-width KANLinear kernels (csrc/kanlinear.cu) through the C ABI."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from oracle import layers as OL

pytestmark = pytest.mark.gpu
RHS_TOL = 1e-5


def _layer_from(params, I, O, device, **kw):
    l = fetode.KANLinear(I, O, **kw)
    missing, unexpected = l.load_state_dict(params, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return l.to(device)


@pytest.mark.parametrize("I,O", [(2, 10), (10, 2), (48, 24)])
def test_kanlinear_matches_reference_golden(cuda, I, O):
    params, d = load_golden(f"kanlinear_{I}x{O}")
    layer = _layer_from(params, I, O, cuda)
    with torch.no_grad():
        got = layer(d["x"].to(cuda))
    assert max_rel(got, d["y"]) < RHS_TOL


def test_kanlinear_adapted_grid_no_branch(cuda):
    params, d = load_golden("kanlinear_nobranch_adapted")
    layer = _layer_from(params, 6, 5, cuda, enable_logistic_basis=False)
    with torch.no_grad():
        got = layer(d["x"].to(cuda))
    assert max_rel(got, d["y"]) < RHS_TOL


@pytest.mark.parametrize("I,O,nb,branch", [(768, 128, 10, "logistic"), (128, 128, 10, "none"),
                                           (64, 130, 12, "logistic"), (100, 10, 8, "logistic"),
                                           (96, 64, 10, "ferro2d"), (7, 3, 5, "ferro2d")])
@pytest.mark.parametrize("B", [1, 200])
def test_kanlinear_matches_oracle_wide(cuda, I, O, nb, branch, B):
    gen = torch.Generator().manual_seed(I * 7 + O)
    p = OL.make_kanlinear_params(I, O, gen, nb=nb, branch=branch)
    kw = dict(num_basis=nb) if branch != "none" else dict(enable_logistic_basis=False)
    layer = fetode.KANLinear(I, O, **kw)
    if branch == "ferro2d":
        layer.logistic_basis = fetode.TwoDimensionFerroelectricBasis(I, nb)
    layer.load_state_dict(p, strict=False)
    layer = layer.to(cuda)
    x = torch.randn(B, I, generator=gen) * 1.1
    st = OL.Ferro2DState({k.split(".", 1)[1]: v for k, v in p.items() if k.startswith("logistic_basis.")}) \
        if branch == "ferro2d" else None
    with torch.no_grad():
        for call in range(2):
            xi = x + 0.1 * call
            want = OL.kanlinear_forward(xi, p, branch=branch, branch_state=st)
            want64 = None if branch == "ferro2d" else OL.kanlinear_forward(
                xi.double(), OL.to_dtype(p, torch.float64), branch=branch)
            got = layer(xi.to(cuda))
            assert max_rel(got, want) < RHS_TOL, (call,)
            if want64 is not None:
                # the reference's own fp32 rounding is of the same order as ours
                assert max_rel(got, want64) < RHS_TOL


def test_leading_dims_and_shape_errors(cuda):
    layer = fetode.KANLinear(6, 4).to(cuda)
    x = torch.randn(3, 5, 6, device=cuda)
    with torch.no_grad():
        y = layer(x)
    assert y.shape == (3, 5, 4)
    with pytest.raises(AssertionError):
        layer(torch.randn(3, 7, device=cuda))
