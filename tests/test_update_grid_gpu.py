"""GPU parity of KANLinear.update_grid / curve2coeff / regularization_loss (efficient_kan/efficientkan.py:133-143,
184-237; SURVEY section 8(b) drop-in surface, 8(f) rank 2) against vectors produced by the live reference
(oracle/make_golden_grid.py).  b_splines runs through fetode_bspline_basis; sort / quantile picks / lstsq are torch
ops on the device, as in the reference.  Tolerances: the knots are plain fp32 arithmetic on sorted data (1e-6); the
refitted coefficients come out of a least-squares solve whose LAPACK (reference, CPU) and cuSOLVER (here) drivers
differ, so they are compared through what they are for -- the layer output after the refit (2e-4 of the max-norm) --
and directly at 2e-3."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel

pytestmark = pytest.mark.gpu


def test_update_grid_and_regularization_match_reference(cuda):
    p, g = load_golden("update_grid_6x4")
    layer = fetode.KANLinear(6, 4)
    layer.load_state_dict(p, strict=False)
    layer = layer.to(cuda)
    probe = g["probe"].to(cuda)
    with torch.no_grad():
        assert max_rel(layer(probe), g["y_before"]) < 1e-5
        assert abs(float(layer.regularization_loss()) - float(g["reg_default"])) < 1e-5 * abs(float(g["reg_default"]))
        reg = layer.regularization_loss(regularize_activation=0.5, regularize_entropy=2.0, regularize_logistic_l1=0.1)
        assert abs(float(reg) - float(g["reg_custom"])) < 1e-5 * abs(float(g["reg_custom"]))
        layer.update_grid(g["x"].to(cuda), margin=0.01)
        assert max_rel(layer.grid, g["grid_after"]) < 1e-6
        assert max_rel(layer.spline_weight, g["spline_weight_after"]) < 2e-3
        assert max_rel(layer(probe), g["y_after"]) < 2e-4


def test_kan_forward_with_update_grid_flag(cuda):
    """KAN.forward(x, update_grid=True) (efficientkan.py:274-279) refits every layer's grid on the way through and must
    leave a network that still evaluates (packed images rebuilt from the new grid)."""
    torch.manual_seed(0)
    net = fetode.KAN([6, 5, 3]).to(cuda)
    x = torch.randn(64, 6, device=cuda) * 0.6
    with torch.no_grad():
        y0 = net(x)
        g0 = net.layers[0].grid.clone()
        y1 = net(x, update_grid=True)
        y2 = net(x)
    assert not torch.equal(net.layers[0].grid, g0)
    assert torch.isfinite(y1).all() and torch.isfinite(y2).all()
    assert max_rel(y2, y0) < 5e-2                       # the refit approximates the same function on the new knots
