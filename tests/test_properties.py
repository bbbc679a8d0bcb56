"""Property tests (SURVEY section 4, item v) of the cubic B-spline basis as the reference evaluates it
(efficientkan.py:117-131): non-negative, zero outside [t_0, t_last), a partition of unity on the inner span
[t_k, t_{-k-1}], at most four non-zero entries per input -- on hypothesis-generated non-uniform knot vectors, for the
oracle (CPU) and for the kernel behind KANLinear.b_splines (GPU)."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import layers as OL


def _knots(steps, lo):
    t = lo + np.cumsum(np.asarray(steps, dtype=np.float64))
    return torch.tensor(np.concatenate([[lo], t]), dtype=torch.float32).unsqueeze(0)      # [1, 12]


def _check(basis, x, knots):
    b = basis[:, 0, :].double()
    t = knots[0].double()
    xv = x[:, 0].double()
    assert (b >= -1e-6).all()
    outside = (xv < t[0]) | (xv >= t[-1])
    assert float(b[outside].abs().max()) == 0.0 if outside.any() else True
    inner = (xv >= t[3]) & (xv < t[8])
    if inner.any():
        assert torch.allclose(b[inner].sum(-1), torch.ones(int(inner.sum()), dtype=torch.float64), atol=2e-5)
    assert int((b.abs() > 0).sum(-1).max()) <= 4


@settings(max_examples=60, deadline=None)
@given(steps=st.lists(st.floats(0.05, 0.9), min_size=11, max_size=11), lo=st.floats(-3.0, 1.0),
       seed=st.integers(0, 2 ** 16))
def test_oracle_bspline_properties(steps, lo, seed):
    knots = _knots(steps, lo)
    gen = torch.Generator().manual_seed(seed)
    span = float(knots[0, -1] - knots[0, 0])
    x = (torch.rand(200, 1, generator=gen) * 1.4 - 0.2) * span + float(knots[0, 0])
    x[:12, 0] = knots[0]                                                   # knot-exact inputs
    _check(OL.bspline_basis(x, knots), x, knots)


@pytest.mark.gpu
def test_kernel_bspline_properties_and_oracle_equality(cuda):
    import fetode
    layer = fetode.KANLinear(3, 2).to(cuda)
    gen = torch.Generator().manual_seed(0)
    for trial in range(20):
        steps = (torch.rand(3, 11, generator=gen) * 0.85 + 0.05)
        lo = torch.rand(3, 1, generator=gen) * 4 - 3
        knots = torch.cat([lo, lo + torch.cumsum(steps, 1)], 1).float().contiguous()
        layer.grid.copy_(knots.to(cuda))
        span = knots[:, -1] - knots[:, 0]
        x = (torch.rand(300, 3, generator=gen) * 1.4 - 0.2) * span + knots[:, 0]
        x[:12] = knots.T
        got = layer.b_splines(x.to(cuda)).cpu()
        want = OL.bspline_basis(x, knots)
        assert float((got - want).abs().max()) < 2e-6
        for i in range(3):
            _check(got[:, i:i + 1, :], x[:, i:i + 1], knots[i:i + 1])
