"""GPU parity of the tcgen05 wide-KANLinear forward (csrc/kanlinear_tc.cu) through the C ABI: against the oracle in
fp32 and fp64 (1e-5 max-norm relative, the north star's per-RHS bar -- this is what the 3xTF32 split is for) and
against the CUDA-core kernel on the same inputs."""
import os

import pytest
import torch

import fetode
from conftest import max_rel
from oracle import layers as OL

pytestmark = pytest.mark.gpu
RHS_TOL = 1e-5


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
    (768, 128, 10, "logistic", 300),      # ECG kan1 shape, ragged last row tile
    (128, 128, 10, "logistic", 128),      # ECG kan2
    (128, 128, 0, "none", 257),           # plain KAN
    (100, 130, 10, "logistic", 77),       # ragged input units and two ragged output tiles
    (96, 64, 10, "ferro2d", 200),         # KANFET branch with hysteresis state
    (256, 10, 8, "logistic", 64),         # MNIST-style head (num_basis 8, N padded to 16)
    (2352, 128, 10, "logistic", 64),      # ETT encoder width (K = 44 688)
])
def test_tc_forward_matches_oracle(cuda, force_tc, I, O, nb, branch, B):
    p, layer, gen = _layer(I, O, nb, branch, cuda, I + O)
    st = OL.Ferro2DState({k.split(".", 1)[1]: v for k, v in p.items() if k.startswith("logistic_basis.")}) \
        if branch == "ferro2d" else None
    x = torch.randn(B, I, generator=gen) * 1.1
    with torch.no_grad():
        for call in range(2):
            xi = x + 0.1 * call
            want = OL.kanlinear_forward(xi, p, branch=branch, branch_state=st)
            got = layer(xi.to(cuda))
            assert torch.isfinite(got).all()
            assert max_rel(got, want) < RHS_TOL, call
            if branch != "ferro2d":
                want64 = OL.kanlinear_forward(xi.double(), OL.to_dtype(p, torch.float64), branch=branch)
                assert max_rel(got, want64) < RHS_TOL


def test_tc_equals_cuda_core_path_at_full_ecg_size(cuda):
    """B = 4096, 768 -> 128 (BASELINE configs[2] kan1): the two kernels agree to fp32 rounding, split-K included."""
    p, layer, gen = _layer(768, 128, 10, "logistic", cuda, 3)
    x = (torch.randn(4096, 768, generator=gen) * 0.9).to(cuda)
    try:
        with torch.no_grad():
            _tc_mode(0)
            ref = layer(x)
            _tc_mode(1)
            got = layer(x)
    finally:
        _tc_mode(-1)
    assert max_rel(got, ref) < 1e-5          # each path is within ~5e-6 of an fp64 evaluation
    # linearity in the rows: any row subset gives the same rows (tiles are independent)
    with torch.no_grad():
        sub = layer(x[1000:1300].contiguous())
    assert max_rel(sub, got[1000:1300]) < 1e-5   # a different split-K factor: rounding only
