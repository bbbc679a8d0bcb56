"""The product has no CPU path: every entry of the Python face refuses CPU tensors loudly instead of computing with
torch ops.  Runs without a GPU."""
import pytest
import torch

import fetode
from fetode import ecg, ops


def test_layers_refuse_cpu_inputs():
    layer = fetode.KANLinear(4, 3)
    with pytest.raises(RuntimeError, match="CUDA only"):
        layer(torch.randn(2, 4))
    basis = ecg.LogisticBasis(4, 3)
    with pytest.raises(RuntimeError):
        basis(torch.randn(2, 4))


def test_glue_ops_refuse_cpu_inputs():
    x = torch.randn(2, 8)
    with pytest.raises(RuntimeError, match="CUDA only"):
        ops.ln_tanh(x, torch.nn.LayerNorm(8), 1.0)
    with pytest.raises(RuntimeError, match="CUDA only"):
        ops.linear(x, torch.nn.Linear(8, 4))


def test_solver_and_optimiser_refuse_cpu_tensors():
    net = fetode.KAN([2, 4, 2])
    with pytest.raises(RuntimeError):
        fetode.odeint(lambda t, y: net(y), torch.ones(1, 2), torch.linspace(0, 1, 3), method="rk4")
    with pytest.raises(RuntimeError, match="no CPU path"):
        fetode.optim.Adam(net.parameters(), lr=1e-3)
