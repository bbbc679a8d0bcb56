"""CPU check of the rule csrc/ferro.cu::hard_gate implements: torch's CPU ``sigmoid(z) > 0.5`` equals
fl32(1 / fl32(1 + fl32(exp(-z)))) > 0.5 with a correctly rounded exp (train_ecg_kan_fet_nn_ode.py:122-123).
The kernel evaluates exactly this expression (double exp rounded once to fp32, IEEE add and divide)."""
import numpy as np
import torch


def gate_rule(z32: np.ndarray) -> np.ndarray:
    e = np.exp(-z32.astype(np.float64)).astype(np.float32)
    s = (np.float32(1.0) + e).astype(np.float32)
    return (np.float32(1.0) / s).astype(np.float32) > np.float32(0.5)


def test_gate_rule_equals_torch_cpu_sigmoid():
    g = torch.Generator().manual_seed(0)
    z = torch.cat([
        torch.linspace(5e-8, 1.5e-7, 100001),
        (10 ** (torch.rand(200000, generator=g) * 14 - 13)) * torch.where(torch.rand(200000, generator=g) < .5, -1., 1.),
        torch.zeros(3), torch.tensor([float("inf"), -float("inf"), 88.0, -88.0, 104.0, -104.0]),
    ]).float()
    want = (torch.sigmoid(z) > 0.5).numpy()
    with np.errstate(over="ignore"):
        got = gate_rule(z.numpy())
    assert np.array_equal(got, want)
    # scalar (non-vectorised) torch path
    idx = list(range(0, 100001, 997))
    want_s = np.array([bool(torch.sigmoid(z[i:i + 1]) > 0.5) for i in idx])
    assert np.array_equal(got[idx], want_s)
    assert not gate_rule(np.zeros(1, np.float32))[0]            # dx == 0 -> down branch
