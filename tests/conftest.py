"""Shared test plumbing.  ``-m "not gpu"`` runs here (no GPU): oracle vs golden vectors, host logic,
C-ABI symbol checks.  ``-m gpu`` runs on the B200 box: parity of the CUDA path (through the C ABI)
against the oracle and the golden vectors."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fet-ode_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    """tests/golden/<name>.npz -> (params dict without the 'p.' prefix, other arrays), as torch tensors."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    params, data = {}, {}
    for k in z.files:
        t = torch.from_numpy(np.array(z[k]))
        (params if k.startswith("p.") else data)[k[2:] if k.startswith("p.") else k] = t
    return params, data


def max_rel(a: torch.Tensor, b: torch.Tensor) -> float:
    """max|a-b| / max|b|  -- the max-norm relative error the north star's 1e-5 / 1e-4 bars use
    (SURVEY section 7 'Tolerance definition')."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="session")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
