"""Regression tests for host-side state rules that only show on a device run: the hysteresis state is OWNED by the
module (the reference stores ``x.detach().clone()``, ferro_class.py:409,584-590), dopri5 stops on a diverging
right-hand side like torchdiffeq does, and eager evaluation between graph replays sees the current weights."""
import copy

import pytest
import torch
import torch.nn.functional as F

import fetode
from conftest import max_rel
from fetode import ecg, mnist_kuramoto_kan as MK

pytestmark = pytest.mark.gpu


def test_state_does_not_alias_the_callers_input(cuda):
    torch.manual_seed(0)
    for mod, width in ((fetode.TwoDimensionFerroelectricBasis(3, 4), 3), (fetode.FerroelectricBasis(3, 2, 4), 3),
                       (fetode.KANFET([3, 5]), 3)):
        mod = mod.to(cuda)
        x = torch.randn(6, width, device=cuda)
        keep = x.clone()
        with torch.no_grad():
            mod(x)
        for m in mod.modules():
            if hasattr(m, "prev_x") and m.prev_x is not None:
                m.prev_x.zero_()                               # reset_stateful_ferro_buffers, train_ecg.py:371-378
        assert torch.equal(x, keep), type(mod).__name__
        with torch.no_grad():
            mod(x)
        x.add_(1.0)                                            # a later in-place edit must not move the state
        for m in mod.modules():
            r = getattr(m, "_rows", None)
            if r is not None:
                assert torch.equal(r, keep)


def test_state_dict_round_trip_after_forward(cuda):
    torch.manual_seed(0)
    net = fetode.KANFET([2, 10, 2]).to(cuda)
    x = torch.rand(9, 2, device=cuda)
    with torch.no_grad():
        net(x)
        best = copy.deepcopy(net.state_dict())
        y1 = net(x * 0.5)
        net.load_state_dict(best)                              # train_kan_fet_ett.py:353-358
        y2 = net(x * 0.5)
    assert torch.equal(y1, y2)


def test_dopri5_raises_on_non_finite_rhs(cuda):
    y0 = torch.ones(4, 2, device=cuda)
    t = torch.tensor([0.0, 1.0], dtype=torch.float64)

    def bad(tt, y):
        return y * float("nan")
    with pytest.raises(AssertionError):
        fetode.odeint(bad, y0, t, method="dopri5", rtol=1e-3, atol=1e-4)


def test_eager_eval_between_replays_sees_current_weights(cuda):
    torch.manual_seed(0)
    m1 = MK.KuramotoKANClassifier(kuramoto_steps=4, num_basis=8)
    m2 = copy.deepcopy(m1)
    m1, m2 = m1.to(cuda), m2.to(cuda)
    gen = torch.Generator().manual_seed(1)
    x = torch.rand(32, 1, 28, 28, generator=gen).to(cuda)
    y = torch.randint(0, 10, (32,), generator=gen).to(cuda)
    o1 = fetode.optim.AdamW(m1.parameters(), lr=1e-2)
    o2 = fetode.optim.AdamW(m2.parameters(), lr=1e-2, capturable=True)
    step = fetode.graphs.GraphedTrainStep(m2, o2, F.cross_entropy, x, y, warmup=2)
    with torch.no_grad():
        m2(x)                                                   # eager eval: caches the packed weights of this version
    for _ in range(4):
        step(x, y)
    for _ in range(6):
        o1.zero_grad()
        F.cross_entropy(m1(x), y).backward()
        o1.step()
    with torch.no_grad():
        a, b = m1(x), m2(x)                                     # must repack: parameters moved under the replays
    assert max_rel(b, a) < 1e-3


def test_gain_node_survives_two_backwards_without_a_step(cuda):
    torch.manual_seed(0)
    f = ecg.MLPKANODEFunc(latent_dim=8, num_basis=4, hidden=16).to(cuda)
    h = torch.randn(5, 8, device=cuda)
    for _ in range(2):                                          # gradient accumulation: no optimiser step in between
        f(None, h).sum().backward()
    assert f.scale.grad is not None and torch.isfinite(f.scale.grad).all()
