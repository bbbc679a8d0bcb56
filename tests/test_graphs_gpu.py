"""Whole-step CUDA-graph replay (fetode/graphs.py) must train exactly like the eager loop it captured: same
parameters after the same number of steps, same hysteresis state carried from step to step
(train_ecg_kan_fet_nn_ode.py:1100-1112 loop body; hard-gate state :131-132)."""
import copy

import pytest
import torch
import torch.nn.functional as F

import fetode
from fetode import ecg, mnist_kuramoto_kan as MK
from conftest import max_rel

pytestmark = pytest.mark.gpu


def _eager_steps(model, opt, x, y, n, max_norm):
    losses = []
    for _ in range(n):
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        if max_norm is not None:
            opt.clip_grad_norm_(max_norm)
        opt.step()
        losses.append(float(loss))
    return losses


def test_graphed_ecg_step_matches_eager(cuda):
    torch.manual_seed(0)
    m1 = ecg.KanFet_MLP_NODE(T=24, num_classes=2, latent_dim=16, num_basis=6, ode_hidden=32, dropout=0.0)
    m1.ode_method = "rk4_classic"
    m2 = copy.deepcopy(m1)
    m1, m2 = m1.to(cuda), m2.to(cuda)
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(48, 24, generator=gen).to(cuda)
    y = torch.randint(0, 2, (48,), generator=gen).to(cuda)
    o1 = fetode.optim.AdamW(m1.parameters(), lr=1e-3)
    o2 = fetode.optim.AdamW(m2.parameters(), lr=1e-3, capturable=True)
    n = 7
    l1 = _eager_steps(m1, o1, x, y, n, 1.0)
    step = fetode.graphs.GraphedTrainStep(m2, o2, F.cross_entropy, x, y, max_norm=1.0, warmup=3)
    l2 = [float(step(x, y)) for _ in range(n - 3)]
    assert o2.step_count == n and int(o2._dev_state[0]) == n
    for a, b in zip(l1[3:], l2):
        assert abs(a - b) < 2e-5 * max(1.0, abs(a))
    for (k, p), q in zip(m1.named_parameters(), m2.parameters()):
        assert max_rel(q.detach(), p.detach()) < 5e-4, k
    # hysteresis state after the last step: the last row of the last call, identical in both runs
    for a, b in zip([m for m in m1.modules() if isinstance(m, ecg.LogisticBasis)],
                    [m for m in m2.modules() if isinstance(m, ecg.LogisticBasis)]):
        assert max_rel(b._row, a._row) < 5e-4
        assert torch.equal(b.prev_x[0, :, 0], b._row)
    # new data goes through the captured input buffers
    x2 = torch.randn(48, 24, generator=gen).to(cuda)
    la = float(step(x2, y))
    lb = _eager_steps(m1, o1, x2, y, 1, 1.0)[0]
    assert abs(la - lb) < 1e-4 * max(1.0, abs(lb))


def test_graphed_mnist_step_matches_eager(cuda):
    torch.manual_seed(0)
    m1 = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8)
    m2 = copy.deepcopy(m1)
    m1, m2 = m1.to(cuda), m2.to(cuda)
    gen = torch.Generator().manual_seed(2)
    x = torch.rand(32, 1, 28, 28, generator=gen).to(cuda)
    y = torch.randint(0, 10, (32,), generator=gen).to(cuda)
    o1 = fetode.optim.AdamW(m1.parameters(), lr=1e-3)
    o2 = fetode.optim.AdamW(m2.parameters(), lr=1e-3, capturable=True)
    l1 = _eager_steps(m1, o1, x, y, 6, None)
    step = fetode.graphs.GraphedTrainStep(m2, o2, F.cross_entropy, x, y, warmup=2)
    l2 = [float(step(x, y)) for _ in range(4)]
    for a, b in zip(l1[2:], l2):
        assert abs(a - b) < 2e-5 * max(1.0, abs(a))
    for (k, p), q in zip(m1.named_parameters(), m2.parameters()):
        assert max_rel(q.detach(), p.detach()) < 5e-4, k


def test_graphed_lv_training_matches_eager(cuda):
    """train_kanfet_node_predprey.py:246-262 loop body (reset_state, fused RK4 solve, MSE, Adam) replayed from a graph."""
    torch.manual_seed(0)
    m1 = fetode.KANFET([2, 10, 2], grid_size=5)
    m2 = copy.deepcopy(m1)
    m1, m2 = m1.to(cuda), m2.to(cuda)
    n = 34
    t = torch.linspace(0, 3.5, n + 1, dtype=torch.float64)                # host grid: no read-back inside the capture
    gen = torch.Generator().manual_seed(1)
    x0 = (torch.rand(16, 2, generator=gen) * 3.0 + 0.5).to(cuda)
    tgt = torch.randn(n + 1, 16, 2, generator=gen).to(cuda)

    def fwd(m, x):
        m.reset_state()
        return fetode.odeint(lambda tt, X: m(X), x, t, method="rk4_classic")

    def mse(pred, y):
        return ((pred - y) ** 2).mean()
    o1 = fetode.optim.Adam(m1.parameters(), lr=2e-3)
    l1 = []
    for _ in range(6):
        o1.zero_grad()
        loss = mse(fwd(m1, x0), tgt)
        loss.backward()
        o1.step()
        l1.append(float(loss.detach()))
    o2 = fetode.optim.Adam(m2.parameters(), lr=2e-3, capturable=True)
    step = fetode.graphs.GraphedTrainStep(m2, o2, mse, x0, tgt, warmup=3, forward=fwd)
    l2 = [float(step(x0, tgt)) for _ in range(3)]
    for a, b in zip(l1[3:], l2):
        assert abs(a - b) < 1e-5 * max(1.0, abs(a))
    for (k, p), q in zip(m1.named_parameters(), m2.parameters()):
        assert max_rel(q.detach(), p.detach()) < 1e-4, k


def test_graphed_step_needs_capturable_optimiser(cuda):
    m = MK.KuramotoKANClassifier(kuramoto_steps=2, num_basis=8).to(cuda)
    opt = fetode.optim.AdamW(m.parameters(), lr=1e-3)
    x = torch.rand(4, 1, 28, 28, device=cuda)
    y = torch.zeros(4, dtype=torch.long, device=cuda)
    with pytest.raises(ValueError):
        fetode.graphs.GraphedTrainStep(m, opt, F.cross_entropy, x, y)
