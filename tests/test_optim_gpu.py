"""GPU parity of the fused flat Adam / AdamW step (csrc/optim.cu, fetode/optim.py) against torch.optim on the CPU --
the optimisers the reference builds (train_kanfet_node_predprey.py:171; train_ecg_kan_fet_nn_ode.py:1093,1110)."""
import copy

import pytest
import torch

import fetode
from conftest import max_rel

pytestmark = pytest.mark.gpu


def _toy(seed):
    gen = torch.Generator().manual_seed(seed)
    shapes = [(7, 5), (13,), (3, 4, 2), (1,), (6, 6)]
    return [torch.nn.Parameter(torch.randn(s, generator=gen)) for s in shapes], gen


@pytest.mark.parametrize("decoupled,wd,clip", [(False, 0.0, None), (False, 0.01, None), (True, 0.01, 1.0),
                                                (True, 0.05, 0.3)])
def test_flat_adam_matches_torch_optim(cuda, decoupled, wd, clip):
    ref_params, gen = _toy(3)
    idle_ref = torch.nn.Parameter(torch.randn(4, generator=gen))          # never receives a gradient
    dev_params = [torch.nn.Parameter(p.detach().clone().to(cuda)) for p in ref_params]
    idle_dev = torch.nn.Parameter(idle_ref.detach().clone().to(cuda))
    cls = torch.optim.AdamW if decoupled else torch.optim.Adam
    ref_opt = cls(ref_params + [idle_ref], lr=2e-3, weight_decay=wd)
    opt = fetode.optim.FlatAdam(dev_params + [idle_dev], lr=2e-3, weight_decay=wd, decoupled=decoupled)
    for step in range(6):
        ref_opt.zero_grad()
        opt.zero_grad()
        grads = [torch.randn(p.shape, generator=gen) * (3.0 if step % 2 else 0.2) for p in ref_params]
        for p, q, g in zip(ref_params, dev_params, grads):
            p.grad = g.clone()
            if q.grad is None:
                q.grad = g.to(cuda)
            else:
                q.grad.add_(g.to(cuda))                                    # accumulate into the attached flat view
        if clip is not None:
            n_ref = torch.nn.utils.clip_grad_norm_(ref_params + [idle_ref], clip)
            n_dev = opt.clip_grad_norm_(clip)
            assert abs(float(n_dev) - float(n_ref)) < 1e-5 * float(n_ref)
        ref_opt.step()
        opt.step()
        if clip is not None:
            for p, q in zip(ref_params, dev_params):
                assert max_rel(q.grad, p.grad) < 1e-6                      # clip_grad_norm_ rescales .grad in place
    for p, q in zip(ref_params, dev_params):
        assert max_rel(q, p.detach()) < 2e-6
        assert q.data_ptr() >= opt.flat_param.data_ptr()                   # parameters are views of the flat buffer
    assert torch.equal(idle_dev.detach().cpu(), idle_ref.detach())         # untouched, as in torch
    assert opt.built and len(opt.params) == len(dev_params)


def test_flat_adam_drives_lv_training_like_torch_adam(cuda):
    """Three Lotka-Volterra training iterations (fused RK4 forward + reverse sweep): FlatAdam vs torch.optim.Adam on
    two copies of the same KANFET; also checks that the packed images are rebuilt after the raw-pointer update."""
    torch.manual_seed(0)
    m1 = fetode.KANFET([2, 10, 2], grid_size=5).to(cuda)
    m2 = copy.deepcopy(m1)
    o1 = torch.optim.Adam(m1.parameters(), lr=2e-3)
    o2 = fetode.optim.Adam(m2.parameters(), lr=2e-3)
    t = torch.linspace(0, 0.7, 8, dtype=torch.float64)
    x0 = torch.tensor([[1.0, 1.0], [2.0, 0.5], [0.4, 3.0]], device=cuda)
    tgt = torch.randn(8, 3, 2, generator=torch.Generator().manual_seed(1)).to(cuda)
    losses = []
    for model, opt in ((m1, o1), (m2, o2)):
        ls = []
        for _ in range(3):
            model.reset_state()
            opt.zero_grad()
            pred = fetode.odeint(lambda tt, X: model(X), x0, t, method="rk4_classic")
            loss = ((pred - tgt) ** 2).mean()
            loss.backward()
            opt.step()
            ls.append(float(loss.detach()))
        losses.append(ls)
    assert losses[0][0] == pytest.approx(losses[1][0], rel=1e-6)
    assert losses[0][2] == pytest.approx(losses[1][2], rel=1e-4)           # later losses see the updated weights
    assert losses[1][2] != losses[1][0]
    for (n1, p1), (n2, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert max_rel(p2, p1.detach()) < 1e-4, n1
