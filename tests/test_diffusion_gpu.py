"""The graph-replayed reverse-diffusion loop over a KANFET denoiser stack (fetode.diffusion.p_sample_loop,
kan_diffusion_ett.py:565-609 / :421-425): graphed == eager == the oracle restatement on the same noise, with the
hysteresis state of every KANFET layer carried from step to step."""
import pytest
import torch

import fetode
from conftest import max_rel
from oracle import layers as OL
from oracle import models as OM

pytestmark = pytest.mark.gpu


def _schedule(T, device=None):
    betas = torch.linspace(1e-4, 0.05, T)
    alphas = 1.0 - betas
    abar = torch.cumprod(alphas, dim=0)
    ts = (betas, alphas, abar, torch.sqrt(abar), torch.sqrt(1.0 - abar))
    return tuple(t.to(device) for t in ts) if device is not None else ts


_FREQ = {}


def _temb(t):
    f = _FREQ.get(t.device)                                  # built once per device: no H2D copy inside a captured step
    if f is None:
        f = _FREQ[t.device] = torch.tensor([0.1, 0.37, 0.9, 1.7], device=t.device)
    return torch.sin(t.float().unsqueeze(1) * f)


def test_graphed_sampler_equals_eager_and_oracle(cuda):
    pred_len, d_in, T, B = 3, 4, 12, 5
    widths = [(pred_len * d_in + d_in + 4, 16), (16, 16), (16, pred_len * d_in)]
    gen = torch.Generator().manual_seed(11)
    ps = [OL.make_kanlinear_params(i, o, gen, nb=10, branch="ferro2d") for i, o in widths]
    nets = torch.nn.ModuleList()
    for (i, o), p in zip(widths, ps):
        n = fetode.KANFET([i, o])
        n.load_state_dict({f"layers.0.{k}": v for k, v in p.items()}, strict=False)
        nets.append(n)
    nets = nets.to(cuda)

    def model(y, past, t):                                   # epsilon network: stack(cat[y, cond, temb]) (:437-443)
        h = torch.cat([y.reshape(y.shape[0], -1), past[:, -1, :], _temb(t)], dim=1)
        for n in nets:
            h = n(h)
        return h.reshape(y.shape[0], pred_len, d_in)

    sts = [OL.kanfet_states([p])[0] for p in ps]

    def model_cpu(y, past, t):
        h = torch.cat([y.reshape(y.shape[0], -1), past[:, -1, :], _temb(t)], dim=1)
        for p, s in zip(ps, sts):
            h = OL.kanlinear_forward(h, p, branch="ferro2d", branch_state=s)
        return h.reshape(y.shape[0], pred_len, d_in)

    past = torch.randn(B, 7, d_in, generator=gen)
    noise = torch.randn(T + 1, B, pred_len, d_in, generator=gen)
    with torch.no_grad():
        want = OM.p_sample_loop(model_cpu, past, _schedule(T), pred_len, d_in, noise)
    sched = _schedule(T, cuda)
    _temb(torch.zeros(1, dtype=torch.long, device=cuda))
    for n in nets:
        n.reset_state()
    eager = fetode.diffusion.p_sample_loop(model, past.to(cuda), sched, pred_len, d_in, graphed=False,
                                           noise=noise.to(cuda), modules=[nets])
    for n in nets:
        n.reset_state()
    graphed = fetode.diffusion.p_sample_loop(model, past.to(cuda), sched, pred_len, d_in, graphed=True,
                                             noise=noise.to(cuda), modules=[nets])
    assert torch.isfinite(graphed).all()
    assert max_rel(eager, want) < 1e-4
    assert max_rel(graphed, eager) < 1e-6
    # default noise path (torch.randn inside the captured step) runs and stays finite
    out = fetode.diffusion.p_sample_loop(model, past.to(cuda), sched, pred_len, d_in, modules=[nets])
    assert out.shape == (B, pred_len, d_in) and torch.isfinite(out).all()
