"""The reverse-diffusion sampling loop around the KANFET denoiser stack, graph-replayed.

  p_sample_loop(model, past, schedule_tensors, pred_len, d_in)      kan_diffusion_ett.py:565-609

Same signature and arithmetic as the reference (epsilon-prediction DDPM: y0_hat, posterior mean and variance per step,
:581-605).  The reference runs T (250) iterations of ~40 tiny launches plus a dozen 0-dim tensor ops from Python; here
ONE step -- denoiser call, update, step counter -- is captured into a CUDA graph (fetode.graphs.GraphedCall) and replayed
T times: the per-step scalars come from device tables indexed by a device-side step counter, so nothing synchronises
with the host.  ``graphed=False`` runs the same step function eagerly (the two are compared in the tests).
``noise`` ([T + 1, B, pred_len, d_in]: initial y, then z for ti = T-1 .. 0) makes the loop deterministic for parity
tests; by default it is drawn with torch.randn like the reference.
"""
from __future__ import annotations

from typing import Optional

import torch

from .graphs import GraphedCall


def _tables(schedule_tensors):
    """Per-step coefficients of :581-600, computed with the reference's expressions (float32, on the device)."""
    betas, alphas, abar, _sqrt_abar, _sqrt_1m_abar = schedule_tensors
    one = torch.ones(1, device=abar.device, dtype=abar.dtype)
    abar_prev = torch.cat([one, abar[:-1]])
    sab = torch.sqrt(abar)
    s1m = torch.sqrt(1.0 - abar)
    beta_tilde = betas * (1.0 - abar_prev) / (1.0 - abar + 1e-8)
    c1 = (torch.sqrt(abar_prev) * betas) / (1.0 - abar + 1e-8)
    c2 = (torch.sqrt(alphas) * (1.0 - abar_prev)) / (1.0 - abar + 1e-8)
    sigma = torch.sqrt(beta_tilde + 1e-8)
    sigma[0] = 0.0                                                    # ti == 0: y = mean (:604-605)
    return torch.stack([sab, s1m, c1, c2, sigma], dim=0).contiguous()  # [5, T]


@torch.no_grad()
def p_sample_loop(model, past, schedule_tensors, pred_len, d_in, *, graphed: bool = True,
                  noise: Optional[torch.Tensor] = None, modules=None):
    """Generate y0 ~ p(y | past) by reverse diffusion; returns (B, pred_len, d_in).  ``model(y, past, t)`` is the
    epsilon network (e.g. KAN_FET_ALL_NODE, kan_diffusion_ett.py:402-444); ``modules`` (default: ``[model]`` when it is an
    nn.Module) are searched for stateful neurons whose hysteresis state must survive from step to step."""
    if not past.is_cuda:
        raise RuntimeError("fetode.diffusion.p_sample_loop: tensors must live on a CUDA device (there is no CPU path)")
    tabs = _tables(schedule_tensors)
    T = tabs.shape[1]
    B = past.shape[0]
    dev = past.device
    if modules is None:
        modules = [model] if isinstance(model, torch.nn.Module) else []
    y = (noise[0] if noise is not None else torch.randn(B, pred_len, d_in, device=dev)).clone().float()
    ti = torch.full((1,), T - 1, device=dev, dtype=torch.long)        # device-side step counter
    zsrc = noise[1:] if noise is not None else None                   # z of step ti lives at row (T - 1 - ti)

    def step(y_buf, ti_buf):
        t = ti_buf.expand(B)
        eps_hat = model(y_buf, past, t)
        c = tabs.index_select(1, ti_buf)                               # [5, 1]
        y0_hat = (y_buf - c[1] * eps_hat) / (c[0] + 1e-8)
        mean = c[2] * y0_hat + c[3] * y_buf
        z = zsrc.index_select(0, (T - 1) - ti_buf)[0] if zsrc is not None else torch.randn_like(y_buf)
        y_buf.copy_(mean + c[4] * z)
        ti_buf.sub_(1)
        return y_buf

    if not graphed:
        for _ in range(T):
            step(y, ti)
        return y
    call = GraphedCall(step, [y, ti], modules=modules, warmup=1)
    # the warm-up and the capture advanced the buffers and the neurons' state: start over from the given y / ti
    call.inputs[0].copy_(y)
    call.inputs[1].fill_(T - 1)
    call.reset_state()
    for _ in range(T):
        call.graph.replay()
    return call.inputs[0].clone()
