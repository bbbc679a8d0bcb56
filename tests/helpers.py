"""Builders shared by the GPU parity tests: identical weights on the oracle side (dict of CPU
tensors) and on the CUDA side (fetode modules)."""
import torch

import fetode
from oracle import layers as OL


def split_layers(params, n_layers):
    """{'layers.0.base_weight': ...} -> [ {base_weight: ...}, ... ] in oracle.layers key convention."""
    out = []
    for li in range(n_layers):
        pre = f"layers.{li}."
        out.append({k[len(pre):]: v for k, v in params.items() if k.startswith(pre)})
    return out


def fetode_kan_from(params, widths, ferro, device, **kw):
    cls = fetode.KANFET if ferro else fetode.KAN
    m = cls(widths, **kw)
    missing, unexpected = m.load_state_dict(params, strict=False)
    assert not unexpected, unexpected
    assert all(k.endswith("prev_x") or k.endswith("branch_sign") for k in missing), missing
    return m.to(device)


def random_kan_params(widths, branch, seed, nb=10):
    gen = torch.Generator().manual_seed(seed)
    layers = [OL.make_kanlinear_params(i, o, gen, nb=nb, branch=branch) for i, o in zip(widths, widths[1:])]
    flat = {}
    for li, p in enumerate(layers):
        for k, v in p.items():
            flat[f"layers.{li}.{k}"] = v
    return layers, flat


def fetode_kan_random(widths, branch, seed, device, nb=10):
    layers, flat = random_kan_params(widths, branch, seed, nb)
    m = fetode.KANFET(widths) if branch == "ferro2d" else fetode.KAN(widths)
    if branch == "none":
        for l in m.layers:
            l.enable_logistic_basis = False
            del l.logistic_basis, l.logistic_weight, l.logistic_scaler
    missing, unexpected = m.load_state_dict(flat, strict=False)
    assert not unexpected, unexpected
    return layers, m.to(device)


def lv_inputs(B, seed, device=None):
    """Synthetic Lotka-Volterra initial conditions X0 ~ U[0.25, 7]^2."""
    gen = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 2, generator=gen) * 6.75 + 0.25
    return x if device is None else x.to(device)
