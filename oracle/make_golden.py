#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING THE LIVE REFERENCE (``/root/reference``, commit
704a0df4) on seeded inputs.  TEST INFRASTRUCTURE -- see ``oracle/__init__.py``.

Run in the build container only (the GPU box has no reference checkout):
    python oracle/make_golden.py
Every file stores the module's state_dict (keys prefixed ``p.``), the inputs and the outputs the
reference produced, so the GPU box can rebuild identical weights without the reference.
The reference has no golden vectors of its own; these files are what pins the oracle
and the CUDA path there.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_live, solvers  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def _np(d):
    return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


def _sd(module, prefix="p."):
    return {prefix + k: v.detach().clone() for k, v in module.state_dict().items()
            if not k.endswith("prev_x") and not k.endswith("branch_sign")}


def _adversarial_inputs(B, I, gen):
    """Random rows + knot-exact, out-of-grid and near-boundary values (SURVEY section 4 list)."""
    x = torch.randn(B, I, generator=gen) * 1.2
    knots = torch.arange(-3, 9, dtype=torch.float32) * 0.4 - 1.0
    specials = torch.cat([knots, torch.tensor([-2.2000000477, 2.2000000477, -5.0, 5.0, 0.0, 1e-8, -1e-8,
                                               2.1999998, -2.1999998, 0.3, -1.0, 1.0])])
    n = min(B, specials.numel())
    for i in range(I):
        x[:n, i] = specials.roll(i)[:n]
    return x


def save(name, **arrays):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **_np(arrays))
    print(f"{name}: {os.path.getsize(path)} bytes, keys={len(arrays)}")


def main():
    assert ref_live.available(), "reference checkout missing"
    ek, fc = ref_live.efficientkan(), ref_live.ferro_class()
    torch.set_num_threads(1)

    # ---- KANLinear with the logistic branch (efficientkan.py:27-182)
    for (I, O) in [(2, 10), (10, 2), (48, 24)]:
        torch.manual_seed(100 + I)
        layer = ek.KANLinear(I, O)
        gen = torch.Generator().manual_seed(7)
        x = _adversarial_inputs(40, I, gen)
        with torch.no_grad():
            y = layer(x)
            basis = layer.b_splines(x)
        save(f"kanlinear_{I}x{O}", x=x, y=y, bsplines=basis, **_sd(layer))

    # ---- KANLinear without branch + non-uniform grid produced by update_grid (:184-221)
    torch.manual_seed(11)
    layer = ek.KANLinear(6, 5, enable_logistic_basis=False)
    gen = torch.Generator().manual_seed(8)
    layer.update_grid(torch.randn(64, 6, generator=gen) * 0.8)
    x = torch.randn(33, 6, generator=gen)
    with torch.no_grad():
        y = layer(x)
    save("kanlinear_nobranch_adapted", x=x, y=y, **_sd(layer))

    # ---- KAN([2,10,2]) (shipped class) : forward + classic RK4 trajectory (the LV network)
    torch.manual_seed(0)
    kan = ek.KAN([2, 10, 2], grid_size=5)
    gen = torch.Generator().manual_seed(1)
    x0 = torch.rand(64, 2, generator=gen) * 6.75 + 0.25
    x0[:8] = (torch.rand(8, 2, generator=gen) * 2 - 1)          # rows that start inside the spline grid
    with torch.no_grad():
        y = kan(x0)
        traj = solvers.rk4_classic(lambda t, h: kan(h), x0, 0.0, 0.14, 10, trajectory=True)
    save("kan_2_10_2", x=x0, y=y, traj=traj, dt=np.float64(0.014), **_sd(kan))

    # ---- KANFET_ref([2,10,2]) : three successive calls (hysteresis state) + RK4 trajectory
    KANFET = ref_live.kanfet_ref_class()
    torch.manual_seed(0)
    kf = KANFET([2, 10, 2], grid_size=5)
    xs = [x0, x0 + 0.05 * torch.randn(64, 2, generator=gen), x0 - 0.02]
    with torch.no_grad():
        ys = [kf(x) for x in xs]
        kf.reset_state()
        traj = solvers.rk4_classic(lambda t, h: kf(h), x0, 0.0, 0.14, 10, trajectory=True)
    save("kanfet_2_10_2", x0=xs[0], x1=xs[1], x2=xs[2], y0=ys[0], y1=ys[1], y2=ys[2], traj=traj,
         dt=np.float64(0.014), **_sd(kf))

    # ---- TwoDimensionFerroelectricBasis (ferro_class.py:526-596): two calls
    torch.manual_seed(3)
    f2 = fc.TwoDimensionFerroelectricBasis(5, 7)
    xa = torch.randn(9, 5, generator=gen) * 2
    xb = xa + 0.3 * torch.randn(9, 5, generator=gen)
    with torch.no_grad():
        oa, ob = f2(xa), f2(xb)
    save("ferro2d_5x7", xa=xa, xb=xb, oa=oa, ob=ob, **_sd(f2))

    # ---- FerroelectricBasis (ferro_class.py:329-424): B=1 first-call (prev=0) and B>1 (prev=x)
    torch.manual_seed(4)
    ff = fc.FerroelectricBasis(6, 4, 5)
    x1 = torch.randn(1, 6, generator=gen) * 2
    x1b = x1 + 0.2
    with torch.no_grad():
        o1, o1b = ff(x1), ff(x1b)
    ff.reset_state()
    xa = torch.randn(7, 6, generator=gen) * 2
    xb = xa + 0.3 * torch.randn(7, 6, generator=gen)
    with torch.no_grad():
        oa, ob = ff(xa), ff(xb)
    save("ferrofull_6x4x5", x1=x1, x1b=x1b, o1=o1, o1b=o1b, xa=xa, xb=xb, oa=oa, ob=ob, **_sd(ff))

    # ---- NoisyFerroelectricBasis with noise_std=0 (real branch_sign state; :427-523)
    torch.manual_seed(5)
    nf = fc.NoisyFerroelectricBasis(6, 4, 5, noise_std=0.0)
    with torch.no_grad():
        na, nb_, nc = nf(xa), nf(xb), nf(xa * 0.5)
    save("ferronoisy_6x4x5", xa=xa, xb=xb, xc=xa * 0.5, oa=na, ob=nb_, oc=nc, **_sd(nf))

    # ---- hard-gate hysteretic LogisticBasis + KANFeatureMixer + MLPKANODEFunc (train_ecg_kan_fet_nn_ode.py)
    ecg = ref_live.script_module("train_ecg_kan_fet_nn_ode")
    torch.manual_seed(6)
    hg = ecg.LogisticBasis(8, 6)
    xa = torch.randn(12, 8, generator=gen)
    sweep = torch.tensor([0.0, 1e-9, -1e-9, 1e-8, -1e-8, 2e-8, -2e-8, 5e-8, 1e-7, -1e-7, 1e-6, -1e-6])
    with torch.no_grad():
        oa = hg(xa)
        prev_after_a = hg.prev_x.clone()
        xb = prev_after_a[0, :, 0].unsqueeze(0) + sweep.unsqueeze(1)       # dx sweep around the gate threshold
        ob = hg(xb)
    save("hardgate_8x6", xa=xa, xb=xb, oa=oa, ob=ob, prev_after_a=prev_after_a, **_sd(hg))

    torch.manual_seed(7)
    rhs = ecg.MLPKANODEFunc(16, 4, 24)
    h = torch.randn(10, 16, generator=gen)
    with torch.no_grad():
        d1 = rhs(torch.tensor(0.0), h)
        d2 = rhs(torch.tensor(0.1), h + 0.1 * d1)
    save("mlpkanrhs_16_4_24", h=h, d1=d1, d2=d2, **_sd(rhs))

    # ---- KANFetODEFunc (train_ecg.py:986-1013): tanh-bound -> ferro -> tanh -> ferro -> clamp
    tecg = ref_live.script_module("train_ecg")
    torch.manual_seed(8)
    kfo = tecg.KANFetODEFunc(6, 9, 4)
    h = torch.randn(5, 6, generator=gen)
    with torch.no_grad():
        d1 = kfo(torch.tensor(0.0), h)
        d2 = kfo(torch.tensor(0.1), h + 0.1 * d1)
    save("kanfetodefunc_6_9_4", h=h, d1=d1, d2=d2, **_sd(kfo))

    # ---- Kuramoto2D + classifier head (mnist_kuramoto_kan.py:145-225)
    mk = ref_live.script_module("mnist_kuramoto_kan")
    torch.manual_seed(9)
    clf = mk.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8)
    with torch.no_grad():
        clf.osc.omega.normal_(0.0, 0.3, generator=gen)      # the zero init would leave omega untested
    img = torch.rand(6, 1, 28, 28, generator=gen)
    with torch.no_grad():
        feat = clf.osc(img)
        logits = clf(img)
    arrays = dict(img=img, logits=logits, **_sd(clf))
    if feat is not None:
        arrays["feat"] = feat
    save("kuramoto_clf", **arrays)

    # ---- in-repo RK loops on a plain MLP right-hand side (pins the stage algebra itself)
    tett = ref_live.script_module("train_kan_fet_ett")
    torch.manual_seed(10)
    dyn = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 4))
    f = lambda t, z: dyn(z)
    z0 = torch.randn(5, 4, generator=gen)
    with torch.no_grad():
        e = ecg.integrate_euler(f, z0, 0.0, 1.0, 8)
        r2 = ecg.integrate_rk2(f, z0, 0.0, 1.0, 8)
        r4 = ecg.integrate_rk4(f, z0, 0.0, 1.0, 8)
        sub = tett.odeint_rk4(f, z0, torch.linspace(0, 1, 4), n_substeps=3)
    save("rk_loops_mlp", z0=z0, euler=e, rk2=r2, rk4=r4, rk4_sub=sub, **_sd(dyn))


if __name__ == "__main__":
    main()
