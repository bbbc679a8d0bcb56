#!/usr/bin/env python
"""Secondary measurements for the BASELINE.json configurations other than the headline ensemble: fwd+bwd(+optimiser)
step times of the training configurations on one B200 next to the reference's CPU arithmetic (oracle port, all host
cores) on a bounded sample.  One JSON object per configuration; `python bench.py` stays the driver's contract.

    python bench_configs.py [--skip-cpu]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import fetode  # noqa: E402
from fetode import ecg, mnist_kuramoto_kan as MK  # noqa: E402

DEV = torch.device("cuda:0")


def gpu_ms(step, iters=5, warm=3):
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def cpu_ms(step, iters=2, warm=1):
    for _ in range(warm):
        step()
    t0 = time.perf_counter()
    for _ in range(iters):
        step()
    return 1e3 * (time.perf_counter() - t0) / iters


def layers_of(model):
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()
          if k.split(".")[-1] not in ("prev_x", "branch_sign", "branch_state")}
    n = len(model.layers)
    return [{k[len(f"layers.{i}."):]: v.requires_grad_(v.is_floating_point() and not k.endswith("grid"))
             for k, v in sd.items() if k.startswith(f"layers.{i}.")} for i in range(n)]


def lv_train(B, skip_cpu):
    from oracle import layers as OL, solvers as OS
    torch.manual_seed(0)
    model = fetode.KANFET([2, 10, 2], grid_size=5)
    lay = layers_of(model)
    model = model.to(DEV)
    opt = torch.optim.Adam(model.parameters(), lr=2e-3)
    n, T = 34, 3.5
    t = torch.linspace(0, T, n + 1, dtype=torch.float64)
    gen = torch.Generator().manual_seed(1)
    x0 = (torch.rand(B, 2, generator=gen) * 6.75 + 0.25) if B > 1 else torch.ones(1, 2)
    target = torch.randn(n + 1, B, 2, generator=gen)
    x0d, tgt = x0.to(DEV), target.to(DEV)

    def step():
        model.reset_state()
        opt.zero_grad(set_to_none=True)
        pred = fetode.odeint(lambda tt, X: model(X), x0d, t, method="rk4_classic")
        ((pred - tgt) ** 2).mean().backward()
        opt.step()
    out = {"config": f"1: LV KANFET([2,10,2]) NODE training, classic RK4 x {n}, MSE, Adam (train_kanfet_node_predprey.py), B={B}",
           "gpu_fwd_bwd_opt_ms": gpu_ms(step), "traj_steps": B * n}
    fopt = fetode.optim.Adam(model.parameters(), lr=2e-3)            # fused flat Adam (csrc/optim.cu), same update rule

    def fstep():
        model.reset_state()
        fopt.zero_grad()
        pred = fetode.odeint(lambda tt, X: model(X), x0d, t, method="rk4_classic")
        ((pred - tgt) ** 2).mean().backward()
        fopt.step()
    out["gpu_fwd_bwd_flatadam_ms"] = gpu_ms(fstep)
    # the same iteration captured into a CUDA graph and replayed (fetode/graphs.py)

    def gfwd(m, x):
        m.reset_state()
        return fetode.odeint(lambda tt, X: m(X), x, t, method="rk4_classic")
    gopt = fetode.optim.Adam(model.parameters(), lr=2e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, lambda pred, y: ((pred - y) ** 2).mean(), x0d, tgt, warmup=2,
                                           forward=gfwd)
    out["gpu_fwd_bwd_graph_ms"] = gpu_ms(lambda: gstep(gstep.x, gstep.y))
    out["gpu_traj_steps_per_s"] = B * n / out["gpu_fwd_bwd_opt_ms"] * 1e3
    if not skip_cpu:
        Bc = min(B, 2048)
        params = [v for l in lay for v in l.values() if v.requires_grad]

        def cstep():
            for p in params:
                p.grad = None
            states = OL.kanfet_states(lay)
            f = lambda tt, h: OL.kan_forward(h, lay, branch="ferro2d", states=states)
            pred = OS.rk4_classic(f, x0[:Bc], 0.0, T, n, trajectory=True)
            ((pred - target[:, :Bc]) ** 2).mean().backward()
        ms = cpu_ms(cstep)
        out.update(cpu_fwd_bwd_ms=ms, cpu_rows=Bc, cpu_traj_steps_per_s=Bc * n / ms * 1e3, cpu_cores=os.cpu_count())
    return out


def ecg_train(B, skip_cpu):
    torch.manual_seed(0)
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    model = model.to(DEV)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(B, 96, generator=gen)
    x = ((x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)).to(DEV)
    y = torch.randint(0, 2, (B,), generator=gen).to(DEV)

    def step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x), y).backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()

    def fwd():
        with torch.no_grad():
            model(x)
    ms, fms = gpu_ms(step, iters=3, warm=2), gpu_ms(fwd, iters=3, warm=2)
    fopt = fetode.optim.AdamW(model.parameters(), lr=1e-3)           # fused flat AdamW + on-device clipping

    def fstep():
        fopt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        fopt.clip_grad_norm_(1.0)
        fopt.step()
    fl = gpu_ms(fstep, iters=3, warm=2)
    # the same step captured once into a CUDA graph and replayed (fetode/graphs.py): no host work per launch
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, max_norm=1.0, warmup=2)
    gl = gpu_ms(lambda: gstep(gstep.x, gstep.y), iters=5, warm=2)
    # the reference's own solver choice for this model: dopri5 (rtol 1e-3, atol 1e-4, :1034-1041) -- adaptive, one
    # host read of the error norm per attempted step, so it runs eagerly
    dp = {}
    try:
        model.ode_method = "dopri5"
        dp["gpu_dopri5_fwd_bwd_opt_ms"] = gpu_ms(fstep, iters=3, warm=2)
    except Exception as e:                                            # noqa: BLE001
        dp["gpu_dopri5_error"] = repr(e)[:200]
    model.ode_method = "rk4_classic"
    return {**dp, "gpu_fwd_bwd_graph_ms": gl, "gpu_fwd_bwd_flatadamw_ms": fl, "config": f"3: ECG KanFet_MLP_NODE(T=96, latent 64, nb 12, hidden 128), classic RK4 x 8, CE, AdamW, clip 1.0, B={B}",
            "gpu_fwd_bwd_opt_ms": ms, "gpu_fwd_ms": fms, "sample_steps_per_s": B * 8 / ms * 1e3}


def ett_encoder(B, skip_cpu):
    from oracle import layers as OL
    torch.manual_seed(0)
    enc = torch.nn.Sequential(torch.nn.Flatten(), fetode.KAN([336 * 7, 128]), torch.nn.ReLU(), fetode.KAN([128, 64])).to(DEV)
    x = torch.randn(B, 336, 7, device=DEV)

    def step():
        enc.zero_grad(set_to_none=True)
        enc(x).square().mean().backward()

    def fwd():
        with torch.no_grad():
            enc(x)
    out = {"config": f"4: ETT encoder Flatten -> KAN([2352,128]) -> ReLU -> KAN([128,64]) (train_kan_fet_ett.py:676-681), B={B}",
           "gpu_fwd_ms": gpu_ms(fwd), "gpu_fwd_bwd_ms": gpu_ms(step, iters=3, warm=2)}
    if not skip_cpu:
        Bc = min(B, 256)
        l1 = {k: v.detach().cpu() for k, v in enc[1].layers[0].state_dict().items()}
        l2 = {k: v.detach().cpu() for k, v in enc[3].layers[0].state_dict().items()}
        xc = x[:Bc].reshape(Bc, -1).cpu()

        def cfwd():
            with torch.no_grad():
                OL.kanlinear_forward(torch.relu(OL.kanlinear_forward(xc, l1)), l2)
        ms = cpu_ms(cfwd)
        out.update(cpu_fwd_ms=ms, cpu_rows=Bc, cpu_fwd_ms_scaled_to_B=ms * B / Bc, cpu_cores=os.cpu_count())
    return out


def mnist_train(B, skip_cpu):
    torch.manual_seed(0)
    model = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8).to(DEV)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    x = torch.rand(B, 1, 28, 28, device=DEV)
    y = torch.randint(0, 10, (B,), device=DEV)

    def step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x), y).backward()
        opt.step()
    ms = gpu_ms(step)
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, warmup=2)
    gl = gpu_ms(lambda: gstep(gstep.x, gstep.y))
    return {"config": f"5: MNIST KuramotoKANClassifier(steps=10, nb=8), CE, AdamW, B={B}", "gpu_fwd_bwd_opt_ms": ms,
            "image_steps_per_s": B / ms * 1e3, "gpu_fwd_bwd_graph_ms": gl, "graph_image_steps_per_s": B / gl * 1e3}


def ett_rnn_encoder(B, skip_cpu):
    """Config 4, "KAN-FET" variant: KANRNNEncoder(7 features, hidden 64, latent 64, nb 10) over a 336-step context
    (train_kan_fet_ett.py:799-818) -- one launch per direction instead of 336 cell calls."""
    from fetode import ett
    from oracle import layers as OL
    torch.manual_seed(0)
    enc = ett.KANRNNEncoder(7, 64, 64, 10)
    p = {k: v.detach().clone() for k, v in enc.state_dict().items()}
    enc = enc.to(DEV)
    x = torch.randn(B, 336, 7, device=DEV)

    def step():
        enc.zero_grad(set_to_none=True)
        enc(x).square().mean().backward()

    def fwd():
        with torch.no_grad():
            enc(x)
    out = {"config": f"4: ETT KANRNNEncoder(F=7, hidden 64, latent 64, nb 10), context 336 (train_kan_fet_ett.py:799-818), B={B}",
           "gpu_fwd_ms": gpu_ms(fwd), "gpu_fwd_bwd_ms": gpu_ms(step)}
    if not skip_cpu:
        Bc = min(B, 64)
        xc = x[:Bc].cpu()

        def cfwd():
            with torch.no_grad():
                OL.kan_rnn_encoder(xc, p)
        ms = cpu_ms(cfwd)
        out.update(cpu_fwd_ms=ms, cpu_rows=Bc, cpu_cores=os.cpu_count())
    return out


def kanfet_stack_sampler(B, skip_cpu):
    """Config 4: the 250-step reverse-diffusion loop over KANFET([928,256]) -> ([256,256]) -> ([256,672])
    (kan_diffusion_ett.py:421-425, 565-609; pred_len 96 x 7 channels, cond 128 + time 128 = 928 inputs), one CUDA-graph
    replay per step against the eager loop."""
    torch.manual_seed(0)
    pred_len, d_in, T = 96, 7, 250
    nets = torch.nn.ModuleList([fetode.KANFET([928, 256]), fetode.KANFET([256, 256]), fetode.KANFET([256, 672])]).to(DEV)
    cond = torch.randn(B, 128, device=DEV)
    freq = torch.exp(torch.arange(64, device=DEV) * (-9.21 / 64))

    def model(y, past, t):
        ang = t.float().unsqueeze(1) * freq
        h = torch.cat([y.reshape(B, -1), cond, torch.sin(ang), torch.cos(ang)], dim=1)
        for n in nets:
            h = n(h)
        return h.reshape(B, pred_len, d_in)
    betas = torch.linspace(1e-4, 0.02, T, device=DEV)
    alphas = 1.0 - betas
    abar = torch.cumprod(alphas, 0)
    sched = (betas, alphas, abar, abar.sqrt(), (1 - abar).sqrt())
    past = torch.randn(B, 336, d_in, device=DEV)

    def run(graphed):
        for n in nets:
            n.reset_state()
        return fetode.diffusion.p_sample_loop(model, past, sched, pred_len, d_in, graphed=graphed, modules=[nets])
    out = {"config": f"4: DDPM p_sample_loop x {T} over the KANFET 928->256->256->672 stack (kan_diffusion_ett.py:421-425, 565-609), B={B}"}
    out["gpu_eager_ms"] = gpu_ms(lambda: run(False), iters=2, warm=1)
    out["gpu_graphed_ms"] = gpu_ms(lambda: run(True), iters=2, warm=1)        # includes the one-off capture
    out["gpu_graphed_ms_per_step"] = out["gpu_graphed_ms"] / T
    out["finite"] = bool(torch.isfinite(run(True)).all())
    return out


def _commit():
    import subprocess
    try:
        return subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip() or None
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-cpu", action="store_true")
    a = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    commit = os.environ.get("FETODE_COMMIT") or _commit()
    for fn, args in ((lv_train, (1,)), (lv_train, (65536,)), (ecg_train, (4096,)), (ett_encoder, (4096,)),
                     (ett_encoder, (64,)), (ett_rnn_encoder, (64,)), (ett_rnn_encoder, (4096,)),
                     (kanfet_stack_sampler, (64,)), (mnist_train, (128,)), (mnist_train, (1024,))):
        try:
            print(json.dumps({"commit": commit, **fn(*args, a.skip_cpu)}), flush=True)
        except Exception as e:                      # keep going: this is a survey, not the contract bench
            print(json.dumps({"config": f"{fn.__name__}{args}", "error": repr(e)[:300]}), flush=True)


if __name__ == "__main__":
    main()
