#!/usr/bin/env python
"""bench.py -- headline benchmark: Lotka-Volterra KAN-FET NODE inference ensemble (BASELINE.json configs[1]) plus the
fwd+bwd training steps of configs[0] and configs[2] in the same JSON line.

Primary workload: KANFET([2,10,2], grid 5) right-hand side, classic RK4, 2^20 synthetic initial conditions x 1000 steps
(dt = 0.014) IN TOTAL, batch-sharded over the N ranks (131 072 rows per GPU at N = 8; STRONG scaling, as the config
states it; no data-path collective).  One "step" of this script = one pass of the hot path over the batch = one launch
of the fused kernel per rank = 2^20 x 1000 trajectory-steps in total.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-train]
    torchrun ... bench.py --gpus N ...      (one rank per GPU)

Prints ONE JSON line (rank 0).
  value      device-timed, inputs resident in HBM (CUDA events around each launch, max over ranks)
  e2e        the same metric through the public API (fetode.integrate_rk4) with the initial conditions copied from pinned
             host memory and the final states copied back inside the timed region
  weak       (N > 1) the same kernel with 2^20 rows PER GPU -- the round-1 figure, kept as a secondary key
  fwd_bwd    BASELINE's "fwd+bwd step ms": LV training (config 1: 34 classic-RK4 steps, MSE, Adam; B = 1 as the script runs
             it and B = 65 536 per GPU) and ECG training (config 3: B = 4096 per GPU, RK4 x 8, cross-entropy, clip 1.0,
             AdamW), whole step replayed from a CUDA graph, the flat-gradient NCCL all-reduce inside the step at N > 1
             (allreduce_ms = the collective alone), each with its roofline fraction and, at N = 1, its CPU leg; plus the
             config-4 ETT encoder (fwd+bwd, B = 4096) and the config-5 MNIST-Kuramoto training step (B = 1024 per GPU)
  roofline   dominant kernel of the primary workload against the FP32 CUDA-core peak measured live (FFMA probe)
  cpu_baseline (N = 1) the oracle port of the reference's PyTorch arithmetic on all host cores, bounded sample
`--impl reference` times that CPU port alone (rank 0), same metric / config, 65 536 trajectories x 10 steps per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

WIDTHS = [2, 10, 2]
RK_STEPS = 1000
DT = 0.014
TRAJ_TOTAL = 1 << 20
FLOP_PER_TRAJ_STEP = 22712          # KANFET([2,10,2]) ferro branch, classic RK4
ECG_FLOP_PER_ROW_RHS = 4.49e6       # ECG right-hand side, logistic branch nb = 10
METRIC = "lv_kanfet_node_traj_steps_per_s"
UNIT = "traj-steps/s"
WORKLOAD = "LV KANFET([2,10,2]) inference ensemble, 2^20 initial conditions x 1000 classic-RK4 steps, dt=0.014 (BASELINE configs[1])"
CPU_SAMPLE = (65536, 10)            # the CPU slice of config 2


def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_model(seed=0):
    """Random-init KANFET of the reference's LV architecture (train_kanfet_node_predprey.py:146)."""
    import fetode
    torch.manual_seed(seed)
    return fetode.KANFET(layers_hidden=WIDTHS, grid_size=5)


def oracle_layers(model, grad=False):
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()
          if not k.endswith("prev_x") and not k.endswith("branch_sign")}
    out = [{k[len(f"layers.{i}."):]: v for k, v in sd.items() if k.startswith(f"layers.{i}.")}
           for i in range(len(WIDTHS) - 1)]
    if grad:
        for lay in out:
            for k, v in lay.items():
                if not k.endswith("grid"):
                    v.requires_grad_(True)
    return out


def synth_x0(n, seed):
    gen = torch.Generator().manual_seed(seed)
    return torch.rand(n, 2, generator=gen) * 6.75 + 0.25


# ------------------------------------------------------------------------------------------
# CPU legs (oracle port of the reference's PyTorch path; test infrastructure used as the reported baseline)
# ------------------------------------------------------------------------------------------
def cpu_reference_run(layers, n_traj, n_steps, seed=0):
    """Seconds to integrate n_traj trajectories for n_steps classic-RK4 steps on the host cores."""
    from oracle import layers as OL
    from oracle import solvers as OS
    x0 = synth_x0(n_traj, seed)
    states = OL.kanfet_states(layers)
    f = lambda t, h: OL.kan_forward(h, layers, branch="ferro2d", states=states)
    with torch.no_grad():
        t0 = time.perf_counter()
        OS.rk4_classic(f, x0, 0.0, n_steps * DT, n_steps)
        return time.perf_counter() - t0


def cpu_lv_train_ms(model, B, iters=2):
    """fwd+bwd of one LV training iteration (34 classic-RK4 steps, MSE) on the host cores, B rows."""
    from oracle import models as OM
    lay = oracle_layers(model, grad=True)
    params = [v for l in lay for v in l.values() if v.requires_grad]
    gen = torch.Generator().manual_seed(1)
    x0 = synth_x0(B, 1) if B > 1 else torch.ones(1, 2)
    target = torch.randn(35, B, 2, generator=gen)

    def step():
        for p in params:
            p.grad = None
        OM.lv_train_loss(lay, x0, target, 3.5, 34).backward()
    step()
    t0 = time.perf_counter()
    for _ in range(iters):
        step()
    return 1e3 * (time.perf_counter() - t0) / iters


def cpu_lv_dopri5_ms(model, rtol, atol):
    """fwd+bwd of one LV training iteration with the adaptive solver (oracle port of dopri5, torch autograd) on the host
    cores, one trajectory (the script's batch); returns (ms, accepted, rejected)."""
    from oracle import layers as OL
    from oracle import solvers as OS
    lay = oracle_layers(model, grad=True)
    gen = torch.Generator().manual_seed(1)
    x0 = torch.ones(1, 2)
    target = torch.randn(35, 1, 2, generator=gen)
    t = torch.linspace(0, 3.5, 35, dtype=torch.float64)
    st = {}
    states = OL.kanfet_states(lay)
    f = lambda tt, h: OL.kan_forward(h, lay, branch="ferro2d", states=states)
    t0 = time.perf_counter()
    pred = OS.odeint(f, x0, t, rtol=rtol, atol=atol, stats=st)
    ((pred - target) ** 2).mean().backward()
    return 1e3 * (time.perf_counter() - t0), st.get("accepted"), st.get("rejected")


def cpu_ecg_train_ms(model, B):
    """fwd+bwd of one ECG training step (RK4 x 8, cross-entropy) on the host cores, B rows (bounded sample)."""
    import torch.nn.functional as F

    from oracle import models as OM
    P = OM.ecg_params(model.state_dict(), torch.float32, requires_grad=True)
    st = OM.EcgStates(P)
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(B, 96, generator=gen)
    x = (x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)
    y = torch.randint(0, 2, (B,), generator=gen)
    t0 = time.perf_counter()
    F.cross_entropy(OM.kanfet_mlp_node_forward(x, P, st), y).backward()
    return 1e3 * (time.perf_counter() - t0)


def run_reference(args):
    rank, world, _ = _dist_env()
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    layers = oracle_layers(make_model())
    n_traj, n_steps = CPU_SAMPLE
    for _ in range(args.warmup):
        cpu_reference_run(layers, n_traj, n_steps)
    secs = [cpu_reference_run(layers, n_traj, n_steps, seed=i) for i in range(args.steps)]
    tot = sum(secs)
    value = n_traj * n_steps * args.steps / tot
    sample = f"{n_traj} trajectories x {n_steps} classic-RK4 steps per step (oracle port, torch CPU, no_grad)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self._stop, self._th = index, [], threading.Event(), None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._th = threading.Thread(target=self._loop, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._th.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# GPU leg
# ------------------------------------------------------------------------------------------
class Ctx:
    """Rank / device / timing helpers shared by the sections of the GPU arm."""

    def __init__(self, args):
        import torch.distributed as dist
        self.dist = dist
        self.rank, self.world, self.local = _dist_env()
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N > 1")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)     # > 126 MB L2

    def barrier(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return v
        t = torch.tensor([v], device=self.dev, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step_fn, k, w, flush=True, grad=False):
        """W warm-up calls, then K calls each bracketed by CUDA events on the launching stream (L2 flushed before each
        when flush); barrier + synchronize on both sides; returns total ms of the K calls, max over ranks."""
        with torch.set_grad_enabled(grad):
            for _ in range(w):
                step_fn()
            self.barrier()
            evs = []
            for _ in range(k):
                if flush:
                    self.flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                step_fn()
                e1.record()
                evs.append((e0, e1))
            self.barrier()
        return self.max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))


def measure_peaks(ctx):
    """FP32 CUDA-core peak (live FFMA probe in libfetode) and dense TF32 tensor peak (cuBLAS through torch.matmul,
    the way MEASURED_PEAKS.json measured bf16) -- the two denominators MEASURED_PEAKS.json does not hold."""
    import ctypes as C

    from fetode._lib import lib
    tf = C.c_double(0.0)
    rc = lib().fetode_probe_fp32_peak(C.byref(tf), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    fp32 = tf.value if rc == 0 and tf.value > 0 else None
    tf32 = None
    try:
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        n = 8192
        a = torch.randn(n, n, device=ctx.dev)
        b = torch.randn(n, n, device=ctx.dev)
        for _ in range(2):
            a @ b
        best = 0.0
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = max(best, 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        tf32 = best
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b
    except Exception:
        pass
    return fp32, tf32


def lv_train_section(ctx, B, fp32_peak, with_cpu):
    """Config 1: LV KANFET([2,10,2]) NODE training iteration (train_kanfet_node_predprey.py:246-257), classic RK4 x 34,
    MSE over all samples, Adam lr 2e-3, B rows per GPU; whole step replayed from a CUDA graph."""
    import fetode
    from fetode import dist as fdist
    model = make_model().to(ctx.dev)
    n, T = 34, 3.5
    t = torch.linspace(0, T, n + 1, dtype=torch.float64)
    gen = torch.Generator().manual_seed(100 + ctx.rank)
    x0 = (torch.rand(B, 2, generator=gen) * 6.75 + 0.25) if B > 1 else torch.ones(1, 2)
    x0d = x0.to(ctx.dev)
    tgt = torch.randn(n + 1, B, 2, generator=gen).to(ctx.dev)
    func = fetode.as_rhs(model)

    def gfwd(m, x):
        m.reset_state()
        return fetode.odeint(func, x, t, method="rk4_classic")
    mse = lambda pred, y: ((pred - y) ** 2).mean()
    eopt = fetode.optim.Adam(model.parameters(), lr=2e-3)

    def eager():
        eopt.zero_grad()
        mse(gfwd(model, x0d), tgt).backward()
        if ctx.world > 1:
            fdist.allreduce_gradients(eopt)
        eopt.step()
    k = 5
    eager_ms = ctx.timed(eager, k, 3, flush=False, grad=True) / k
    gopt = fetode.optim.Adam(model.parameters(), lr=2e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, mse, x0d, tgt, warmup=2, forward=gfwd, allreduce=ctx.world > 1)
    k = 10
    ms = ctx.timed(lambda: gstep(gstep.x, gstep.y), k, 3, flush=False, grad=True) / k
    ar_ms = None
    if ctx.world > 1:
        ar_ms = ctx.timed(lambda: ctx.dist.all_reduce(gopt.flat_grad), 20, 5, flush=False) / 20
    loss = float(gstep.loss)
    units = B * n * ctx.world
    ach = 3.0 * FLOP_PER_TRAJ_STEP * B * n / (ms * 1e-3) / 1e12
    out = {"workload": f"config 1: LV KANFET([2,10,2]) NODE training, classic RK4 x {n}, MSE, fused Adam "
                       "(train_kanfet_node_predprey.py:246-257)",
           "rows_per_gpu": B, "ms_per_step": ms, "eager_ms_per_step": eager_ms, "graph_replayed": True,
           "traj_steps_per_s": units / (ms * 1e-3), "loss_finite": loss == loss,
           "allreduce_ms": ar_ms, "allreduce_bytes": int(gopt.flat_grad.numel()) * 4 if ctx.world > 1 else 0,
           "roofline": {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
                        "frac": ach / fp32_peak if fp32_peak else None,
                        "flop_per_traj_step": 3 * FLOP_PER_TRAJ_STEP, "convention": "fwd+bwd = 3 x the SURVEY 8(d) forward count"}}
    if with_cpu:
        Bc = min(B, 2048)
        cms = cpu_lv_train_ms(make_model(), Bc)                          # same seed: the same weights on the host
        out["cpu"] = {"ms_per_step": cms, "rows": Bc, "traj_steps_per_s": Bc * n / (cms * 1e-3), "cores": os.cpu_count(),
                      "kind": "port", "sample": f"fwd+bwd of {Bc} rows x {n} classic-RK4 steps (oracle port, torch autograd)"}
    del gstep, gopt
    return out


def lv_dopri5_section(ctx, B, with_cpu=False, rtol=1e-7, atol=1e-9):
    """Config 1 with the script's own solver call -- ``odeint(calDeriv, X0, t)`` = torchdiffeq's default dopri5 at its
    default tolerances rtol 1e-7 / atol 1e-9 (train_kanfet_node_predprey.py:252).  B <= 12: the accept / reject controller
    runs inside the kernel (fetode_smallnet_dp5_solve_fwd / _bwd), one launch per 4096 accepted steps; larger batches: one
    launch per attempted step and one 8-byte read-back (eager either way, not graph-replayed)."""
    import fetode
    model = make_model().to(ctx.dev)
    n, T = 34, 3.5
    t = torch.linspace(0, T, n + 1, dtype=torch.float64)
    gen = torch.Generator().manual_seed(100 + ctx.rank)
    x0 = (torch.rand(B, 2, generator=gen) * 6.75 + 0.25) if B > 1 else torch.ones(1, 2)
    x0d = x0.to(ctx.dev)
    tgt = torch.randn(n + 1, B, 2, generator=gen).to(ctx.dev)
    func = fetode.as_rhs(model)
    opt = fetode.optim.Adam(model.parameters(), lr=2e-3)
    st = {}

    def it():
        opt.zero_grad()
        model.reset_state()
        pred = fetode.odeint(func, x0d, t, rtol=rtol, atol=atol, options={"stats": st})
        ((pred - tgt) ** 2).mean().backward()
        opt.step()
    k = 3
    ms = ctx.timed(it, k, 1, flush=False, grad=True) / k
    att = st.get("accepted", 0) + st.get("rejected", 0)
    out = {"workload": "config 1 with the script's solver call: LV KANFET([2,10,2]) NODE training iteration, adaptive dopri5 "
                        f"(rtol {rtol:g}, atol {atol:g}), t in [0, {T}] x {n + 1} samples, MSE, fused Adam "
                        "(train_kanfet_node_predprey.py:246-257)",
            "rows_per_gpu": B, "ms_per_step": ms, "graph_replayed": False, "fused_attempt_kernels": bool(st.get("fused")),
            "controller_on_device": bool(st.get("device_controller")),
            "accepted_steps": st.get("accepted"), "rejected_steps": st.get("rejected"),
            "us_per_attempt_fwd_bwd": ms * 1e3 / att if att else None,
            "other_paths": "profiles/r2_dopri5_fused_vs_staged.jsonl (host-controlled attempt launches, per-stage launches)"}
    if with_cpu and B == 1:
        cms, acc, rej = cpu_lv_dopri5_ms(make_model(), rtol, atol)
        out["cpu"] = {"ms_per_step": cms, "accepted_steps": acc, "rejected_steps": rej, "cores": os.cpu_count(), "kind": "port",
                      "sample": "one iteration, one trajectory: oracle dopri5 + torch autograd (same weights, same tolerances)"}
    return out


def ecg_train_section(ctx, B, tf32_peak, with_cpu):
    """Config 3: KanFet_MLP_NODE(T=96, latent 64, nb 12, hidden 128) training step (train_ecg_kan_fet_nn_ode.py:1100-1112),
    classic RK4 x 8, cross-entropy, clip_grad_norm_(1.0), AdamW; B rows per GPU.  At N > 1 every rank but the last
    appends the GLOBAL last row (sentinel) so the hard gate sees the reference's prev_x (:129-132) without communication."""
    import torch.nn.functional as F

    import fetode
    from fetode import ecg
    torch.manual_seed(0)
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    cpu_sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    model = model.to(ctx.dev)

    def rows_of(rank):
        gen = torch.Generator().manual_seed(200 + rank)
        x = torch.randn(B, 96, generator=gen)
        x = (x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)               # per-row z-score :44
        return x, torch.randint(0, 2, (B,), generator=gen)
    x, y = rows_of(ctx.rank)
    sentinel = ctx.world > 1 and ctx.rank != ctx.world - 1
    if sentinel:
        x = torch.cat([x, rows_of(ctx.world - 1)[0][-1:]], dim=0)
    x, y = x.to(ctx.dev), y.to(ctx.dev)
    fwd = (lambda m, inp: m(inp)[:B]) if sentinel else (lambda m, inp: m(inp))
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, max_norm=1.0, warmup=2, forward=fwd,
                                           allreduce=ctx.world > 1)
    k = 5
    ms = ctx.timed(lambda: gstep(gstep.x, gstep.y), k, 2, flush=False, grad=True) / k
    ar_ms = None
    if ctx.world > 1:
        ar_ms = ctx.timed(lambda: ctx.dist.all_reduce(gopt.flat_grad), 20, 5, flush=False) / 20
    loss = float(gstep.loss)
    n_rk = 8
    ach = 3.0 * 4 * n_rk * ECG_FLOP_PER_ROW_RHS * B / (ms * 1e-3) / 1e12
    peak = tf32_peak / 3.0 if tf32_peak else None
    out = {"workload": "config 3: ECG KanFet_MLP_NODE(T=96, latent 64, nb 12, hidden 128), classic RK4 x 8, CE, clip 1.0, "
                       "fused AdamW (train_ecg_kan_fet_nn_ode.py:1100-1112)",
           "rows_per_gpu": B, "ms_per_step": ms, "graph_replayed": True, "sample_steps_per_s": B * n_rk * ctx.world / (ms * 1e-3),
           "loss_finite": loss == loss, "hardgate_rule": "sentinel row (global last row on every rank)" if ctx.world > 1 else "single rank",
           "allreduce_ms": ar_ms, "allreduce_bytes": int(gopt.flat_grad.numel()) * 4 if ctx.world > 1 else 0,
           "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                        "frac": ach / peak if peak else None,
                        "peak_source": "cuBLAS TF32 8192^3 through torch.matmul, measured in this run, / 3 (3xTF32 split)",
                        "tf32_dense_tflops": tf32_peak, "flop_per_row_rhs": ECG_FLOP_PER_ROW_RHS,
                        "convention": "useful FLOP: fwd+bwd = 3 x 32 right-hand sides x SURVEY 8(d) count"}}
    if with_cpu:
        Bc = 128
        model_cpu = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
        model_cpu.load_state_dict(cpu_sd)
        cms = cpu_ecg_train_ms(model_cpu, Bc)
        out["cpu"] = {"ms_per_step": cms, "rows": Bc, "sample_steps_per_s": Bc * n_rk / (cms * 1e-3), "cores": os.cpu_count(),
                      "kind": "port", "sample": f"fwd+bwd of {Bc} rows, RK4 x 8 (oracle port, torch autograd); per-row cost is flat in B"}
    del gstep, gopt
    return out


def ett_encoder_section(ctx, B, tf32_peak):
    """Config 4: the wide KAN encoder of the ETT forecaster, Flatten -> KAN([336*7, 128]) -> ReLU -> KAN([128, 64])
    (train_kan_fet_ett.py:676-681), fwd+bwd of a squared-output loss, B rows per GPU (no optimiser: the encoder is one
    block of the forecaster)."""
    import fetode
    torch.manual_seed(0)
    enc = torch.nn.Sequential(torch.nn.Flatten(), fetode.KAN([336 * 7, 128]), torch.nn.ReLU(), fetode.KAN([128, 64])).to(ctx.dev)
    gen = torch.Generator().manual_seed(300 + ctx.rank)
    x = torch.randn(B, 336, 7, generator=gen).to(ctx.dev)

    def step():
        enc.zero_grad(set_to_none=True)
        enc(x).square().mean().backward()

    def fwd():
        with torch.no_grad():
            enc(x)
    k = 10
    ms = ctx.timed(step, k, 3, flush=False, grad=True) / k
    fms = ctx.timed(fwd, k, 3, flush=False) / k
    flop_row = 12.0e6                                              # encoder with the logistic branch
    ach = 3.0 * flop_row * B / (ms * 1e-3) / 1e12
    peak = tf32_peak / 3.0 if tf32_peak else None
    return {"workload": "config 4: ETT encoder KAN([2352,128]) -> ReLU -> KAN([128,64]) (train_kan_fet_ett.py:676-681), fwd+bwd",
            "rows_per_gpu": B, "ms_per_step": ms, "fwd_ms": fms, "graph_replayed": False,
            "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak else None,
                         "flop_per_row": flop_row, "convention": "useful FLOP, fwd+bwd = 3 x forward; TF32 dense (cuBLAS, this run) / 3"}}


def mnist_train_section(ctx, B, fp32_peak):
    """Config 5: KuramotoKANClassifier(steps=10, num_basis=8) training step (mnist_kuramoto_kan.py:228-283): 10 Euler
    steps of the phase lattice + KANLinear(1568 -> 10) head, cross-entropy, fused AdamW; B images per GPU,
    graph-replayed, flat-gradient all-reduce inside at N > 1."""
    import torch.nn.functional as F

    import fetode
    from fetode import mnist_kuramoto_kan as MK
    torch.manual_seed(0)
    model = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8).to(ctx.dev)
    gen = torch.Generator().manual_seed(400 + ctx.rank)
    x = torch.rand(B, 1, 28, 28, generator=gen).to(ctx.dev)
    y = torch.randint(0, 10, (B,), generator=gen).to(ctx.dev)
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, warmup=2, allreduce=ctx.world > 1)
    k = 20
    ms = ctx.timed(lambda: gstep(gstep.x, gstep.y), k, 3, flush=False, grad=True) / k
    ar_ms = None
    if ctx.world > 1:
        ar_ms = ctx.timed(lambda: ctx.dist.all_reduce(gopt.flat_grad), 20, 5, flush=False) / 20
    ach = 3.0 * 0.83e6 * B / (ms * 1e-3) / 1e12
    out = {"workload": "config 5: MNIST KuramotoKANClassifier(steps=10, nb=8), CE, fused AdamW (mnist_kuramoto_kan.py:228-283)",
           "rows_per_gpu": B, "ms_per_step": ms, "graph_replayed": True, "images_per_s": B * ctx.world / (ms * 1e-3),
           "allreduce_ms": ar_ms, "allreduce_bytes": int(gopt.flat_grad.numel()) * 4 if ctx.world > 1 else 0,
           "roofline": {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
                        "frac": ach / fp32_peak if fp32_peak else None, "flop_per_image": 0.83e6,
                        "convention": "SURVEY 8(d) count x 3; at 1024 images the step is ~90 launches of a few microseconds: latency-bound"}}
    del gstep, gopt
    return out


def inspan_fractions(x0, yT, state, D):
    """Share of layer inputs inside the knot span [-2.2, 2.2) (where the spline term is non-zero): layer 1 = the ODE
    state (at t = 0 and at the horizon), layer 2 = the hidden pre-activations of the last right-hand-side call (the
    kernel's hysteresis state holds them).  The second-generation kernel evaluates the spline cubic unconditionally
    (masked outside the span), so the timed work no longer depends on these fractions; they say how much of the
    SURVEY 8(d) spline FLOP count is 'live'."""
    f = lambda t: float(((t >= -2.2) & (t < 2.2)).float().mean())
    out = {"layer1_t0": f(x0), "layer1_horizon": f(yT)}
    if state is not None and state.shape[1] > D:
        out["layer2_horizon"] = f(state[:, D:])
    return out


def run_ours(args):
    import fetode
    from fetode import dist as fdist
    from fetode._lib import TABLEAU

    # libraries (NCCL's version banner) write to fd 1; keep stdout clean for the one JSON line
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    ctx = Ctx(args)
    rank, world, dev = ctx.rank, ctx.world, ctx.dev

    # ---------------- primary: config 2, 2^20 trajectories in total, sharded
    model = make_model().to(dev)
    plan = model.small_plan()
    assert plan is not None, "fused small-net kernel unavailable"
    total = args.trajectories
    lo, hi = fdist.shard_bounds(total, rank, world)
    B = hi - lo
    n_rk = args.rk_steps
    tab = TABLEAU["rk4_classic"]
    plan.image()                                                       # pack once (weights are static)
    func = fetode.as_rhs(model)

    def make_case(x_cpu):
        rows = x_cpu.shape[0]
        x0_host = x_cpu.contiguous().pin_memory()
        x0 = x0_host.to(dev)
        y = torch.empty_like(x0)
        state = torch.empty(rows, plan.state_floats, device=dev)
        out_host = torch.empty(rows, 2).pin_memory()

        def kernel_step():
            plan.rk_fwd(tab, x0, DT, n_rk, state=state, state_is_zero=True, y_out=y)

        def e2e_step():
            model.reset_state()
            xd = x0_host.to(dev, non_blocking=True)
            yT = fetode.integrate_rk4(func, xd, 0.0, n_rk * DT, n_rk)
            out_host.copy_(yT, non_blocking=True)
        return kernel_step, e2e_step, x0, y, state

    # every rank draws the SAME global batch and takes its slice, so the union over ranks is the N = 1 workload
    kernel_step, e2e_step, x0, y, state = make_case(synth_x0(total, seed=1000)[lo:hi])

    with ClockSampler(ctx.local) as clk:
        ms_kernel = ctx.timed(kernel_step, args.steps, args.warmup)
    finite = bool(torch.isfinite(y).all())
    inspan = inspan_fractions(x0, y, state, 2)
    ms_e2e = ctx.timed(e2e_step, args.steps, max(1, min(args.warmup, 2)))
    units_per_step = total * n_rk
    value = units_per_step * args.steps / (ms_kernel * 1e-3)
    e2e_value = units_per_step * args.steps / (ms_e2e * 1e-3)

    # ---------------- secondary: weak scaling (2^20 rows per GPU), N > 1 only
    weak = None
    if world > 1 and not args.no_weak:
        del kernel_step, e2e_step, x0, y, state
        torch.cuda.empty_cache()
        wk, _, wx0, wy, wstate = make_case(synth_x0(TRAJ_TOTAL, seed=1000 + rank))
        wms = ctx.timed(wk, 2, 1)
        weak = {"value": TRAJ_TOTAL * world * n_rk * 2 / (wms * 1e-3), "unit": UNIT, "trajectories_per_gpu": TRAJ_TOTAL,
                "ms_per_step": wms / 2, "steps": 2, "warmup": 1}
        del wk, wx0, wy, wstate
        torch.cuda.empty_cache()

    # ---------------- peaks + the training configurations (BASELINE metric: "fwd+bwd step ms")
    fp32_peak, tf32_peak = measure_peaks(ctx)
    fwd_bwd = None
    if not args.no_train:
        with_cpu = world == 1 and not args.no_cpu
        if with_cpu:
            torch.set_num_threads(os.cpu_count() or 1)
        fwd_bwd = {}
        for name, fn, a in (("lv_train_B1", lv_train_section, (1, fp32_peak, with_cpu)),
                            ("lv_train_B65536", lv_train_section, (65536, fp32_peak, with_cpu)),
                            ("lv_train_dopri5_B1", lv_dopri5_section, (1, with_cpu)),
                            ("ecg_train_B4096", ecg_train_section, (4096, tf32_peak, with_cpu)),
                            ("ett_encoder_B4096", ett_encoder_section, (4096, tf32_peak)),
                            ("mnist_train_B1024", mnist_train_section, (1024, fp32_peak))):
            try:
                fwd_bwd[name] = fn(ctx, *a)
            except Exception as e:                                   # noqa: BLE001 -- the primary number must still print
                fwd_bwd[name] = {"error": repr(e)[:300]}
            torch.cuda.empty_cache()

    if rank == 0:
        per_launch_ms = ms_kernel / args.steps
        peak_tf = fp32_peak if fp32_peak else 2 * 128 * 148 * 1.965e9 / 1e12
        ach_tf = FLOP_PER_TRAJ_STEP * B * n_rk / (per_launch_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_bytes = B * (8 + 8 + 4 * plan.state_floats)              # y0 in, y out, hysteresis state out
        traffic, traffic_src = None, None
        try:
            import csv
            prof = os.path.join(ROOT, "profiles", "r2_smallnet_rk2_split_ncu_full.csv")
            vals = {r[0]: (r[1], float(r[2])) for r in csv.reader(open(prof)) if len(r) == 3 and r[0].startswith("dram__bytes")}
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            traffic = sum(v * scale[u] for u, v in vals.values())
            traffic_src = "profiles/r2_smallnet_rk2_split_ncu_full.csv (dram read+write of one launch at 2^20 rows)"
        except Exception:
            pass
        roofline = {
            "bound": "fp32", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
            "traffic": traffic, "traffic_source": traffic_src,
            "peak_source": "live FFMA probe (fetode_probe_fp32_peak); MEASURED_PEAKS.json holds HBM/bf16 only" if fp32_peak
                           else "fallback 2*128*148*1.965 GHz",
            "kernel": "smallnet_rk2_kernel<2,10,10,5,FERRO2D,TP=1,64 threads,MINB=10,SPLIT=2>", "rows_per_launch": B,
            "flop_per_traj_step": FLOP_PER_TRAJ_STEP,
            "convention": "MODEL FLOP of SURVEY 8(d) (30 per branch basis, dense 8-coefficient spline); the kernel "
                          "executes fewer (factored exponentials, 4-term cubic): see profiles/ for executed-instruction rates",
            "inspan_fraction": inspan,
            "hbm": {"achieved_gbs": hbm_bytes / (per_launch_ms * 1e-3) / 1e9,
                    "peak_gbs": peaks.get("hbm_gbs"), "bytes_per_launch": hbm_bytes},
        }
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_launch_ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "trajectories_total": total, "trajectories_per_gpu": B,
                       "rk4_steps_per_step": n_rk, "l2": "flushed (256 MiB write) between timed iterations",
                       "weights": "random init, torch.manual_seed(0)",
                       "parallelism": f"dp{world} (batch-sharded, no collective on this path)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": total * 8, "d2h_bytes_per_step": total * 8,
                    "ms_per_step": ms_e2e / args.steps, "api": "fetode.integrate_rk4(as_rhs(KANFET), x0, 0, 14, 1000)"},
            "gpu_launches": args.steps * world,
            "outputs_finite": finite,
            "clocks": clk.summary(),
            "roofline": roofline,
            "peaks_measured_here": {"fp32_tflops": fp32_peak, "tf32_dense_tflops": tf32_peak},
        }
        if weak is not None:
            line["weak"] = weak
        if fwd_bwd is not None:
            line["fwd_bwd"] = fwd_bwd
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            layers = oracle_layers(model)
            n_traj, n_steps = CPU_SAMPLE
            cpu_reference_run(layers, n_traj, 1)                     # warm-up
            secs = min(cpu_reference_run(layers, n_traj, n_steps) for _ in range(2))
            line["cpu_baseline"] = {"value": n_traj * n_steps / secs, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{n_traj} trajectories x {n_steps} classic-RK4 steps (oracle port, torch CPU, no_grad)"}
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        import gc
        import signal
        gc.collect()
        torch.cuda.synchronize()
        signal.signal(signal.SIGALRM, lambda *_: os._exit(0))        # the line is printed: never hang in teardown
        signal.alarm(60)
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--trajectories", type=int, default=TRAJ_TOTAL, help="total over all ranks")
    ap.add_argument("--rk-steps", type=int, default=RK_STEPS)
    ap.add_argument("--no-train", action="store_true", help="skip the fwd_bwd block")
    ap.add_argument("--no-weak", action="store_true", help="skip the weak-scaling secondary measurement (N > 1)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU legs (N = 1)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
