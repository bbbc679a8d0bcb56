#!/usr/bin/env python
"""bench.py -- headline benchmark: Lotka-Volterra KAN-FET NODE inference ensemble.

Workload (BASELINE.json configs[1]): KANFET([2,10,2], grid 5) right-hand side, classic RK4,
2^20 synthetic initial conditions x 1000 steps (dt = 0.014) PER GPU (weak scaling: every rank
integrates its own shard, no data-path collective).  One "step" of this script = one pass of the
hot path over one batch = one launch of the fused kernel = 2^20 x 1000 trajectory-steps.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun ... bench.py --gpus N ...      (one rank per GPU)

Prints ONE JSON line (rank 0).  `value` is device-timed with inputs resident in HBM; `e2e` is the same
metric through the public API (fetode.integrate_rk4) with the initial conditions copied from pinned
host memory and the final states copied back inside the timed region.
`--impl reference` times the CPU oracle (the reference's PyTorch arithmetic restated in oracle/,
bit-identical to the live reference) on the host cores on a bounded sample of the same workload.
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
TRAJ_PER_GPU = 1 << 20
FLOP_PER_TRAJ_STEP = 22712          # KANFET([2,10,2]) ferro branch, classic RK4
METRIC = "lv_kanfet_node_traj_steps_per_s"
UNIT = "traj-steps/s"


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


def oracle_layers(model):
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()
          if not k.endswith("prev_x") and not k.endswith("branch_sign")}
    return [{k[len(f"layers.{i}."):]: v for k, v in sd.items() if k.startswith(f"layers.{i}.")}
            for i in range(len(WIDTHS) - 1)]


def synth_x0(n, seed):
    gen = torch.Generator().manual_seed(seed)
    return torch.rand(n, 2, generator=gen) * 6.75 + 0.25


# ------------------------------------------------------------------------------------------
# CPU leg (oracle port of the reference's PyTorch path)
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


def cpu_sample_size(layers, budget_s):
    """Pick (n_traj, n_steps) so one sample costs about budget_s seconds on this host."""
    n_traj = 16384
    t = cpu_reference_run(layers, n_traj, 1)            # also the warm-up
    t = min(t, cpu_reference_run(layers, n_traj, 1))
    n_steps = max(1, min(50, int(budget_s / max(t, 1e-4))))
    return n_traj, n_steps


def run_reference(args):
    rank, world, _ = _dist_env()
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    layers = oracle_layers(make_model())
    total_budget = 100.0
    per_step = total_budget / max(1, args.steps + args.warmup)
    n_traj, n_steps = cpu_sample_size(layers, min(per_step, 20.0))
    for _ in range(args.warmup):
        cpu_reference_run(layers, n_traj, n_steps)
    secs = [cpu_reference_run(layers, n_traj, n_steps, seed=i) for i in range(args.steps)]
    tot = sum(secs)
    value = n_traj * n_steps * args.steps / tot
    sample = f"{n_traj} trajectories x {n_steps} classic-RK4 steps per step (oracle port, torch CPU, no_grad)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "LV KANFET([2,10,2]) inference ensemble, classic RK4, dt=0.014 (BASELINE configs[1])",
                   "sample": sample},
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
def run_ours(args):
    import torch.distributed as dist

    import fetode
    from fetode._lib import TABLEAU, lib

    rank, world, local = _dist_env()
    # libraries (NCCL's version banner) write to fd 1; keep stdout clean for the one JSON line
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun --nproc-per-node N for --gpus N > 1")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    model = make_model().to(dev)
    plan = model.small_plan()
    assert plan is not None, "fused small-net kernel unavailable"
    B = args.traj_per_gpu
    n_rk = args.rk_steps
    x0_host = synth_x0(B, seed=1000 + rank).pin_memory()
    x0 = x0_host.to(dev)
    y = torch.empty_like(x0)
    state = torch.empty(B, plan.state_floats, device=dev)
    out_host = torch.empty(B, 2).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    tab = TABLEAU["rk4_classic"]
    plan.image()                                                       # pack once (weights are static)

    def kernel_step():
        plan.rk_fwd(tab, x0, DT, n_rk, state=state, state_is_zero=True, y_out=y)

    func = fetode.as_rhs(model)

    def e2e_step():
        model.reset_state()
        xd = x0_host.to(dev, non_blocking=True)
        yT = fetode.integrate_rk4(func, xd, 0.0, n_rk * DT, n_rk)
        out_host.copy_(yT, non_blocking=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, k, w):
        with torch.no_grad():
            for _ in range(w):
                step_fn()
            barrier()
            evs = []
            t_wall0 = time.perf_counter()
            for _ in range(k):
                flush.zero_()                                            # evict L2 between timed iterations
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                step_fn()
                e1.record()
                evs.append((e0, e1))
            barrier()
            wall = time.perf_counter() - t_wall0
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, wall

    with ClockSampler(local) as clk:
        ms_kernel, _ = timed(kernel_step, args.steps, args.warmup)
    ms_e2e, _ = timed(e2e_step, args.steps, max(1, min(args.warmup, 2)))

    units_per_step = B * n_rk * world
    value = units_per_step * args.steps / (ms_kernel * 1e-3)
    e2e_value = units_per_step * args.steps / (ms_e2e * 1e-3)

    if rank == 0:
        # roofline of the dominant (only) kernel: FP32 CUDA-core pipe, peak measured live by an FFMA probe
        import ctypes as C
        tf = C.c_double(0.0)
        rc = lib().fetode_probe_fp32_peak(C.byref(tf), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        peak_tf = tf.value if rc == 0 and tf.value > 0 else 2 * 128 * 148 * 1.965e9 / 1e12
        ach_tf = FLOP_PER_TRAJ_STEP * B * n_rk / (ms_kernel / args.steps * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_bytes = B * (8 + 8 + 4 * plan.state_floats)              # y0 in, y out, hysteresis state out
        # DRAM bytes per launch of the same kernel from the committed `ncu --set full` capture (profiles/)
        traffic, traffic_src = None, None
        try:
            import csv
            prof = os.path.join(ROOT, "profiles", "r1_smallnet_rk_ncu_full.csv")
            vals = {r[0]: (r[1], float(r[2])) for r in csv.reader(open(prof)) if len(r) == 3 and r[0].startswith("dram__bytes")}
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            traffic = sum(v * scale[u] for u, v in vals.values())
            traffic_src = "profiles/r1_smallnet_rk_ncu_full.csv (dram read+write of one launch at this size; mostly register-spill lines)"
        except Exception:
            pass
        roofline = {
            "bound": "fp32", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
            "traffic": traffic, "traffic_source": traffic_src,
            "peak_source": "live FFMA probe (fetode_probe_fp32_peak); MEASURED_PEAKS.json holds HBM/bf16 only" if rc == 0
                           else "fallback 2*128*148*1.965 GHz",
            "kernel": "smallnet_rk_kernel<2,10,5,FERRO2D>",
            "flop_per_traj_step": FLOP_PER_TRAJ_STEP,
            "hbm": {"achieved_gbs": hbm_bytes / (ms_kernel / args.steps * 1e-3) / 1e9,
                    "peak_gbs": peaks.get("hbm_gbs"), "bytes_per_launch": hbm_bytes},
        }
        # CPU baseline: oracle port on a bounded sample (about 10-20 s)
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        layers = oracle_layers(model)
        n_traj, n_steps = cpu_sample_size(layers, 6.0)
        secs = min(cpu_reference_run(layers, n_traj, n_steps) for _ in range(2))
        cpu_val = n_traj * n_steps / secs
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_kernel / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "LV KANFET([2,10,2]) inference ensemble, classic RK4, dt=0.014 (BASELINE configs[1])",
                       "trajectories_per_gpu": B, "rk4_steps_per_step": n_rk, "l2": "flushed (256 MiB write) between timed iterations",
                       "weights": "random init, torch.manual_seed(0)", "parallelism": f"dp{world} (batch-sharded, no collective)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * 8 * world, "d2h_bytes_per_step": B * 8 * world,
                    "ms_per_step": ms_e2e / args.steps, "api": "fetode.integrate_rk4(as_rhs(KANFET), x0, 0, 14, 1000)"},
            "gpu_launches": args.steps,
            "clocks": clk.summary(),
            "roofline": roofline,
            "cpu_baseline": {"value": cpu_val, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{n_traj} trajectories x {n_steps} classic-RK4 steps (oracle port, torch CPU, no_grad)"},
        }
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--traj-per-gpu", type=int, default=TRAJ_PER_GPU)
    ap.add_argument("--rk-steps", type=int, default=RK_STEPS)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
