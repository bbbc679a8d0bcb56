#!/usr/bin/env python
"""Data-parallel training throughput of the BASELINE.json training configurations (3: ECG, 5: MNIST-Kuramoto) at
N GPUs: one process per GPU (torchrun), full parameter replica, fixed rows per GPU (weak scaling), ONE flat fp32
NCCL all-reduce of the parameter gradients per optimiser step (fetode.dist.allreduce_gradients), clipping after it.

    python bench_train_dp.py                      # N = 1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench_train_dp.py

One JSON line per configuration (rank 0): device-timed step (CUDA events, max over ranks), samples/s over all ranks.
`python bench.py` stays the driver's contract; this script is the scaling evidence for the training path.
"""
import gc
import json
import os
import signal
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import fetode  # noqa: E402
from fetode import dist as fdist, ecg, mnist_kuramoto_kan as MK  # noqa: E402


def timed(step, world, dev, iters=5, warm=3):
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)                                   # NCCL's banner goes to stderr, JSON lines to the real stdout
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lines = []

    # ---- config 5: MNIST-shaped Kuramoto-KAN classifier, 1024 images per GPU
    torch.manual_seed(0)
    model = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8).to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    B = 1024
    gen = torch.Generator().manual_seed(100 + rank)
    x = torch.rand(B, 1, 28, 28, generator=gen).to(dev)
    y = torch.randint(0, 10, (B,), generator=gen).to(dev)

    def mnist_step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x), y).backward()
        fdist.allreduce_gradients(model)
        opt.step()
    ms = timed(mnist_step, world, dev)
    # the same step (all-reduce included) captured into one CUDA graph per rank and replayed
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, allreduce=world > 1, warmup=2)
    gms = timed(lambda: gstep(gstep.x, gstep.y), world, dev)
    lines.append({"config": "5: MNIST KuramotoKANClassifier(steps=10, nb=8), CE, AdamW", "n_gpus": world,
                  "rows_per_gpu": B, "ms_per_step": ms, "images_per_s": B * world / ms * 1e3,
                  "graph_ms_per_step": gms, "graph_images_per_s": B * world / gms * 1e3,
                  "allreduce_floats": sum(p.numel() for p in model.parameters())})
    del gstep, gopt

    # ---- config 3: ECG KanFet_MLP_NODE, classic RK4 x 8, 4096 series per GPU
    torch.manual_seed(0)
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    model = model.to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    B = 4096
    x = torch.randn(B, 96, generator=gen)
    x = ((x - x.mean(1, keepdim=True)) / x.std(1, keepdim=True)).to(dev)
    y = torch.randint(0, 2, (B,), generator=gen).to(dev)

    def ecg_step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x), y).backward()
        fdist.allreduce_gradients(model)
        fdist.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
    ms = timed(ecg_step, world, dev, iters=3, warm=2)
    gopt = fetode.optim.AdamW(model.parameters(), lr=1e-3, capturable=True)
    gstep = fetode.graphs.GraphedTrainStep(model, gopt, F.cross_entropy, x, y, max_norm=1.0, allreduce=world > 1, warmup=2)
    gms = timed(lambda: gstep(gstep.x, gstep.y), world, dev, iters=5, warm=2)
    lines.append({"config": "3: ECG KanFet_MLP_NODE(T=96, latent 64, nb 12, hidden 128), classic RK4 x 8, CE, AdamW, clip 1.0",
                  "n_gpus": world, "rows_per_gpu": B, "ms_per_step": ms, "sample_steps_per_s": B * 8 * world / ms * 1e3,
                  "graph_ms_per_step": gms, "graph_sample_steps_per_s": B * 8 * world / gms * 1e3,
                  "allreduce_floats": sum(p.numel() for p in model.parameters())})
    if rank == 0:
        sys.stdout.flush()
        os.dup2(saved, 1)
        for ln in lines:
            print(json.dumps(ln), flush=True)
    # graphs that captured NCCL kernels must be gone before the communicator is torn down
    del gstep, gopt
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        signal.signal(signal.SIGALRM, lambda *_: os._exit(0))     # results are printed: never hang in teardown
        signal.alarm(30)
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
