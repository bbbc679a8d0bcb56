"""N ranks == 1 rank on the REAL fetode.ecg modules (SURVEY section 4 item iv, section 8e): the hard-gated basis gates
on the last row of the GLOBAL batch (train_ecg_kan_fet_nn_ode.py:129-132), so a batch-sharded run must either carry
that row on every rank (sentinel row, zero communication) or broadcast it after every call.  Both are checked against
the single-rank run: losses of two consecutive steps (the row is carried across steps), all-reduced gradients of
every parameter, and the hysteresis rows left behind, to 1e-5.  Needs 2 GPUs (NCCL); skipped otherwise."""
import copy
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_steps(model, batches, shard):
    """Two forward/backward passes without a parameter update; returns (losses, flat gradient, rows)."""
    from fetode import dist as fd
    from fetode import ecg
    model.zero_grad()
    losses = []
    for x, y in batches:
        if shard is None:
            loss = F.cross_entropy(model(x), y)
            loss.backward()
            losses.append(loss.detach().double())
            continue
        lo, hi = fd.shard_bounds(x.shape[0])
        if shard == "sentinel":
            xa, n = fd.shard_rows_with_sentinel(x)
            logits = model(xa)[:n]
        else:
            logits = model(x[lo:hi])
        # mean over the GLOBAL batch = sum over ranks of (local sum / global rows)
        loss = F.cross_entropy(logits, y[lo:hi], reduction="sum") / x.shape[0]
        loss.backward()
        tot = loss.detach().double().clone()
        dist.all_reduce(tot)
        losses.append(tot)
    if shard is not None:
        fd.allreduce_gradients(model, average=False)
    flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in model.parameters()])
    rows = [m._row.clone() for m in model.modules() if isinstance(m, ecg.LogisticBasis)]
    return torch.stack(losses), flat, rows


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from fetode import dist as fd
    from fetode import ecg
    torch.manual_seed(0)
    base = ecg.KanFet_MLP_NODE(T=24, num_classes=2, latent_dim=16, num_basis=6, ode_hidden=32, dropout=0.0)
    base.ode_method = "rk4_classic"
    with torch.no_grad():
        base.odefunc.out.weight.mul_(100.0)
    gen = torch.Generator().manual_seed(1)
    batches = [(torch.randn(51, 24, generator=gen).to(dev), torch.randint(0, 2, (51,), generator=gen).to(dev))
               for _ in range(2)]                                    # 51 rows: unequal shards (26 / 25)
    res = {}
    res["single"] = _run_steps(copy.deepcopy(base).to(dev), batches, None)
    res["sentinel"] = _run_steps(copy.deepcopy(base).to(dev), batches, "sentinel")
    m = copy.deepcopy(base).to(dev)
    assert fd.enable_hardgate_sync(m) == 2
    res["broadcast"] = _run_steps(m, batches, "broadcast")
    res["unsynced"] = _run_steps(copy.deepcopy(base).to(dev), batches, "none")
    torch.save({k: (v[0].cpu(), v[1].cpu(), [r.cpu() for r in v[2]]) for k, v in res.items()},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def test_two_ranks_equal_one_rank_on_ecg_model(tmp_path):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    for r in range(world):
        l1, g1, rows1 = res[r]["single"]
        for mode in ("sentinel", "broadcast"):
            l, g, rows = res[r][mode]
            assert _rel(l, l1) < 1e-5, (r, mode, l, l1)
            assert _rel(g, g1) < 1e-5, (r, mode)
            for a, b in zip(rows, rows1):
                assert _rel(a, b) < 1e-5, (r, mode)
    # without either, rank 0 gates on its own last row: a different function (this is what the sync is for)
    lu, gu, _ = res[0]["unsynced"]
    assert _rel(gu, res[0]["single"][1]) > 1e-5
