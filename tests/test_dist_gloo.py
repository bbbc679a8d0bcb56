"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: row sharding, the single flat gradient all-reduce
(N ranks reproduce the 1-rank gradient of a mean loss, equal and unequal shards), the dopri5 norm exchange."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_rows, out_dir):
    for p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fetode import dist as fd
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 2))
    model[2].bias.requires_grad_(False)                       # a frozen parameter must be skipped
    gen = torch.Generator().manual_seed(1)
    x, y = torch.randn(n_rows, 6, generator=gen), torch.randn(n_rows, 2, generator=gen)
    lo, hi = fd.shard_bounds(n_rows)
    xs = fd.shard_rows(x)
    assert xs.shape[0] == hi - lo
    loss = ((model(xs) - y[lo:hi]) ** 2).mean()
    loss.backward()
    fd.allreduce_gradients(model, local_rows=hi - lo, global_rows=n_rows)
    pair = torch.tensor([float(rank + 1), float(hi - lo)], dtype=torch.float64)
    fd.norm_sync(pair)
    torch.save({"grads": [p.grad.clone() for p in model.parameters() if p.requires_grad], "pair": pair,
                "bounds": (lo, hi)}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_rows", [8, 7])
def test_two_ranks_reproduce_single_rank_gradient(tmp_path, n_rows):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_rows, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 2))
    model[2].bias.requires_grad_(False)
    gen = torch.Generator().manual_seed(1)
    x, y = torch.randn(n_rows, 6, generator=gen), torch.randn(n_rows, 2, generator=gen)
    ((model(x) - y) ** 2).mean().backward()
    want = [p.grad for p in model.parameters() if p.requires_grad]
    for r in range(world):
        for g, w in zip(res[r]["grads"], want):
            assert torch.allclose(g, w, rtol=1e-5, atol=1e-7)
        assert res[r]["pair"].tolist() == [3.0, float(n_rows)]
    assert res[0]["bounds"][0] == 0 and res[0]["bounds"][1] == res[1]["bounds"][0] and res[1]["bounds"][1] == n_rows


def test_shard_bounds_cover_everything():
    sys.path.insert(0, os.path.join(ROOT, "fet-ode_b200"))
    from fetode import dist as fd
    for n in (0, 1, 5, 8, 1 << 20):
        for w in (1, 2, 3, 8):
            b = [fd.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(hi - lo for lo, hi in b) - min(hi - lo for lo, hi in b) <= 1


def _sync_worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fetode import dist as fd
    from fetode import ecg
    # a hard-gated basis whose state row differs per rank: after the sync every rank holds the LAST rank's row
    m = torch.nn.Sequential(ecg.LogisticBasis(5, 3))
    m[0]._row = torch.full((5,), float(rank + 1))
    fd.hardgate_prev_sync(m)
    assert fd.enable_hardgate_sync(m) == 1 and m[0]._dist_sync
    x = torch.arange(7 * 2, dtype=torch.float32).reshape(7, 2)
    rows, n = fd.shard_rows_with_sentinel(x)
    torch.save({"row": m[0]._row.clone(), "rows": rows.clone(), "n": n}, os.path.join(out_dir, f"s{rank}.pt"))
    dist.destroy_process_group()


def test_hardgate_row_sync_and_sentinel_rows(tmp_path):
    world = 2
    mp.spawn(_sync_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(tmp_path, f"s{r}.pt")) for r in range(world)]
    x = torch.arange(7 * 2, dtype=torch.float32).reshape(7, 2)
    for r in range(world):
        assert torch.equal(res[r]["row"], torch.full((5,), float(world)))        # the last rank's row everywhere
    assert res[0]["n"] == 4 and res[1]["n"] == 3
    assert torch.equal(res[0]["rows"], torch.cat([x[:4], x[-1:]]))                # rank 0: its rows + the global last row
    assert torch.equal(res[1]["rows"], x[4:])                                     # the last rank already ends on it
