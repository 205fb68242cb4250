"""CPU, world_size 2 over gloo: the data-parallel exchange of the train step (solver.FlatGradients):
every rank ends with the average of the per-rank gradients, clips identically and takes the same
NaN-guard branch (reference solver.py:178-182 semantics under SURVEY.md section 8e's DP scheme)."""
import importlib
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    solver = importlib.import_module("end-to-end-asr-pytorch_b200.solver")
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    fg = solver.FlatGradients(net.parameters())
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(4, 6, generator=g)
    fg.zero()
    (net(x).pow(2).sum() * (10.0 if rank == 1 else 1.0)).backward()
    local = fg.flat.clone()
    fg.allreduce_mean()
    norm = fg.clip_(0.5)
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    expect = sum(gathered) / world
    n_expect = expect.norm()
    expect_clipped = expect * torch.clamp(0.5 / (n_expect + 1e-6), max=1.0)
    views_ok = all(p.grad.data_ptr() >= fg.flat.data_ptr() and p.grad.data_ptr() < fg.flat.data_ptr() + fg.flat.numel() * 4
                   for p in net.parameters())
    q.put((rank, bool(torch.allclose(fg.flat, expect_clipped, atol=1e-6)), float(norm), float(n_expect), views_ok))
    dist.destroy_process_group()


def test_flat_gradient_allreduce_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _, _, _ in res), res
    assert all(v for *_, v in res)
    norms = [n for _, _, n, _, _ in res]
    assert abs(norms[0] - norms[1]) < 1e-7            # identical norm on every rank -> identical clip / NaN branch
    assert all(abs(n - e) < 1e-5 for _, _, n, e, _ in res)
