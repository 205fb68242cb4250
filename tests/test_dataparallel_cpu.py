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


# ---- Tester.exec: utterance-sharded decode, rank-0 ordered write (replaces the joblib fan-out of solver.py:342-367) ----
def _canned_tester(solver, ckpdir):
    """A Tester whose beam_decode is canned (no GPU): 7 utterances, n-best of 3, the reference's `gt<TAB>hyp` files."""
    t = object.__new__(solver.Tester)
    t.config = {'solver': {'decode_beam_size': 3}}
    t.paras = type('P', (), {'verbose': False})()
    t.ckpdir, t.decode_file = ckpdir, 'decode_beam_3_len_0.1'
    t.test_set = [([torch.zeros(1, 4, 2)], [torch.tensor([[0, i + 2, 1]])]) for i in range(7)]
    t.verbose = lambda msg: None

    class M:
        @staticmethod
        def translate(seq, return_string=True):
            return ' '.join(str(int(v)) for v in seq)

    t.mapper = M()

    def beam_decode(x, y, write=True):
        if y[1] == 5:
            raise ValueError('over-long utterance')                # one utterance the kernels refuse: must not lose the rest
        return (M.translate(y), ['hyp %d rank-independent #%d' % (y[1], k) for k in range(3)])

    t.beam_decode = beam_decode
    return t


def _tester_worker(rank, world, port, ckpdir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    solver = importlib.import_module("end-to-end-asr-pytorch_b200.solver")
    _canned_tester(solver, ckpdir).exec()
    dist.destroy_process_group()


def test_tester_exec_writes_identical_files_for_world_1_and_2(tmp_path):
    sys.path.insert(0, ROOT)
    solver = importlib.import_module("end-to-end-asr-pytorch_b200.solver")
    d1, d2 = str(tmp_path / "w1"), str(tmp_path / "w2")
    os.makedirs(d1); os.makedirs(d2)
    _canned_tester(solver, d1).exec()                              # world 1, this process
    ctx = mp.get_context("spawn")
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_tester_worker, args=(r, 2, port, d2)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for name in ('decode_beam_3_len_0.1.txt', 'decode_beam_3_len_0.1_nbest.txt'):
        a, b = open(os.path.join(d1, name), 'rb').read(), open(os.path.join(d2, name), 'rb').read()
        assert a == b and len(a) > 0, name
    best = open(os.path.join(d1, 'decode_beam_3_len_0.1.txt')).read().splitlines()
    assert len(best) == 7 and best[0] == '0 2 1\thyp 2 rank-independent #0'     # `gt<TAB>hyp`, utterance order
    assert best[3] == '0 5 1\t'                                                 # the refused utterance: empty hypothesis
    assert sorted(os.listdir(d2)) == sorted(os.listdir(d1))                     # the rank journals are gone
