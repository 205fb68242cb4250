"""GPU parity of the one-launch optimizer step (csrc/optim.cu, _optim.Adadelta) with torch.optim.Adadelta -- the
optimizer of reference src/solver.py:101-106 -- and with the numpy oracle, through the C ABI."""
import copy
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
pytestmark = pytest.mark.gpu

SHAPES = [(31,), (1280, 120), (5,), (2560, 640), (3, 7, 11), (40000,)]     # odd sizes: padded flat offsets, tails


def _setup(dev, seed=0):
    torch.manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, device=dev)) for s in SHAPES]


@pytest.mark.parametrize("weight_decay", [0.0, 1e-3])
def test_fused_adadelta_matches_torch_and_oracle(weight_decay):
    solver = importlib.import_module(PKG + ".solver")
    optim = importlib.import_module(PKG + "._optim")
    from oracle import optim_oracle as oo
    dev = torch.device("cuda:0")
    mine, ref = _setup(dev), _setup(dev)
    fg = solver.FlatGradients(mine)
    a = optim.Adadelta(mine, lr=1.0, eps=1e-8, weight_decay=weight_decay)
    a.bind(fg)
    b = torch.optim.Adadelta(ref, lr=1.0, eps=1e-8, weight_decay=weight_decay)
    cpu = [(p.detach().cpu().numpy().copy(), np.zeros(p.shape, np.float32), np.zeros(p.shape, np.float32)) for p in ref]
    for step in range(5):
        torch.manual_seed(100 + step)
        grads = [torch.randn(*s, device=dev) * (10.0 ** (step - 2)) for s in SHAPES]
        fg.zero()
        for p, q, g in zip(mine, ref, grads):
            p.grad.copy_(g)                                   # the flat views stay attached
            q.grad = g.clone()
        assert a._flat_ok()
        a.step()
        b.step()
        for (w, sq, acc), g in zip(cpu, grads):
            oo.adadelta_step(w, g.cpu().numpy(), sq, acc, lr=1.0, rho=0.9, eps=1e-8, weight_decay=weight_decay)
    torch.cuda.synchronize()
    for p, q, (w, sq, acc) in zip(mine, ref, cpu):
        torch.testing.assert_close(p, q, rtol=2e-6, atol=1e-7)
        torch.testing.assert_close(a.state[p]["square_avg"], b.state[q]["square_avg"], rtol=2e-6, atol=0)
        torch.testing.assert_close(a.state[p]["acc_delta"], b.state[q]["acc_delta"], rtol=4e-6, atol=1e-30)
        np.testing.assert_allclose(p.detach().cpu().numpy(), w, rtol=2e-6, atol=1e-7)
        assert float(a.state[p]["step"]) == float(b.state[q]["step"]) == 5.0


def test_state_dict_round_trips_with_the_stock_optimizer():
    solver = importlib.import_module(PKG + ".solver")
    optim = importlib.import_module(PKG + "._optim")
    dev = torch.device("cuda:0")
    mine, ref = _setup(dev, 1), _setup(dev, 1)
    fg = solver.FlatGradients(mine)
    a = optim.Adadelta(mine, lr=1.0, eps=1e-8)
    a.bind(fg)
    b = torch.optim.Adadelta(ref, lr=1.0, eps=1e-8)

    def one(step):
        torch.manual_seed(7 + step)
        grads = [torch.randn(*s, device=dev) for s in SHAPES]
        for p, q, g in zip(mine, ref, grads):
            p.grad.copy_(g)
            q.grad = g.clone()
        a.step()
        b.step()

    one(0)
    one(1)
    # fused -> stock and stock -> fused through state_dict, then both keep stepping identically
    # (state_dict() hands out references; a checkpoint written with torch.save holds copies)
    sd_a, sd_b = copy.deepcopy(a.state_dict()), copy.deepcopy(b.state_dict())
    b.load_state_dict(sd_a)
    a.load_state_dict(sd_b)
    assert a._flat_ok()
    for p in mine:
        assert a.state[p]["square_avg"].data_ptr() >= a._sq.data_ptr()          # state is a flat view again
        assert a.state[p]["square_avg"].data_ptr() < a._sq.data_ptr() + 4 * a._sq.numel()
    one(2)
    torch.cuda.synchronize()
    for p, q in zip(mine, ref):
        torch.testing.assert_close(p, q, rtol=2e-6, atol=1e-7)
        torch.testing.assert_close(a.state[p]["acc_delta"], b.state[q]["acc_delta"], rtol=4e-6, atol=1e-30)


def test_replaced_gradients_take_the_stock_path():
    solver = importlib.import_module(PKG + ".solver")
    optim = importlib.import_module(PKG + "._optim")
    dev = torch.device("cuda:0")
    mine, ref = _setup(dev, 2), _setup(dev, 2)
    fg = solver.FlatGradients(mine)
    a = optim.Adadelta(mine, lr=1.0, eps=1e-8)
    a.bind(fg)
    b = torch.optim.Adadelta(ref, lr=1.0, eps=1e-8)
    torch.manual_seed(9)
    for p, q in zip(mine, ref):
        g = torch.randn_like(p)
        p.grad = g.clone()                                    # NOT the flat view
        q.grad = g.clone()
    assert not a._flat_ok()
    a.step()
    b.step()
    for p, q in zip(mine, ref):
        torch.testing.assert_close(p, q, rtol=0, atol=0)
    a.zero_grad()                                             # re-attaches the flat views
    assert a._flat_ok()
