"""GPU parity: the attention branch's sequence-loss kernel (csrc/ce.cu) vs oracle/ce_oracle.py and the golden
vectors the reference's own call sequence produced (tests/golden/seqloss.pt).

Tolerances (north_star: 1e-4 relative in fp32): scalar and per-row losses |a-b| <= 1e-4*|b| (+1e-6 absolute for rows
whose loss is ~0); gradient |a-b| <= 1e-4*|b| + 1e-4*row_scale with row_scale = upstream/(len_b*B), the largest
magnitude an entry of that row can take; argmax identical (ties: first index, as numpy/torch).
"""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import ce_oracle as ceo

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _mod():
    return importlib.import_module("end-to-end-asr-pytorch_b200.ce")


def make_case(B, L, V, seed, scale=3.0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, L, V, generator=g) * scale
    lens = [L] + [int(v) for v in torch.randint(1, L + 1, (B - 1,), generator=g)]
    y = torch.zeros(B, L + 1, dtype=torch.long)
    for b, n in enumerate(lens):
        y[b, 1:n] = torch.randint(2, V, (n - 1,), generator=g)
        y[b, n] = 1
    label = y[:, 1:L + 1].contiguous()
    return x, y, label, torch.sum(y != 0, dim=-1)


def check(x, label, tgt_len, upstream, cuda, prescale=None):
    ce = _mod()
    xd = x.to(cuda).requires_grad_(True)
    loss, rows, argmax = ce.seq_loss_mean(xd, label.to(cuda), tgt_len.to(cuda), 0,
                                          upstream if prescale is None else prescale, return_aux=True)
    (upstream * loss).backward()
    o_loss, o_rows, o_grad, o_argmax = ceo.seq_loss_mean(x.numpy(), label.numpy(), tgt_len.numpy(), 0, upstream)
    assert abs(float(loss) - o_loss) <= 1e-4 * abs(o_loss)
    assert np.all(np.abs(rows.cpu().numpy() - o_rows) <= 1e-4 * np.abs(o_rows) + 1e-6)
    assert np.array_equal(argmax.cpu().numpy(), o_argmax)
    B = x.shape[0]
    row_scale = (upstream / (tgt_len.numpy().astype(np.float64) * B))[:, None, None]
    err = np.abs(xd.grad.cpu().numpy() - o_grad)
    assert np.all(err <= 1e-4 * np.abs(o_grad) + 1e-4 * row_scale)
    assert np.all(xd.grad.cpu().numpy()[label.numpy() == 0] == 0.0)     # padded steps: exactly zero rows
    return float(loss)


@pytest.mark.parametrize("name", ["small", "wide"])
def test_matches_reference_golden(cuda, name):
    g = torch.load(os.path.join(G, "seqloss.pt"), weights_only=False)[name]
    ce = _mod()
    xd = g["att_pred"].to(cuda).requires_grad_(True)
    tgt_len = torch.sum(g["y"] != 0, dim=-1).to(cuda)
    loss, rows, argmax = ce.seq_loss_mean(xd, g["label"].to(cuda), tgt_len, 0, g["upstream"], return_aux=True)
    (g["upstream"] * loss).backward()
    assert abs(float(loss) - float(g["att_loss"])) <= 1e-4 * abs(float(g["att_loss"]))
    assert torch.allclose(rows.cpu(), g["rows"], rtol=1e-4, atol=1e-6)
    assert torch.equal(argmax.cpu(), g["argmax"])
    scale = (g["upstream"] / (tgt_len.cpu().double() * xd.shape[0]))[:, None, None]
    assert torch.all((xd.grad.cpu().double() - g["grad"].double()).abs() <= 1e-4 * g["grad"].double().abs() + 1e-4 * scale)
    pp = importlib.import_module("end-to-end-asr-pytorch_b200.postprocess")
    assert abs(pp.cal_acc(xd.detach(), g["label"].to(cuda), hyp=argmax) - g["acc"]) < 1e-9
    assert abs(pp.cal_acc(xd.detach(), g["label"].to(cuda)) - g["acc"]) < 1e-9


@pytest.mark.parametrize("B,L,V", [(1, 1, 3), (3, 7, 31), (5, 13, 63), (2, 9, 1024), (4, 11, 1025), (3, 5, 4999),
                                   (6, 20, 5000), (2, 3, 40000)])
def test_matches_oracle(cuda, B, L, V):
    x, y, label, tgt_len = make_case(B, L, V, seed=B * 1000 + V)
    check(x, label, tgt_len, 0.5, cuda)


def test_wrong_prescale_is_rescaled(cuda):
    x, y, label, tgt_len = make_case(4, 9, 5000, seed=3)
    check(x, label, tgt_len, 0.7, cuda, prescale=0.25)


def test_extreme_logits_and_ties(cuda):
    x, y, label, tgt_len = make_case(3, 6, 257, seed=9)
    x[0, 0] = 0.0                                         # all equal: argmax 0, loss log V
    x[0, 1, 100] = 80.0; x[0, 1, 7] = 80.0                # tie between two large entries
    x[1, 0] = -1.0e4; x[1, 0, 3] = 1.0e4                  # one-hot-like row
    x[2, 1, :] = torch.linspace(-60, 60, 257)
    check(x, label, tgt_len, 1.0, cuda)


def test_strided_logits_and_scalar_is_reproducible(cuda):
    ce = _mod()
    x, y, label, tgt_len = make_case(4, 8, 96, seed=5)
    big = torch.zeros(4, 8, 128)
    big[:, :, :96] = x
    view = big.to(cuda)[:, :, :96]                        # non-contiguous: row stride 128
    a = ce.seq_loss_mean(view, label.to(cuda), tgt_len.to(cuda))
    b = ce.seq_loss_mean(x.to(cuda), label.to(cuda), tgt_len.to(cuda))
    c = ce.seq_loss_mean(x.to(cuda), label.to(cuda), tgt_len.to(cuda))
    assert float(a) == float(b) == float(c)


def test_module_matches_torch_cross_entropy_none(cuda):
    """The reduction='none' module of solver.py:90 as validation calls it (solver.py:239), with autograd."""
    ce = _mod()
    x, y, label, tgt_len = make_case(3, 10, 500, seed=8)
    xd = x.view(-1, 500).to(cuda).requires_grad_(True)
    rows = ce.CrossEntropyLoss(ignore_index=0, reduction='none')(xd, label.view(-1).to(cuda))
    w = torch.linspace(0.5, 2.0, rows.numel(), device=cuda)
    (rows * w).sum().backward()
    x64 = x.view(-1, 500).double().requires_grad_(True)
    ref = torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none')(x64, label.view(-1))
    (ref * w.cpu().double()).sum().backward()
    assert torch.allclose(rows.cpu().double(), ref.detach(), rtol=1e-4, atol=1e-6)
    assert torch.allclose(xd.grad.cpu().double(), x64.grad, rtol=1e-4, atol=1e-6)


def test_cpu_tensor_is_refused(cuda):
    ce = _mod()
    x, y, label, tgt_len = make_case(2, 3, 8, seed=1)
    with pytest.raises(Exception):
        ce.seq_loss_mean(x, label, tgt_len)
