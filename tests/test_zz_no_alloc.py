"""The allocation-free CTC entry point (b2asr_ctc_loss_fwd_bwd_static, include/b2asr.h) against the existing
entry point b2asr_ctc_loss_fwd_bwd, both called through the C ABI with raw pointers.

The static variant keeps its scratch in a file-scope device array inside libb2asr.so that every call shares and
that keeps the previous call's contents, so the check runs on two calls in a row with DIFFERENT shapes (the second
call finds the first one's tables, rows and ticket word in the array).  Tolerance: the one tests/test_ctc_gpu.py
uses (1e-4 relative on nll/loss; gradient 1e-4 relative + 1e-4 of the row scale) -- the two entry points run the
same kernels, so the results are expected to agree far inside it.

This file is named so that pytest collects it last.
"""
import importlib

import pytest
import torch


def _lib():
    importlib.import_module("end-to-end-asr-pytorch_b200.build").build()
    return importlib.import_module("end-to-end-asr-pytorch_b200._lib")


def _case(B, T, V, Lmax, seed, dev):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(B, T, V, generator=g) * 2.0).to(dev)
    tl = torch.randint(max(1, Lmax // 2), Lmax + 1, (B,), generator=g)
    tl[0] = Lmax
    y = torch.randint(1, V, (B, Lmax), generator=g)
    y[:, 1::3] = y[:, 0::3][:, :y[:, 1::3].shape[1]]          # adjacent repeats
    il = torch.randint(2 * Lmax + 1, T + 1, (B,), generator=g)
    il[0] = T
    return x, y.to(dev), il.to(dev), tl.to(dev)


def _run(m, static, x, y, il, tl, weight):
    """One call through the C ABI (batch-first logits, fused log-softmax, 'mean'); -> nll, loss, grad."""
    L = m.lib()
    B, T, V = x.shape
    Lmax = y.shape[1]
    nll = torch.full((B,), float("nan"), device=x.device)
    loss = torch.full((1,), float("nan"), device=x.device)
    grad = torch.full_like(x, float("nan"))
    head = (x.data_ptr(), x.stride(1), x.stride(0), y.data_ptr(), y.stride(0), il.data_ptr(), tl.data_ptr(),
            T, B, V, Lmax, 0, 1, 1, weight, nll.data_ptr(), loss.data_ptr(), grad.data_ptr(), grad.stride(1),
            grad.stride(0))
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        rc = L.b2asr_ctc_loss_fwd_bwd_static(*head, stream)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
    else:
        nbytes = L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, 1)
        assert nbytes > 0
        ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        rc = L.b2asr_ctc_loss_fwd_bwd(*head, ws.data_ptr(), nbytes, stream)
        torch.cuda.synchronize()
        free0 = free1 = 0
    m.check(rc, "ctc static" if static else "ctc")
    return nll, loss, grad, free0 - free1


def _assert_same(a, b, tl, weight):
    nll_a, loss_a, g_a = a
    nll_b, loss_b, g_b = b
    assert torch.isfinite(nll_b).all() and torch.isfinite(g_b).all()
    assert torch.allclose(nll_a, nll_b, rtol=1e-4, atol=1e-4)
    assert torch.allclose(loss_a, loss_b, rtol=1e-4, atol=1e-5)
    B = g_a.shape[0]
    row_scale = (weight / (B * tl.clamp(min=1).float())).view(B, 1, 1)
    err = (g_a - g_b).abs()
    assert bool((err <= 1e-4 * g_b.abs() + 1e-4 * row_scale).all()), float(err.max())


@pytest.mark.gpu
def test_static_workspace_entry_point_matches_the_caller_workspace_one(cuda):
    m = _lib()
    w = 0.5
    # two calls in a row on different shapes: char-sized vocabulary (whole-vocabulary table), then a subword
    # vocabulary (gathered table) with shorter utterances; then the second shape again with the memory watermark read
    shapes = [(6, 120, 31, 24, 11), (3, 90, 1200, 17, 12), (3, 90, 1200, 17, 13)]
    leaked = None
    for B, T, V, Lmax, seed in shapes:
        x, y, il, tl = _case(B, T, V, Lmax, seed, cuda)
        want = _run(m, False, x, y, il, tl, w)
        got = _run(m, True, x, y, il, tl, w)
        _assert_same(got[:3], want[:3], tl, w)
        leaked = got[3]
    # third call: every kernel of the call has been loaded by then, so the device's free memory must not move
    assert leaked == 0, "the allocation-free entry point changed the device's free memory by %d bytes" % leaked


@pytest.mark.gpu
def test_static_workspace_entry_point_is_independent_of_leftovers(cuda):
    """Scribble over nothing we can reach -- instead run a big call, then a small one, then compare the small one
    with a fresh-workspace run: the leftovers of the big call must not leak into the small call's result."""
    m = _lib()
    big = _case(8, 200, 63, 40, 21, cuda)
    small = _case(2, 30, 9, 5, 22, cuda)
    _run(m, True, *big, 1.0)
    got = _run(m, True, *small, 1.0)
    want = _run(m, False, *small, 1.0)
    _assert_same(got[:3], want[:3], small[3], 1.0)


def test_static_workspace_entry_point_rejects_oversized_shapes():
    """No GPU needed: the size check comes before any CUDA call."""
    m = _lib()
    L = m.lib()
    T, B, V, Lmax = 3000, 64, 5000, 600
    assert L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, 1) > (32 << 20)
    rc = L.b2asr_ctc_loss_fwd_bwd_static(0, 0, 0, 0, 0, 0, 0, T, B, V, Lmax, 0, 1, 1, 1.0, 0, 0, 1, 0, 0, 0)
    assert rc == -3
    msg = L.b2asr_last_error().decode()
    assert "built-in array holds 33554432" in msg and "needs" in msg
    with pytest.raises(ValueError):
        m.check(rc, "b2asr_ctc_loss_fwd_bwd_static")
