"""GPU parity: kernel (1) CTC loss + gradient vs the oracle (oracle/ctc_oracle.py).

Tolerances (north_star: 1e-4 relative in fp32):
  * loss / per-utterance nll: |a-b| <= 1e-4*|b| against torch CTCLoss fp32 (the reference's
    own call sequence) AND against its float64 evaluation;
  * gradient: |a-b| <= 1e-4*|b| + 1e-4*row_scale against the float64 evaluation of the
    reference call, where row_scale = upstream/(B*target_len_b) is the largest magnitude a
    gradient entry of that utterance can take (softmax and occupancy both lie in [0,1]).
    The fp32 ATen kernel the reference runs is itself 1e-3..2e-2*row_scale away from float64
    at T>=300 (fp32 log-space cancellation; measured with a one-off script in round 1), so float64 is the
    judge; against torch fp32 we only require that we are no further from it than it is from
    float64 (x4).
"""
import importlib

import numpy as np
import pytest
import torch

from oracle import ctc_oracle as co

pytestmark = pytest.mark.gpu


def _mod():
    return importlib.import_module("end-to-end-asr-pytorch_b200.ctc")


def make_case(B, T, V, Lmax, seed, repeats=0.1, full_len=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, T, V, generator=g) * 2.0
    tl = torch.randint(max(1, Lmax // 2), Lmax + 1, (B,), generator=g)
    tl[0] = Lmax
    y = torch.randint(1, V, (B, Lmax), generator=g)
    rep = torch.rand(B, Lmax, generator=g) < repeats
    for i in range(1, Lmax):
        y[:, i] = torch.where(rep[:, i], y[:, i - 1], y[:, i])
    need = tl + (y[:, 1:] == y[:, :-1]).sum(-1)            # generous feasibility bound
    il = torch.empty(B, dtype=torch.long)
    for b in range(B):
        lo = min(T, int(need[b]) + 1)
        il[b] = T if full_len else int(torch.randint(lo, T + 1, (1,), generator=g))
    il[0] = T
    for b in range(B):
        y[b, tl[b]:] = 0
    return x, y, il, tl


CASES = [
    # B, T,   V,   Lmax
    (3, 12, 6, 4),
    (4, 50, 31, 20),
    (16, 100, 63, 30),       # timit-shaped
    (8, 400, 31, 150),       # libri char-shaped (M2)
    (4, 200, 5000, 40),      # subword vocabulary -> gathered table
    (2, 300, 1500, 149),     # V > 1024 but whole-vocabulary table
    (1, 7, 3, 3),
    (5, 33, 9, 16),          # L close to T/2
]


@pytest.mark.parametrize("B,T,V,Lmax", CASES)
@pytest.mark.parametrize("pair", [1, 0])
def test_fused_logits_matches_reference_call(cuda, monkeypatch, B, T, V, Lmax, pair):
    monkeypatch.setenv("B2ASR_CTC_PAIR", str(pair))
    m = _mod()
    x, y, il, tl = make_case(B, T, V, Lmax, seed=B * 1000 + T)
    w = 0.5
    # reference call sequence, fp32 and fp64
    loss32, nll32, g32 = co.reference_call(x, y, il.tolist(), tl, torch.float32, weight=w)
    loss64, nll64, g64 = co.reference_call(x, y, il.tolist(), tl, torch.float64, weight=w)
    xd = x.to(cuda).requires_grad_(True)
    loss, nll = m.ctc_loss(xd, y.to(cuda), il.tolist(), tl.to(cuda), from_logits=True, batch_first=True,
                           grad_prescale=w, return_nll=True)
    (w * loss).backward()
    torch.cuda.synchronize()
    nll_d, g_d = nll.cpu().double(), xd.grad.cpu().double()
    assert torch.allclose(nll_d, nll64, rtol=1e-4, atol=1e-4), (nll_d, nll64)
    assert torch.allclose(nll_d, nll32.double(), rtol=1e-4, atol=1e-3)
    assert abs(float(loss.detach()) - float(loss64)) <= 1e-4 * abs(float(loss64)) + 1e-5
    scale = w / (B * tl.clamp_min(1).double())
    err = (g_d - g64).abs()
    tol = 1e-4 * g64.abs() + 1e-4 * scale.view(B, 1, 1)
    worst = float((err / tol).max())
    ref_gap = float((g32.double() - g64).abs().max())
    assert worst <= 1.0, "grad mismatch vs float64 oracle: worst err/tol=%.2f max|err|=%.3e (torch fp32 gap %.3e)" % (
        worst, float(err.max()), ref_gap)
    assert float((g_d - g32.double()).abs().max()) <= 4 * ref_gap + 1e-6 * float(scale.max())
    # rows past in_len are exactly zero
    for b in range(B):
        assert float(g_d[b, int(il[b]):].abs().sum()) == 0.0


def test_log_probs_api_is_torch_compatible(cuda):
    m = _mod()
    x, y, il, tl = make_case(4, 60, 40, 12, seed=7)
    lp = torch.log_softmax(x.transpose(0, 1), -1).contiguous()
    loss_ref, nll_ref, g_ref = co.torch_ctc_from_log_probs(lp, y, il, tl, torch.float64)
    lpd = lp.to(cuda).requires_grad_(True)
    crit = m.CTCLoss(blank=0, reduction="mean")
    loss = crit(lpd, y.to(cuda), torch.LongTensor(il.tolist()), tl.to(cuda))   # CPU input_lengths as solver.py:160
    loss.backward()
    assert abs(float(loss) - float(loss_ref)) <= 1e-4 * abs(float(loss_ref))
    err = (lpd.grad.cpu().double() - g_ref).abs()
    assert float(err.max()) <= 1e-4 * float(g_ref.abs().max()) + 1e-7
    # same thing through a transposed (non-contiguous) view, as the reference builds it
    xb = x.to(cuda).requires_grad_(True)
    loss2 = crit(torch.log_softmax(xb.transpose(0, 1), -1), y.to(cuda), torch.LongTensor(il.tolist()), tl.to(cuda))
    loss2.backward()
    _, _, g_logits = co.reference_call(x, y, il.tolist(), tl, torch.float64)
    assert float((xb.grad.cpu().double() - g_logits).abs().max()) <= 1e-4 * float(g_logits.abs().max()) + 1e-7


def test_upstream_other_than_prescale_is_rescaled(cuda):
    m = _mod()
    x, y, il, tl = make_case(3, 40, 20, 8, seed=11)
    _, _, g64 = co.reference_call(x, y, il.tolist(), tl, torch.float64, weight=0.3)
    xd = x.to(cuda).requires_grad_(True)
    loss = m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True, grad_prescale=1.0)
    (0.3 * loss).backward()
    assert float((xd.grad.cpu().double() - g64).abs().max()) <= 1e-4 * float(g64.abs().max()) + 1e-7


def test_reduction_none_and_sum(cuda):
    m = _mod()
    x, y, il, tl = make_case(5, 30, 12, 6, seed=3)
    lp = torch.log_softmax(x.transpose(0, 1), -1)
    for red in ("none", "sum"):
        ref_in = lp.double().clone().requires_grad_(True)
        ref = torch.nn.functional.ctc_loss(ref_in, y, il, tl, blank=0, reduction=red)
        ref.sum().backward()
        d = lp.to(cuda).contiguous().requires_grad_(True)
        out = m.ctc_loss(d, y.to(cuda), il, tl, reduction=red)
        out.sum().backward()
        assert torch.allclose(out.cpu().double(), ref.detach(), rtol=1e-4, atol=1e-5)
        assert float((d.grad.cpu().double() - ref_in.grad).abs().max()) <= 1e-4 * float(ref_in.grad.abs().max()) + 1e-7


def test_edge_cases_empty_target_infeasible_and_loss_only(cuda):
    m = _mod()
    T, B, V = 6, 4, 5
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, T, V, generator=g)
    y = torch.tensor([[1, 1, 2, 3], [2, 3, 0, 0], [0, 0, 0, 0], [4, 4, 4, 4]])
    tl = torch.tensor([4, 2, 0, 4])
    il = torch.tensor([6, 4, 5, 6])          # utt 3: 4 equal labels need 7 frames -> infeasible
    loss64, nll64, g64 = co.reference_call(x, y, il.tolist(), tl, torch.float64)
    xd = x.to(cuda).requires_grad_(True)
    loss, nll = m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True, return_nll=True)
    loss.backward()
    nll_c = nll.cpu().double()
    assert torch.isinf(nll_c[3]) and torch.isinf(nll64[3])
    assert torch.allclose(nll_c[:3], nll64[:3], rtol=1e-4, atol=1e-5)
    gd = xd.grad.cpu().double()
    assert torch.isnan(gd[3, : il[3]]).all() and torch.isnan(g64[3, : il[3]]).all()
    assert float((gd[:3] - g64[:3]).abs().max()) <= 1e-5
    # loss-only path (no grad requested) agrees with the grad path
    with torch.no_grad():
        _, nll2 = m.ctc_loss(x.to(cuda), y.to(cuda), il, tl, from_logits=True, batch_first=True, return_nll=True)
    assert torch.allclose(nll2.cpu()[:3], nll.detach().cpu()[:3], rtol=1e-5, atol=1e-5) and torch.isinf(nll2[3])


def test_long_sequence_is_closer_to_float64_than_fp32_aten(cuda):
    """T=1500, V=2000: fp32 log-space without renormalisation loses ~3 digits; ours must not."""
    m = _mod()
    x, y, il, tl = make_case(2, 1500, 2000, 200, seed=21, full_len=True)
    _, nll64, g64 = co.reference_call(x, y, il.tolist(), tl, torch.float64)
    _, _, g32 = co.reference_call(x, y, il.tolist(), tl, torch.float32)
    xd = x.to(cuda).requires_grad_(True)
    loss, nll = m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True, return_nll=True)
    loss.backward()
    ours = float((xd.grad.cpu().double() - g64).abs().max())
    aten = float((g32.double() - g64).abs().max())
    assert torch.allclose(nll.cpu().double(), nll64, rtol=1e-5)
    assert ours <= aten / 10, (ours, aten)
    err = (xd.grad.cpu().double() - g64).abs()
    scale = 1.0 / (2 * tl.double())
    # T=1500: 5e-4 of the row scale (fp32 log-space spread between forward-best and path states)
    assert float((err / (1e-4 * g64.abs() + 5e-4 * scale.view(2, 1, 1))).max()) <= 1.0


def test_m5_maximum_gradient_against_float64(cuda):
    """BASELINE.json config 5 at its largest point (T'=3000, L=600, V=5000), two utterances: nll and the WHOLE
    gradient against the float64 evaluation of the reference's call sequence (solver.py:160 + torch CTCLoss).
    Tolerance: nll 1e-5 relative; gradient 1e-4 relative + 1e-3 of the row scale 1/(B*L_b) (3000 fp32 log-space
    steps; the fp32 ATen kernel the reference runs is required to be at least 10x further from float64)."""
    m = _mod()
    x, y, il, tl = make_case(2, 3000, 5000, 600, seed=33)
    il[1] = 2400
    _, nll64, g64 = co.reference_call(x, y, il.tolist(), tl, torch.float64)
    _, _, g32 = co.reference_call(x, y, il.tolist(), tl, torch.float32)
    xd = x.to(cuda).requires_grad_(True)
    loss, nll = m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True, return_nll=True)
    loss.backward()
    g = xd.grad.cpu().double()
    assert torch.allclose(nll.cpu().double(), nll64, rtol=1e-5)
    err = (g - g64).abs()
    scale = (1.0 / (2 * tl.double())).view(2, 1, 1)
    ratio = float((err / (1e-4 * g64.abs() + 1e-3 * scale)).max())
    aten = float((g32.double() - g64).abs().max())
    print("T'=3000 gradient: worst err/tol %.3f, max abs err ours %.3e, ATen fp32 %.3e" % (ratio, float(err.max()), aten))
    assert ratio <= 1.0, ratio
    assert float(err.max()) <= aten / 10, (float(err.max()), aten)
    assert float(g[1, 2400:].abs().max()) == 0.0


def test_determinism(cuda):
    m = _mod()
    x, y, il, tl = make_case(6, 120, 31, 40, seed=9, repeats=0.4)
    outs = []
    for _ in range(3):
        xd = x.to(cuda).requires_grad_(True)
        m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True).backward()
        outs.append(xd.grad.clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_second_backward_over_the_same_graph_rescales_a_copy(cuda):
    """the stored gradient is rescaled in place when the upstream gradient differs from the prescale
    folded in at forward time; a SECOND backward (retain_graph) with yet another upstream must not compound it."""
    m = _mod()
    ce = importlib.import_module("end-to-end-asr-pytorch_b200.ce")
    x, y, il, tl = make_case(3, 40, 20, 8, seed=12)
    _, _, g1 = co.reference_call(x, y, il.tolist(), tl, torch.float64, weight=1.0)
    xd = x.to(cuda).requires_grad_(True)
    loss = m.ctc_loss(xd, y.to(cuda), il, tl, from_logits=True, batch_first=True, grad_prescale=1.0)
    (0.3 * loss).backward(retain_graph=True)
    first = xd.grad.clone()
    xd.grad = None
    (2.0 * loss).backward()
    assert float((first.cpu().double() - 0.3 * g1).abs().max()) <= 1e-4 * float(g1.abs().max())
    assert float((xd.grad.cpu().double() - 2.0 * g1).abs().max()) <= 2e-4 * float(g1.abs().max())
    # the sequence loss takes the same route
    logits = torch.randn(2, 5, 9, device=cuda, requires_grad=True)
    lab = torch.randint(1, 9, (2, 5), device=cuda)
    tlen = torch.tensor([5, 5], device=cuda)
    l2 = ce.seq_loss_mean(logits, lab, tlen, grad_prescale=0.5)
    (0.5 * l2).backward(retain_graph=True)
    a = logits.grad.clone(); logits.grad = None
    (1.5 * l2).backward()
    assert torch.allclose(logits.grad, 3.0 * a, rtol=1e-5, atol=1e-8)
