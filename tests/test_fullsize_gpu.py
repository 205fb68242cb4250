"""BASELINE.json full-size configurations through size-independent properties (the CPU oracle would take minutes):
CTC at the M5 maximum (T'=3000, L=600, V=5000, B=64) and the attention step at M2 (B=32, T'=400, D=640, A=300)."""
import importlib

import pytest
import torch

PKG = "end-to-end-asr-pytorch_b200"
pytestmark = pytest.mark.gpu


def test_ctc_m5_maximum_properties(cuda):
    ctc = importlib.import_module(PKG + ".ctc")
    B, T, V, L = 64, 3000, 5000, 600
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, T, V, device=cuda)
    il = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); il[0] = T
    tl = torch.randint(L // 2, L + 1, (B,), generator=g); tl[0] = L
    y = torch.randint(1, V, (B, L), generator=g)
    y[:, 1::10] = y[:, 0::10][:, : y[:, 1::10].shape[1]]          # 10 % forced adjacent repeats
    for b in range(B):
        y[b, tl[b]:] = 0
    yd, tld, ild = y.to(cuda), tl.to(cuda), il.to(cuda)
    xr = x.requires_grad_(True)
    loss = ctc.ctc_loss(xr, yd, ild, tld, from_logits=True, batch_first=True)
    loss.backward()
    grad = xr.grad
    assert torch.isfinite(loss) and torch.isfinite(grad).all()
    # rows past the input length are exactly zero; every valid row's gradient sums to 0 (softmax sums to the row
    # scale, the occupancies of a frame sum to the same scale).  Tolerance: 3e-3 of the row scale 1/(B*L_b) -- after
    # 3000 fp32 log-space steps the occupancies of a frame sum to 1 within ~1e-3 (ATen's own fp32 kernel is further
    # from the float64 result, see test_ctc_gpu.py::test_long_sequence_is_closer_to_float64_than_fp32_aten)
    t_idx = torch.arange(T, device=cuda)[None, :]
    valid = t_idx < ild[:, None]
    assert float(grad[~valid].abs().max()) == 0.0
    scale = 1.0 / (B * tld.float())
    rowsum = grad.sum(-1)
    assert float((rowsum.abs() / scale[:, None])[valid].max()) <= 3e-3
    # blank column: grad = scale*(softmax - occupancy) >= -scale and <= scale everywhere
    assert float((grad.abs().amax(-1) / scale[:, None]).max()) <= 1.0 + 3e-3
    # a subset of utterances against ATen's CUDA CTC on the same logits (fp32), 1e-4 relative on nll
    sub = [0, 7, 31, 63]
    lp = torch.log_softmax(x.detach()[sub].double(), -1).transpose(0, 1)
    ref = torch.nn.functional.ctc_loss(lp, yd[sub], ild[sub], tld[sub], blank=0, reduction="none")
    nll = ctc.ctc_loss(x.detach()[sub].contiguous(), yd[sub], ild[sub], tld[sub], from_logits=True, batch_first=True, reduction="none")
    assert torch.allclose(nll.double(), ref, rtol=1e-4)
    # linearity in the prescale / determinism: a second evaluation is bit-identical
    xr.grad = None
    loss2 = ctc.ctc_loss(xr, yd, ild, tld, from_logits=True, batch_first=True)
    loss2.backward()
    assert float(loss2) == float(loss) and torch.equal(xr.grad, grad)


def test_attention_m2_properties(cuda):
    attn = importlib.import_module(PKG + "._attn")
    B, T, D, A, dec = 32, 400, 640, 300, 320
    g = torch.Generator().manual_seed(2)
    h = torch.randn(B, T, D, generator=g).to(cuda)
    k = torch.tanh(torch.randn(B, T, A, generator=g)).to(cuda)
    lens_list = sorted([int(v) for v in torch.randint(250, T + 1, (B,), generator=g)], reverse=True)
    lens = attn.lens_i32(lens_list, cuda)
    s = torch.randn(B, dec, generator=g).to(cuda)
    W = dict(phi=torch.randn(A, dec, generator=g) * dec ** -0.5, conv=torch.randn(10, 1, 201, generator=g) * 201 ** -0.5,
             lproj=torch.randn(A, 10, generator=g) * 10 ** -0.5, gw=torch.randn(1, A, generator=g) * A ** -0.5, gb=torch.zeros(1))
    W = {n: v.to(cuda) for n, v in W.items()}
    prev = None
    for _ in range(3):
        att, ctx, _ = attn.step_forward("loc", s, W["phi"], k, h, lens, prev, W["conv"], W["lproj"], W["gw"], W["gb"], 2.0)
        mask = torch.arange(T, device=cuda)[None, :] >= lens[:, None]
        assert float(att[mask].abs().max()) == 0.0                               # padded frames get exactly 0 (asr.py:454)
        assert torch.allclose(att.sum(-1), torch.ones(B, device=cuda), atol=1e-5)  # softmax over the valid frames
        assert torch.allclose(ctx, torch.bmm(att.unsqueeze(1), h).squeeze(1), rtol=1e-4, atol=1e-5)   # asr.py:457
        prev = att
