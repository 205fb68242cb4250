"""b2asr_gemm_tn_3xtf32 (csrc/keys2.cu: the key projection's tcgen05 pipeline without the tanh) against float64.

Tolerance: the Listener products feed kernels held to 1e-4 relative; here |got - ref| <= 2e-5 * (1 + |ref|) with
operands scaled so that results are O(1) -- cuBLAS fp32 itself is at ~2.5e-6 there, one TF32 pass at ~1.5e-3."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def _gemm():
    return importlib.import_module("end-to-end-asr-pytorch_b200._gemm")


@pytest.mark.parametrize("M,N,K,bias", [(12800, 2560, 640, False), (51200, 2560, 120, False), (12800, 640, 2560, False),
                                        (1000, 336, 100, True), (129, 81, 36, True), (128, 80, 32, False),
                                        (4096, 2560, 640, True)])
def test_gemm_matches_float64(cuda, M, N, K, bias):
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g).to(cuda)
    b = (torch.randn(N, K, generator=g) * K ** -0.5).to(cuda)
    c = (torch.randn(N, generator=g) * 0.1).to(cuda) if bias else None
    mod = _gemm()
    assert mod.ENABLED
    old, mod.MIN_WORK = mod.MIN_WORK, 0
    try:
        got = mod.mm_tn(a, b, c)
    finally:
        mod.MIN_WORK = old
    ref = a.double() @ b.double().t()
    if bias:
        ref = ref + c.double()
    flag = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    lib32 = torch.nn.functional.linear(a, b, c)
    torch.backends.cuda.matmul.allow_tf32 = flag
    err = float((got.double() - ref).abs().max())
    print("M=%d N=%d K=%d: max abs err %.2e (cuBLAS fp32 %.2e)" % (M, N, K, err, float((lib32.double() - ref).abs().max())))
    assert float(((got.double() - ref).abs() / (1 + ref.abs())).max()) <= 2e-5


def test_gemm_unsupported_shapes_take_the_library(cuda):
    mod = _gemm()
    a = torch.randn(300, 30, device=cuda)          # K % 4 != 0
    b = torch.randn(50, 30, device=cuda)
    got = mod.mm_tn(a, b)
    assert torch.allclose(got, a @ b.t(), atol=1e-5)
    a2 = torch.randn(300, 33, device=cuda)[:, 1:]  # unaligned, non-contiguous view
    got = mod.mm_tn(a2, torch.randn(50, 32, device=cuda))
    assert got.shape == (300, 50)


@pytest.mark.parametrize("Kr,Ma,Nb", [(12800, 2560, 640), (51200, 2560, 120), (4000, 1280, 320), (1000, 132, 72), (96, 128, 64)])
def test_gemm_nt_matches_float64(cuda, Kr, Ma, Nb):
    """weight-gradient form: a^T b with the reduction over the rows of both operands"""
    g = torch.Generator().manual_seed(Kr + Ma + Nb)
    a = (torch.randn(Kr, Ma, generator=g) * Kr ** -0.5).to(cuda)
    b = torch.randn(Kr, Nb, generator=g).to(cuda)
    mod = _gemm()
    old, mod.MIN_WORK = mod.MIN_WORK, 0
    try:
        got = mod.mm_nt(a, b)
    finally:
        mod.MIN_WORK = old
    assert got is not None and got.shape == (Ma, Nb)
    ref = a.double().t() @ b.double()
    flag = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    lib32 = a.t().mm(b)
    torch.backends.cuda.matmul.allow_tf32 = flag
    err = float((got.double() - ref).abs().max())
    print("Kr=%d Ma=%d Nb=%d: max abs err %.2e (cuBLAS fp32 %.2e)" % (Kr, Ma, Nb, err, float((lib32.double() - ref).abs().max())))
    assert float(((got.double() - ref).abs() / (1 + ref.abs())).max()) <= 2e-5


@pytest.mark.parametrize("B,T,H,nd", [(32, 400, 320, 2), (5, 37, 64, 2), (3, 100, 32, 1)])
def test_gemm_nt_shifted_views_with_utterance_mask(cuda, B, T, H, nd):
    """dW_hh = sum_b sum_t dG[b,t+1]^T h[b,t] (forward direction) and dG[b,t]^T h[b,t+1] (reverse) from strided views
    of the [B,T,nd,4H] gate gradients and the [B,T,nd*H] outputs, rows that pair two utterances excluded."""
    g = torch.Generator().manual_seed(B * T + H)
    dG = (torch.randn(B, T, nd, 4 * H, generator=g) * (B * T) ** -0.5).to(cuda)
    out = torch.randn(B, T, nd * H, generator=g).to(cuda)
    dG2, out2 = dG.view(B * T, nd * 4 * H), out.view(B * T, nd * H)
    mod = _gemm()
    old, mod.MIN_WORK = mod.MIN_WORK, 0
    try:
        for d in range(nd):
            if d == 0:
                a, b = dG2[1:, :4 * H], out2[:-1, :H]
                ref = torch.bmm(dG[:, 1:, 0].double().transpose(1, 2), out[:, :T - 1, :H].double()).sum(0)
            else:
                a, b = dG2[:-1, 4 * H:], out2[1:, H:]
                ref = torch.bmm(dG[:, :T - 1, 1].double().transpose(1, 2), out[:, 1:, H:].double()).sum(0)
            got = mod.mm_nt(a, b, period=T)
            assert got is not None
            assert float(((got.double() - ref).abs() / (1 + ref.abs())).max()) <= 2e-5, (d, float((got.double() - ref).abs().max()))
    finally:
        mod.MIN_WORK = old


def test_gemm_nt_capped_grid_and_slices_agree(cuda):
    g = torch.Generator().manual_seed(3)
    a = (torch.randn(20000, 1280, generator=g) * 0.01).to(cuda)
    b = torch.randn(20000, 320, generator=g).to(cuda)
    mod = _gemm()
    full = mod.mm_nt(a, b)
    capped = mod.mm_nt(a, b, max_ctas=48)
    ref = a.double().t() @ b.double()
    for got in (full, capped):
        assert float(((got.double() - ref).abs() / (1 + ref.abs())).max()) <= 2e-5


def test_wide_linear_autograd_matches_torch(cuda):
    """_gemm.linear (the V=5000 output layers of the shipped libri config): forward, dx, dW, db against nn.Linear in
    float64."""
    mod = _gemm()
    torch.manual_seed(0)
    lin = torch.nn.Linear(320, 1000).to(cuda)
    x = torch.randn(6, 50, 320, device=cuda, requires_grad=True)
    old, mod.MIN_WORK = mod.MIN_WORK, 0
    try:
        y = mod.linear(lin, x)
        assert y.grad_fn is not None and "Linear" in type(y.grad_fn).__name__
        gy = torch.randn_like(y)
        gx, gw, gb = torch.autograd.grad(y, (x, lin.weight, lin.bias), gy)
    finally:
        mod.MIN_WORK = old
    lin64 = torch.nn.Linear(320, 1000).double().to(cuda)
    lin64.load_state_dict({k: v.double() for k, v in lin.state_dict().items()})
    x64 = x.detach().double().requires_grad_(True)
    y64 = lin64(x64)
    rx, rw, rb = torch.autograd.grad(y64, (x64, lin64.weight, lin64.bias), gy.double())
    # the yardstick is the library's own fp32 rounding on the same products (sums of K terms of size O(1)-O(10))
    flag = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    x32 = x.detach().clone().requires_grad_(True)
    y32 = lin(x32)
    lx, lw, lb = torch.autograd.grad(y32, (x32, lin.weight, lin.bias), gy)
    torch.backends.cuda.matmul.allow_tf32 = flag
    for got, lib32, ref in ((y, y32, y64), (gx, lx, rx), (gw, lw, rw), (gb, lb, rb)):
        err = float((got.detach().double() - ref).abs().max())
        err_lib = float((lib32.detach().double() - ref).abs().max())
        assert err <= max(6 * err_lib, 2e-5), (err, err_lib)
    small = torch.nn.Linear(320, 31).to(cuda)                  # narrow layers stay on the library
    assert "_Linear" not in type(mod.linear(small, x).grad_fn).__name__
