"""Dense products next to the Listener recurrence (reference src/asr.py:466-501: nn.LSTM's input projection and its
input gradient) on the tcgen05 pipeline of csrc/keys2.cu instead of cuBLAS's fp32 SIMT kernels.

    mm_tn(a, b, bias)   a [M,K] . b[N,K]^T (+ bias[N])  ->  [M,N]      (torch.nn.functional.linear's operand form)

Error-compensated TF32 (three tensor-core products per algorithmic one, fp32 accumulation in tensor memory): the max
abs error against float64 is ~2x cuBLAS fp32's (tests/test_gemm_gpu.py), far inside the 1e-4 parity bar, at about three
times the SIMT kernels' rate.  Shapes the tensor copies cannot address (K % 4 != 0, unaligned views) take the library
GEMM; B2ASR_GEMM=cublas selects the library everywhere (A/B runs).
"""
import os

import torch

from . import _lib

ENABLED = os.environ.get("B2ASR_GEMM", "tcgen05") != "cublas"
# SMs a weight-gradient product may take while a Listener recurrence kernel (96 of the 148 SMs at the shipped shapes)
# runs next to it; a persistent grid wider than the free SMs would make the recurrence's clusters wait for it
SIDE_CTAS = int(os.environ.get("B2ASR_SIDE_CTAS", "52"))
FLOPS = {"b2asr_gemm_tn_3xtf32": 0.0, "b2asr_gemm_nt_3xtf32": 0.0}    # algorithmic 2*M*N*K issued so far (bench.py reads it)
MIN_WORK = 1 << 22            # M*N*K below which a library GEMM launch is as good (tiny test shapes, decoder steps)


def mm_tn(a, b, bias=None):
    M, K = a.shape
    N = b.shape[0]
    if (ENABLED and a.is_cuda and a.dtype == torch.float32 and b.dtype == torch.float32 and K % 4 == 0
            and M * N * K >= MIN_WORK and a.is_contiguous() and b.is_contiguous()
            and a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0):
        out = torch.empty(M, N, dtype=torch.float32, device=a.device)
        FLOPS["b2asr_gemm_tn_3xtf32"] += 2.0 * M * N * K
        if bias is not None:
            bias = bias.contiguous()
        _lib.call("b2asr_gemm_tn_3xtf32", 1, _lib.lib().b2asr_gemm_tn_3xtf32, a.data_ptr(), b.data_ptr(), _lib.ptr(bias),
                  out.data_ptr(), M, N, K, _lib.stream_ptr())
        return out
    return torch.nn.functional.linear(a, b, bias)


def mm_nt(a, b, period=0, max_ctas=0):
    """a [Kr,Ma], b [Kr,Nb] (row-major views, unit column stride) -> a^T b [Ma,Nb], the reduction over the rows; rows
    with index % period == period - 1 do not count (see include/b2asr.h).  Returns None when the tensor-core kernel
    does not take the operands (the caller then uses its library route)."""
    Kr, Ma = a.shape
    Nb = b.shape[1]
    if not (ENABLED and a.is_cuda and a.dtype == torch.float32 and b.dtype == torch.float32 and b.shape[0] == Kr
            and a.stride(1) == 1 and b.stride(1) == 1 and a.stride(0) % 4 == 0 and b.stride(0) % 4 == 0
            and a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0 and (period == 0 or period >= 32)
            and Kr * Ma * Nb >= MIN_WORK):
        return None
    L = _lib.lib()
    S = L.b2asr_gemm_nt_3xtf32_slices(Ma, Nb, Kr, max_ctas)
    out = torch.empty(S, Ma, Nb, dtype=torch.float32, device=a.device)
    FLOPS["b2asr_gemm_nt_3xtf32"] += 2.0 * Kr * Ma * Nb
    _lib.call("b2asr_gemm_nt_3xtf32", 1, L.b2asr_gemm_nt_3xtf32, a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0),
              out.data_ptr(), Ma, Nb, Kr, period, S, max_ctas, _lib.stream_ptr())
    return out[0] if S == 1 else out.sum(0)


class _Linear(torch.autograd.Function):
    """y = x W^T + b for a wide output layer (the V=5000 `ctc_layer` / `char_trans` of the shipped libri config,
    reference asr.py:38-44): the forward product, dx = dy W and dW = dy^T x on the tensor-core pipeline."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1])
        ctx.save_for_backward(x2, w)
        ctx.shape, ctx.has_bias = x.shape, b is not None
        return mm_tn(x2.contiguous(), w.contiguous(), b).view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, g):
        x2, w = ctx.saved_tensors
        g2 = g.reshape(-1, g.shape[-1]).contiguous()
        g_x = g_w = g_b = None
        if ctx.needs_input_grad[0]:
            g_x = mm_tn(g2, w.t().contiguous()).view(ctx.shape)
        if ctx.needs_input_grad[1]:
            g_w = mm_nt(g2, x2.contiguous())
            if g_w is None:
                g_w = g2.t().mm(x2)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            g_b = g2.sum(0)
        return g_x, g_w, g_b


def linear(layer, x):
    """layer(x) for an nn.Linear; wide layers on CUDA fp32 go through the tensor-core products, everything else (and
    B2ASR_GEMM=cublas) through the library."""
    w = layer.weight
    if (ENABLED and x.is_cuda and x.dtype == torch.float32 and w.shape[0] >= 256 and w.shape[0] % 4 == 0
            and w.shape[1] % 4 == 0 and x.numel() // x.shape[-1] * w.shape[0] * w.shape[1] >= MIN_WORK):
        return _Linear.apply(x, w, layer.bias)
    return layer(x)
