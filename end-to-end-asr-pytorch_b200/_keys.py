"""Kernel group (2a): key projection K = tanh(h psi^T + b), once per batch (reference asr.py:412-419).

Forward runs on the tcgen05 kernel when csrc provides b2asr_attn_keys (3xTF32 split for fp32-class
accuracy); backward is two plain library GEMMs around the tanh derivative.
"""
import torch
import torch.nn.functional as F

from . import _lib


class _KeysFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, w, b):
        keys = _forward(h, w, b)
        ctx.save_for_backward(h, w, keys)
        return keys

    @staticmethod
    def backward(ctx, g):
        h, w, keys = ctx.saved_tensors
        dz = g * (1.0 - keys * keys)                       # [B,T,A]
        dz2 = dz.reshape(-1, dz.shape[-1])
        g_h = dz2.mm(w).view_as(h) if ctx.needs_input_grad[0] else None
        g_w = dz2.t().mm(h.reshape(-1, h.shape[-1])) if ctx.needs_input_grad[1] else None
        g_b = dz2.sum(0) if ctx.needs_input_grad[2] else None
        return g_h, g_w, g_b


def _forward(h, w, b):
    L = _lib.lib()
    if hasattr(L, "b2asr_attn_keys"):
        from . import _keys_native
        return _keys_native.forward(h, w, b)
    return torch.tanh(F.linear(h, w, b))                   # library GEMM until the tcgen05 kernel lands


def project(h, w, b):
    _lib.require_cuda(h, w, b)
    if torch.is_grad_enabled() and (h.requires_grad or w.requires_grad or b.requires_grad):
        return _KeysFunction.apply(h, w, b)
    return _forward(h.detach(), w.detach(), b.detach())
