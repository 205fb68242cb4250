"""Kernel group (2a): key projection K = tanh(h psi^T + b), once per batch (reference asr.py:412-419).

Forward runs on the tcgen05 kernel b2asr_attn_keys (3xTF32 split for fp32-class accuracy); backward is
two plain library GEMMs around the tanh derivative (cuBLAS, as SURVEY.md section 2 scopes dense GEMMs).
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
    """tcgen05 kernel (csrc/keys.cu): h [..., D] contiguous -> keys [..., A]."""
    h = h.contiguous()
    w = w.contiguous()
    b = b.contiguous()
    D, A = h.shape[-1], w.shape[0]
    M = h.numel() // D
    out = torch.empty(h.shape[:-1] + (A,), dtype=torch.float32, device=h.device)
    _lib.call("b2asr_attn_keys", 1, _lib.lib().b2asr_attn_keys, h.data_ptr(), w.data_ptr(), b.data_ptr(),
              out.data_ptr(), M, D, A, _lib.stream_ptr())
    return out


def project(h, w, b):
    _lib.require_cuda(h, w, b)
    if torch.is_grad_enabled() and (h.requires_grad or w.requires_grad or b.requires_grad):
        return _KeysFunction.apply(h, w, b)
    return _forward(h.detach(), w.detach(), b.detach())
