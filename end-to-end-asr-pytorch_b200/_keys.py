"""Kernel group (2a): key projection K = tanh(h psi^T + b), once per batch (reference asr.py:412-419).

Forward runs on the tcgen05 kernel b2asr_attn_keys (3xTF32 split for fp32-class accuracy); backward is the tanh
derivative and two products on the same pipeline (_gemm.mm_tn / mm_nt; library GEMMs for shapes those do not take).
The Listener's tanh(Linear) after every layer (reference asr.py:313-316) is the same function and uses it too.
"""
import torch
import torch.nn.functional as F

from . import _enc, _gemm, _lib


class _KeysFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, w, b):
        keys = _forward(h, w, b)
        ctx.save_for_backward(h, w, keys)
        ctx.params = (w, b)
        return keys

    @staticmethod
    def backward(ctx, g):
        h, w, keys = ctx.saved_tensors
        dz = g * (1.0 - keys * keys)                       # [B,T,A]
        dz2 = dz.reshape(-1, dz.shape[-1])
        h2 = h.reshape(-1, h.shape[-1])
        g_h = g_w = None
        if ctx.needs_input_grad[0]:
            g_h = _gemm.mm_tn(dz2, w.t().contiguous()).view_as(h)          # dz W
        pw, pb = ctx.params
        if (_enc._armed[0] and not torch.is_grad_enabled() and ctx.needs_input_grad[1] and pw.is_leaf
                and (pb is None or pb.is_leaf)):
            # inside solver.TrainStep's backward pass (see _enc.deferred_weight_grads): nothing downstream reads the
            # weight gradients, so they join the Listener's on the side stream, next to the next recurrence kernel
            ready = torch.cuda.Event()
            ready.record()
            want_b = ctx.needs_input_grad[2] and pb is not None

            def work(side, shared):
                side.wait_event(ready)
                for t in (dz2, h2):
                    t.record_stream(side)
                gw = _gemm.mm_nt(dz2, h2, max_ctas=_gemm.SIDE_CTAS if shared else 0)
                if gw is None:
                    gw = dz2.t().mm(h2)
                for p, g in ((pw, gw), (pb, dz2.sum(0) if want_b else None)):
                    if g is None:
                        continue
                    if p.grad is None:
                        p.grad = g.clone()
                    else:
                        p.grad.add_(g)

            _enc._deferred.append(work)
            if not _enc._join_queued[0]:
                torch.autograd.Variable._execution_engine.queue_callback(_enc._join)
                _enc._join_queued[0] = True
            return g_h, None, None
        if ctx.needs_input_grad[1]:
            g_w = _gemm.mm_nt(dz2, h2)                                     # dz^T h
            if g_w is None:
                g_w = dz2.t().mm(h2)
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
