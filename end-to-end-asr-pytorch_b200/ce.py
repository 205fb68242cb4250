"""Attention-branch sequence loss on the sm_100a kernel of csrc/ce.cu.

Mirrors what the reference does with ``torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none')``
(src/solver.py:90) at its two call sites:

* training, solver.py:149-155 -- ``seq_loss_mean``: per-row CE, summed per utterance, divided by the target
  length, averaged over the batch.  One pass over the logits produces the scalar, the finished gradient w.r.t. the logits
  (the upstream factor the caller expects is folded in, as for the CTC loss) and the row argmax that
  ``cal_acc`` / ``cal_cer`` (postprocess.py:121-146) would otherwise recompute with another pass;
* validation, solver.py:239 -- ``CrossEntropyLoss``: the ``reduction='none'`` vector with ordinary autograd.

No CPU path: tensors must live on a CUDA device and libb2asr.so must be built.
"""
import torch

from . import _lib


def _launch(logits, labels, rows_per_group, group_len, ignore_index, prescale, want_grad, want_scalar, want_argmax):
    _lib.require_cuda(logits, labels, group_len)
    if logits.dtype != torch.float32:
        raise ValueError("ce loss: fp32 logits required, got %s" % logits.dtype)
    if logits.dim() != 2:
        raise ValueError("ce loss: logits must be [N,V]")
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    N, V = logits.shape
    labels = labels.reshape(-1).to(torch.int64).contiguous()
    if labels.numel() != N:
        raise ValueError("ce loss: %d labels for %d rows" % (labels.numel(), N))
    if group_len is not None:
        group_len = group_len.to(torch.int64).contiguous()
        if group_len.numel() * rows_per_group != N:
            raise ValueError("ce loss: %d groups of %d rows do not cover %d rows" % (group_len.numel(), rows_per_group, N))
    dev = logits.device
    loss_rows = torch.empty(N, dtype=torch.float32, device=dev)
    argmax = torch.empty(N, dtype=torch.int64, device=dev) if want_argmax else None
    grad = torch.empty(N, V, dtype=torch.float32, device=dev) if want_grad else None
    scalar = torch.empty(1, dtype=torch.float32, device=dev) if want_scalar else None
    _lib.call("b2asr_ce_loss_fwd_bwd", 2 if want_scalar else 1, _lib.lib().b2asr_ce_loss_fwd_bwd,
              logits.data_ptr(), logits.stride(0), labels.data_ptr(), N, V, int(rows_per_group), _lib.ptr(group_len),
              int(ignore_index), float(prescale), loss_rows.data_ptr(), _lib.ptr(argmax), _lib.ptr(grad),
              V, _lib.ptr(scalar), _lib.stream_ptr())
    return loss_rows, argmax, grad, (scalar[0] if want_scalar else None)


class _SeqLossMean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, att_pred, label, tgt_len, ignore_index, prescale):
        B, L, V = att_pred.shape
        want_grad = bool(ctx.needs_input_grad[0])
        x = att_pred if att_pred.is_contiguous() else att_pred.contiguous()
        rows, argmax, grad, loss = _launch(x.view(B * L, V), label, L, tgt_len, ignore_index, prescale,
                                           want_grad, True, True)
        ctx.prescale = float(prescale)
        ctx.shape = (B, L, V)
        if want_grad:
            ctx.save_for_backward(grad)
        rows, argmax = rows.view(B, L), argmax.view(B, L)
        ctx.mark_non_differentiable(rows, argmax)
        return loss, rows, argmax

    @staticmethod
    def backward(ctx, g_out, _g_rows, _g_argmax):
        (grad,) = ctx.saved_tensors
        g_out = g_out.reshape(1).to(torch.float32).contiguous()
        last = getattr(ctx, 'folded', None)
        if last is not None:
            # a second backward over the same graph (retain_graph, a loss term used twice): the buffer already holds
            # the FIRST call's upstream, so scale a copy by the ratio instead of rescaling the buffer again
            return (grad * (g_out / last)).view(ctx.shape), None, None, None, None
        _lib.call("b2asr_rescale_if_needed", 1, _lib.lib().b2asr_rescale_if_needed, grad.data_ptr(),
                  grad.numel(), g_out.data_ptr(), ctx.prescale, _lib.stream_ptr())
        ctx.folded = g_out.clone()                        # the upstream gradient now folded into the buffer
        return grad.view(ctx.shape), None, None, None, None


def seq_loss_mean(att_pred, label, tgt_len, ignore_index=0, grad_prescale=1.0, return_aux=False):
    """solver.py:149-155 in one pass over the logits.

    att_pred [B,L,V] fp32 logits, label [B,L] int64 (0 = padding), tgt_len [B] = ``torch.sum(y != 0, -1)``.
    grad_prescale is the upstream gradient the caller expects (``1 - ctc_weight`` at solver.py:163): it is folded
    into the stored gradient, so backward is free when the guess is right and one rescale pass when it is not.
    -> att_loss (0-dim); with return_aux also (loss_rows [B,L], argmax [B,L])."""
    loss, rows, argmax = _SeqLossMean.apply(att_pred, label, tgt_len, ignore_index, grad_prescale)
    return (loss, rows, argmax) if return_aux else loss


class _CERows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        want_grad = bool(ctx.needs_input_grad[0])
        rows, _, grad, _ = _launch(logits, labels, 1, None, ignore_index, float(logits.shape[0]), want_grad, False, False)
        if want_grad:
            ctx.save_for_backward(grad)                    # softmax - onehot per row (prescale N cancels 1/G)
        return rows

    @staticmethod
    def backward(ctx, g_rows):
        (grad,) = ctx.saved_tensors
        return grad * g_rows.reshape(-1, 1), None, None


class CrossEntropyLoss(torch.nn.Module):
    """Call-compatible with ``torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none')`` as the reference
    builds it (src/solver.py:90) and calls it (solver.py:151,239): logits [N,V], labels [N] -> [N]."""

    def __init__(self, ignore_index=-100, reduction='none'):
        super().__init__()
        if reduction != 'none':
            raise ValueError("CrossEntropyLoss: only reduction='none' (the reference's setting) is supported")
        self.ignore_index = ignore_index

    def forward(self, logits, labels):
        return _CERows.apply(logits, labels, self.ignore_index)
