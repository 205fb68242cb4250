"""Host-side mirror of the reference's CTC pieces, routed through the sm_100a kernels.

* `CTCLoss` / `ctc_loss`  -- call-compatible with `torch.nn.CTCLoss(blank=0, reduction='mean')`
  as used at reference src/solver.py:93,160,253-254 (CPU `input_lengths` accepted), plus a
  fused-logits mode that swallows the `F.log_softmax(ctc_pred.transpose(0,1), -1)` of
  src/solver.py:160 without materialising the transpose or the log-probs.
* `CTCPrefixScore`        -- reference src/ctc.py:9-101 (see below in this file).

No CPU path: tensors must live on a CUDA device and libb2asr.so must be built.
"""
import torch

from . import _lib

_REDUCTIONS = {"none": 0, "mean": 1, "sum": 2}


def _lengths_on(device, v):
    if torch.is_tensor(v):
        return v.to(device=device, dtype=torch.int64, non_blocking=True).contiguous()
    return torch.as_tensor([int(i) for i in v], dtype=torch.int64).to(device, non_blocking=True)


class _CTCFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, targets, in_len, tgt_len, blank, reduction, from_logits, batch_first, prescale):
        _lib.require_cuda(x, targets, in_len, tgt_len)
        if x.dtype != torch.float32:
            raise ValueError("ctc_loss: fp32 input required, got %s" % x.dtype)
        if x.dim() != 3 or x.stride(2) != 1:
            x = x.contiguous()
        if batch_first:
            B, T, V = x.shape
            sx_b, sx_t = x.stride(0), x.stride(1)
        else:
            T, B, V = x.shape
            sx_t, sx_b = x.stride(0), x.stride(1)
        if targets.dim() != 2:
            raise ValueError("ctc_loss: targets must be [B, Lmax] (padded); 1-D concatenated targets are not supported")
        targets = targets.to(torch.int64)
        if targets.stride(1) != 1:
            targets = targets.contiguous()
        Lmax = int(targets.shape[1])
        want_grad = bool(ctx.needs_input_grad[0])
        L = _lib.lib()
        ws_bytes = L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, int(want_grad))
        if ws_bytes == 0:
            _lib.check(-2, "b2asr_ctc_workspace_bytes")
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        nll = torch.empty(B, dtype=torch.float32, device=x.device)
        red = _REDUCTIONS[reduction]
        loss = torch.empty(1, dtype=torch.float32, device=x.device) if red != 0 else None
        grad = torch.empty_like(x, memory_format=torch.contiguous_format) if want_grad else None
        if want_grad:
            sg_b, sg_t = (grad.stride(0), grad.stride(1)) if batch_first else (grad.stride(1), grad.stride(0))
        else:
            sg_b = sg_t = 0
        _lib.call("b2asr_ctc_loss_fwd_bwd", 3 if want_grad else 2, L.b2asr_ctc_loss_fwd_bwd,
            x.data_ptr(), sx_t, sx_b, targets.data_ptr(), targets.stride(0),
            in_len.data_ptr(), tgt_len.data_ptr(), T, B, V, Lmax, int(blank),
            int(bool(from_logits)), red, float(prescale),
            nll.data_ptr(), _lib.ptr(loss), _lib.ptr(grad), sg_t, sg_b,
            ws.data_ptr(), ws_bytes, _lib.stream_ptr())
        ctx.red = red
        ctx.prescale = float(prescale)
        ctx.batch_first = batch_first
        if want_grad:
            ctx.save_for_backward(grad)
        ctx.mark_non_differentiable(nll)
        out = loss.reshape(()) if red != 0 else nll.clone()
        return out, nll

    @staticmethod
    def backward(ctx, g_out, _g_nll):
        (grad,) = ctx.saved_tensors
        if ctx.red == 0:
            # per-utterance upstream: scale rows (not the hot path)
            shape = (-1, 1, 1) if ctx.batch_first else (1, -1, 1)
            g = grad * (g_out.reshape(shape) / ctx.prescale)
        else:
            g_out = g_out.reshape(1).to(torch.float32).contiguous()
            last = getattr(ctx, 'folded', None)
            if last is not None:                           # second backward over the graph: see ce.py
                g = grad * (g_out / last)
            else:
                _lib.call("b2asr_rescale_if_needed", 1, _lib.lib().b2asr_rescale_if_needed, grad.data_ptr(),
                          grad.numel(), g_out.data_ptr(), ctx.prescale, _lib.stream_ptr())
                ctx.folded = g_out.clone()
                g = grad
        return g, None, None, None, None, None, None, None, None


def ctc_loss(x, targets, input_lengths, target_lengths, blank=0, reduction="mean",
             from_logits=False, batch_first=False, grad_prescale=1.0, return_nll=False):
    """CTC loss on the sm_100a kernels.

    x: [T,B,V] (or [B,T,V] with batch_first=True; any batch/time strides, V contiguous).
    from_logits=False : x are log-probabilities -> drop-in for torch.nn.functional.ctc_loss.
    from_logits=True  : x are raw logits; log-softmax is fused and the gradient is w.r.t. logits.
    grad_prescale     : the upstream gradient the caller expects (e.g. the joint-CTC weight of
                        solver.py:163); it is folded into the stored gradient so that backward is
                        free when the guess is right and a rescale pass when it is not.
    """
    if reduction not in _REDUCTIONS:
        raise ValueError("ctc_loss: unknown reduction %r" % (reduction,))
    in_len = _lengths_on(x.device, input_lengths)
    tgt_len = _lengths_on(x.device, target_lengths)
    loss, nll = _CTCFunction.apply(x, targets, in_len, tgt_len, blank, reduction, from_logits,
                                   batch_first, grad_prescale)
    return (loss, nll) if return_nll else loss


class CTCLoss(torch.nn.Module):
    """Call-compatible with torch.nn.CTCLoss as the reference uses it (src/solver.py:93,160)."""

    def __init__(self, blank=0, reduction="mean", zero_infinity=False, from_logits=False,
                 batch_first=False, grad_prescale=1.0):
        super().__init__()
        if zero_infinity:
            raise ValueError("CTCLoss: zero_infinity=True is not supported (the reference uses the default False)")
        self.blank, self.reduction = blank, reduction
        self.from_logits, self.batch_first, self.grad_prescale = from_logits, batch_first, grad_prescale

    def forward(self, log_probs, targets, input_lengths, target_lengths):
        return ctc_loss(log_probs, targets, input_lengths, target_lengths, self.blank, self.reduction,
                        self.from_logits, self.batch_first, self.grad_prescale)


class CTCPrefixScore():
    """CTC prefix scorer with the reference's interface (src/ctc.py:9-101), computed by the sm_100a
    kernels of csrc/beam.cu instead of numpy on the host.

    `x` is the [1,T,V] CTC log-posterior tensor (asr.py:180-181) and stays on the device; states and
    scores are returned as numpy float32 arrays with the reference's shapes (r [T,2]; psi [K],
    r [K,T,2]) so existing callers keep working.  The device-resident beam search (beam.py) does not
    go through this class - it keeps the states on the device - but uses the same device code.
    """

    def __init__(self, x):
        _lib.require_cuda(x)
        self.logzero = -100000000.0
        self.blank = 0
        self.eos = 1
        self.x = x.detach()[0].to(torch.float32).contiguous()       # [T,V] on the device
        self.odim = x.shape[-1]
        self.input_length = int(self.x.shape[0])

    def init_state(self):
        """r[t,1] = sum_{i<=t} x[i,blank] (sequential float32 sum), r[t,0] = logzero (ctc.py:19-27)."""
        r = torch.empty(self.input_length, 2, dtype=torch.float32, device=self.x.device)
        _lib.call("b2asr_ctc_prefix_init", 1, _lib.lib().b2asr_ctc_prefix_init, self.x.data_ptr(),
                  self.input_length, self.odim, self.blank, r.data_ptr(), _lib.stream_ptr())
        return r.cpu().numpy()

    def _score(self, g, r_prev, candidates):
        dev = self.x.device
        T = self.input_length
        cands = [int(c) for c in candidates]
        rp = torch.as_tensor(r_prev, dtype=torch.float32).to(dev).contiguous()
        last = int(g[-1]) if len(g) > 0 else 0
        psi = torch.empty(len(cands), dtype=torch.float32, device=dev)
        r = torch.empty(len(cands), T, 2, dtype=torch.float32, device=dev)
        cd = torch.as_tensor(cands, dtype=torch.int32).to(dev)
        for k0 in range(0, len(cands), 32):                         # <= 32 candidates per launch
            k1 = min(len(cands), k0 + 32)
            _lib.call("b2asr_ctc_prefix_score", 1, _lib.lib().b2asr_ctc_prefix_score, self.x.data_ptr(), T, self.odim,
                      len(g), last, rp.data_ptr(), cd[k0:k1].data_ptr(), k1 - k0, psi[k0:k1].data_ptr(),
                      r[k0:k1].data_ptr(), _lib.stream_ptr())
        return psi.cpu().numpy(), r.cpu().numpy()

    def cheap_compute(self, g, r_prev, candidates):
        """Scores of g+[c] for the given candidates only (ctc.py:65-101) -> (psi [K], r [K,T,2])."""
        return self._score(g, r_prev, candidates)

    def full_compute(self, g, r_prev):
        """Scores of g+[c] for every c in the vocabulary (ctc.py:29-63) -> (psi [V], r [V,T,2])."""
        return self._score(g, r_prev, range(self.odim))
