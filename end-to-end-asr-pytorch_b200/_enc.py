"""One Listener layer's (bi)LSTM over a padded batch (reference src/asr.py:466-501: pack_padded_sequence ->
nn.LSTM -> pad_packed_sequence) as ONE autograd node around the persistent recurrence kernels of csrc/enc.cu.

Per layer the host does three things the library path did per frame:

    forward   G = x W_ih^T for every frame and both directions (one GEMM)  ->  b2asr_lstm_seq_fwd walks the frames
    backward  b2asr_lstm_seq_bwd walks them back and leaves dG             ->  dx = dG W_ih, dW_ih = dG^T x,
              dW_hh = sum_t dG_t^T h_{t-1} (one strided bmm per direction), db = sum dG

Frames past a row's length produce zero outputs and zero gradients, exactly like the packed sequence of the
reference; there is no dependence on the rows being sorted by length.
"""
import ctypes
import os

import torch

from . import _gemm, _lib

_status = {}

# The recurrence kernels occupy 96 of the 148 SMs (7 resident 16-CTA clusters).  A layer's weight-gradient GEMMs feed
# nothing downstream in the backward pass, so they are held back until the NEXT (lower) layer's recurrence kernel has
# been launched and then run on a side stream next to it; the gradients are accumulated into .grad there and the
# main stream joins at the end of the backward pass (autograd engine callback).  B2ASR_ENC_OVERLAP=0: everything in
# stream order, gradients returned through autograd.
# Returning None for the weight gradients and adding them to .grad from a side stream is only legitimate where the
# caller owns .grad and nobody asks autograd for those gradients (torch.autograd.grad, backward(inputs=...), hooks, DDP
# would all see None): the path is therefore taken only inside `deferred_weight_grads()`, which solver.TrainStep
# enters around its backward pass; everywhere else the gradients are returned through autograd in stream order.
OVERLAP = os.environ.get("B2ASR_ENC_OVERLAP", "1") != "0"
_side = {}
_deferred = []
_join_queued = [False]
_armed = [False]


class deferred_weight_grads:
    """Context manager around a backward pass whose LSTM weight gradients may be accumulated into .grad from the side
    stream.  If the backward pass raises, the held-back closures are dropped (they would otherwise run during the next
    backward pass on tensors of a dead graph) and the side stream is joined."""

    def __enter__(self):
        self._prev = _armed[0]
        _armed[0] = OVERLAP
        return self

    def __exit__(self, exc_type, exc, tb):
        _armed[0] = self._prev
        if exc_type is not None:
            del _deferred[:]
            _join_queued[0] = False
            if _side and torch.cuda.is_available():
                device = torch.device("cuda", torch.cuda.current_device())
                torch.cuda.current_stream(device).wait_stream(_side_stream(device))
        return False


def _side_stream(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _side:
        _side[key] = torch.cuda.Stream(device=device)
    return _side[key]


def _flush(device, after=None):
    """Enqueue the held-back weight-gradient work on the side stream; `after` = event the side stream waits for first
    (recorded just before the recurrence launch it is meant to run next to, so it cannot start ahead of it)."""
    if not _deferred:
        return
    work = list(_deferred)
    del _deferred[:]
    side = _side_stream(device)
    if after is not None:
        side.wait_event(after)
    with torch.cuda.stream(side):
        for fn in work:
            fn(side, after is not None)         # after: the work runs next to the recurrence launched behind that event


def _join():
    """End of a backward pass: nothing may read .grad on the caller's stream before the side stream is done."""
    _join_queued[0] = False
    if not _side and not _deferred:
        return
    device = torch.device("cuda", torch.cuda.current_device())
    _flush(device)
    torch.cuda.current_stream(device).wait_stream(_side_stream(device))


def status_word(device):
    """The per-device int32 words of the kernels: [0] raised when a launch gives up waiting for a peer CTA,
    [1] / [2] counts of failed exchange polls forward / backward."""
    key = (device.type, device.index)
    if key not in _status:
        _status[key] = torch.zeros(16, dtype=torch.int32, device=device)
    return _status[key]


def check_status(device):
    """Synchronising check: raises (and re-arms the word) if a recurrence launch gave up waiting for a peer CTA, after
    which its outputs are undefined.  solver.TrainStep reads the word together with its NaN-skip counter
    (`TrainStep.check()`, called wherever the trainer already reads a value back), Trainer.valid and beam_decode call
    it after their forward passes."""
    s = status_word(device)
    if int(s[0].item()) != 0:
        s.zero_()
        raise _lib.B2asrError("b2asr_lstm_seq: a persistent launch timed out waiting for a peer CTA; its outputs "
                              "are undefined (profiler, sanitizer or a shared GPU can cause this)")


def supported(layer, x):
    """LSTM, one layer, no projection, fp32 on CUDA, H a multiple of 4 (csrc/enc.cu: bulk-copy alignment)."""
    return (isinstance(layer, torch.nn.LSTM) and layer.num_layers == 1 and getattr(layer, "proj_size", 0) == 0
            and layer.batch_first and x.is_cuda and x.dtype == torch.float32 and layer.hidden_size % 4 == 0
            and layer.hidden_size <= 1024)


def _p(t):
    return None if t is None else t.data_ptr()


def _descr(B, T, H, ND, gates, w_hh, b_ih, b_hh, lens, cells, dev):
    io = _lib.LstmSeq()
    io.B, io.T, io.H, io.ND = B, T, H, ND
    io.gates = _p(gates)
    for d in range(ND):
        io.w_hh[d] = _p(w_hh[d])
        io.b_ih[d] = _p(b_ih[d])
        io.b_hh[d] = _p(b_hh[d])
    io.lens = _p(lens)
    io.cells = _p(cells)
    io.status = _p(status_word(dev))
    return io


class LstmSeq(torch.autograd.Function):
    """(x [B,T,in], lens int32 [B], per direction w_ih, w_hh, b_ih, b_hh) -> out [B,T,ND*H]."""

    @staticmethod
    def forward(ctx, x, lens, nd, *weights):
        _lib.require_cuda(x, lens)
        w_ih = [weights[4 * d].detach() for d in range(nd)]
        w_hh = [weights[4 * d + 1].detach().contiguous() for d in range(nd)]
        b_ih = [None if weights[4 * d + 2] is None else weights[4 * d + 2].detach().contiguous() for d in range(nd)]
        b_hh = [None if weights[4 * d + 3] is None else weights[4 * d + 3].detach().contiguous() for d in range(nd)]
        x = x.detach().contiguous()
        B, T, n_in = x.shape
        H = w_hh[0].shape[1]
        dev = x.device
        w_cat = torch.cat(w_ih, 0) if nd > 1 else w_ih[0].contiguous()               # [ND*4H, in]
        gates = _gemm.mm_tn(x.view(B * T, n_in), w_cat).view(B, T, nd, 4 * H)        # activations in place below
        out = torch.empty(B, T, nd * H, dtype=torch.float32, device=dev)
        cells = torch.empty(B, T, nd, H, dtype=torch.float32, device=dev)
        io = _descr(B, T, H, nd, gates, w_hh, b_ih, b_hh, lens, cells, dev)
        io.out = _p(out)
        _lib.call("b2asr_lstm_seq_fwd", 1, _lib.lib().b2asr_lstm_seq_fwd, ctypes.byref(io), _lib.stream_ptr())
        ctx.nd = nd
        ctx.keep = (x, lens, w_cat, w_hh, b_ih, b_hh, gates, cells, out)
        ctx.has_bias = [b is not None for b in b_ih]
        ctx.params = weights
        return out

    @staticmethod
    def backward(ctx, d_out):
        x, lens, w_cat, w_hh, b_ih, b_hh, gates, cells, out = ctx.keep
        nd = ctx.nd
        B, T, n_in = x.shape
        H = w_hh[0].shape[1]
        dev = x.device
        d_out = d_out.contiguous()
        d_gates = torch.empty(B, T, nd, 4 * H, dtype=torch.float32, device=dev)
        io = _descr(B, T, H, nd, gates, w_hh, b_ih, b_hh, lens, cells, dev)
        io.d_out = _p(d_out)
        io.d_gates = _p(d_gates)
        before = None
        if _armed[0] and _deferred:
            before = torch.cuda.Event()
            before.record()
        _lib.call("b2asr_lstm_seq_bwd", 1, _lib.lib().b2asr_lstm_seq_bwd, ctypes.byref(io), _lib.stream_ptr())
        need = ctx.needs_input_grad
        dG2 = d_gates.view(B * T, nd * 4 * H)
        params = ctx.params
        has_bias = ctx.has_bias

        def weight_grads(max_ctas=0):
            """max_ctas > 0: the products share the device with a recurrence kernel and take at most that many SMs."""
            x2, out2 = x.view(B * T, n_in), out.view(B * T, nd * H)
            g_wcat = _gemm.mm_nt(dG2, x2, max_ctas=max_ctas)                          # [ND*4H, in]
            if g_wcat is None:
                g_wcat = dG2.t().mm(x2)
            g_b = dG2.sum(0)
            grads = []
            for d in range(nd):
                g_wih = g_wcat[d * 4 * H:(d + 1) * 4 * H]
                if T > 1:
                    # h_{t-1} of the forward walk is out[:, t-1, :H]; of the reverse walk out[:, t+1, H:].  Over the
                    # flattened [B*T] rows that pairs row r with row r+1 (a view offset); the rows that would pair two
                    # utterances (r % T == T-1) are excluded inside the kernel.
                    if d == 0:
                        a2, b2 = dG2[1:, :4 * H], out2[:-1, :H]
                    else:
                        a2, b2 = dG2[:-1, d * 4 * H:(d + 1) * 4 * H], out2[1:, d * H:(d + 1) * H]
                    g_whh = _gemm.mm_nt(a2, b2, period=T, max_ctas=max_ctas)
                    if g_whh is None:
                        dg = d_gates[:, 1:, d] if d == 0 else d_gates[:, :T - 1, d]   # [B,T-1,4H] strided
                        hp = out[:, :T - 1, d * H:(d + 1) * H] if d == 0 else out[:, 1:, d * H:(d + 1) * H]
                        g_whh = torch.bmm(dg.transpose(1, 2), hp).sum(0)
                else:
                    g_whh = torch.zeros_like(w_hh[d])
                g_bias = g_b[d * 4 * H:(d + 1) * 4 * H]
                grads += [g_wih, g_whh, g_bias if has_bias[d] else None, g_bias if has_bias[d] else None]
            return grads

        leaf = all(p is None or (p.is_leaf and p.requires_grad) for p in params)
        if _armed[0] and leaf and torch.is_grad_enabled() is False:
            ready = torch.cuda.Event()
            ready.record()                                  # d_gates complete (this stream)
            _flush(dev, before)                             # upper layers' weight gradients run next to THIS recurrence

            def work(side, shared):
                side.wait_event(ready)
                for t in (d_gates, x, out):
                    t.record_stream(side)
                for p, g in zip(params, weight_grads(_gemm.SIDE_CTAS if shared else 0)):
                    if p is None or g is None:
                        continue
                    if p.grad is None:
                        p.grad = g.clone()
                    else:
                        p.grad.add_(g)

            _deferred.append(work)
            if not _join_queued[0]:
                torch.autograd.Variable._execution_engine.queue_callback(_join)
                _join_queued[0] = True
            g_x = _gemm.mm_tn(dG2, w_cat.t().contiguous()).view(B, T, n_in) if need[0] else None
            ctx.keep = None
            return (g_x, None, None) + (None,) * len(params)
        g_x = _gemm.mm_tn(dG2, w_cat.t().contiguous()).view(B, T, n_in) if need[0] else None
        grads = weight_grads()
        ctx.keep = None
        return (g_x, None, None, *grads)


def lstm_layer(layer, x, state_len):
    """Drop-in for `pad_packed_sequence(layer(pack_padded_sequence(x, state_len)))[0]` on a CUDA batch."""
    nd = 2 if layer.bidirectional else 1
    tmax = int(max(state_len))
    if x.shape[1] != tmax:
        x = x[:, :tmax]
    lens = torch.as_tensor(list(state_len), dtype=torch.int32).to(x.device, non_blocking=True)
    weights = []
    for d in range(nd):
        sfx = "_l0" + ("_reverse" if d else "")
        weights += [getattr(layer, "weight_ih" + sfx), getattr(layer, "weight_hh" + sfx),
                    getattr(layer, "bias_ih" + sfx) if layer.bias else None,
                    getattr(layer, "bias_hh" + sfx) if layer.bias else None]
    return LstmSeq.apply(x, lens, nd, *weights)
