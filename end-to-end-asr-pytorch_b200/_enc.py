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

import torch

from . import _lib

_status = {}


def status_word(device):
    """The per-device int32 words of the kernels: [0] raised when a launch gives up waiting for a peer CTA,
    [1] / [2] counts of failed exchange polls forward / backward."""
    key = (device.type, device.index)
    if key not in _status:
        _status[key] = torch.zeros(16, dtype=torch.int32, device=device)
    return _status[key]


def check_status(device):
    """Synchronising check used by tests and by the trainer's NaN guard."""
    s = status_word(device)
    if int(s[0].item()) != 0:
        s.zero_()
        raise _lib.B2asrError("b2asr_lstm_seq: a persistent launch timed out waiting for a peer CTA")


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
        gates = torch.mm(x.view(B * T, n_in), w_cat.t()).view(B, T, nd, 4 * H)       # activations in place below
        out = torch.empty(B, T, nd * H, dtype=torch.float32, device=dev)
        cells = torch.empty(B, T, nd, H, dtype=torch.float32, device=dev)
        io = _descr(B, T, H, nd, gates, w_hh, b_ih, b_hh, lens, cells, dev)
        io.out = _p(out)
        _lib.call("b2asr_lstm_seq_fwd", 1, _lib.lib().b2asr_lstm_seq_fwd, ctypes.byref(io), _lib.stream_ptr())
        ctx.nd = nd
        ctx.keep = (x, lens, w_cat, w_hh, b_ih, b_hh, gates, cells, out)
        ctx.has_bias = [b is not None for b in b_ih]
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
        _lib.call("b2asr_lstm_seq_bwd", 1, _lib.lib().b2asr_lstm_seq_bwd, ctypes.byref(io), _lib.stream_ptr())
        need = ctx.needs_input_grad
        dG2 = d_gates.view(B * T, nd * 4 * H)
        g_x = dG2.mm(w_cat).view(B, T, n_in) if need[0] else None
        g_wcat = dG2.t().mm(x.view(B * T, n_in))                                      # [ND*4H, in]
        g_b = dG2.sum(0)
        grads = []
        for d in range(nd):
            g_wih = g_wcat[d * 4 * H:(d + 1) * 4 * H]
            if T > 1:
                # h_{t-1} of the forward walk is out[:, t-1, :H]; of the reverse walk out[:, t+1, H:]
                dg = d_gates[:, 1:, d] if d == 0 else d_gates[:, :T - 1, d]           # [B,T-1,4H] strided
                hp = out[:, :T - 1, d * H:(d + 1) * H] if d == 0 else out[:, 1:, d * H:(d + 1) * H]
                g_whh = torch.bmm(dg.transpose(1, 2), hp).sum(0)
            else:
                g_whh = torch.zeros_like(w_hh[d])
            g_bias = g_b[d * 4 * H:(d + 1) * 4 * H]
            grads += [g_wih, g_whh, g_bias if ctx.has_bias[d] else None, g_bias if ctx.has_bias[d] else None]
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
