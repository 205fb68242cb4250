"""Decoder-loop fusion for the Speller's single-layer LSTMCell.

The reference steps ``torch.nn.LSTMCell`` on ``cat([last_char, context])`` once per decoder step
(src/asr.py:84-104, 352-357): per step that is a cat, two GEMMs, a pointwise kernel, and in backward four
more GEMMs plus the running sums of every weight gradient (about 25 launches).  Here the same arithmetic is
regrouped around what actually recurs:

  W_ih [emb | ctx] x [last_char ; context] = W_emb last_char + W_ctx context

* the embedding half of every teacher-forced step is one GEMM before the loop (``E`` [B,L,4H]);
* per step: ``gates = E[:,t] + context W_ctx^T + h W_hh^T`` (two GEMM launches) and one pointwise kernel
  (csrc/lstm.cu) -> 3 launches forward, 3 backward (pointwise, d_context, d_h);
* the weight gradients dW_ctx, dW_hh, db are one GEMM each after the loop, from the [B,L,4H] buffer the
  per-step backward kernels write into; dE is that buffer itself.

Same pattern as the attention session in asr.py: an entry node whose backward runs after all step backwards.
"""
import torch

from . import _lib


def cell_forward(gates, b_ih, b_hh, c_prev):
    """gates [B,4H] pre-activation (overwritten with the activations) -> h, c [B,H]."""
    B, H4 = gates.shape
    H = H4 // 4
    h = torch.empty(B, H, dtype=torch.float32, device=gates.device)
    c = torch.empty_like(h)
    _lib.call("b2asr_lstm_cell_fwd", 1, _lib.lib().b2asr_lstm_cell_fwd,
              gates.data_ptr(), _lib.ptr(b_ih), _lib.ptr(b_hh), c_prev.data_ptr(), h.data_ptr(), c.data_ptr(),
              B, H, _lib.stream_ptr())
    return h, c


def cell_backward(d_h, d_c, act, c_prev, c, d_gates_out):
    """-> d_c_prev; writes the pre-activation gate gradients into d_gates_out ([B,4H] view, any row stride)."""
    B, H4 = act.shape
    H = H4 // 4
    assert d_gates_out.stride(1) == 1
    d_c_prev = torch.empty(B, H, dtype=torch.float32, device=act.device)
    _lib.call("b2asr_lstm_cell_bwd", 1, _lib.lib().b2asr_lstm_cell_bwd,
              _lib.ptr(d_h), _lib.ptr(d_c), act.data_ptr(), c_prev.data_ptr(), c.data_ptr(),
              d_gates_out.data_ptr(), d_gates_out.stride(0), d_c_prev.data_ptr(), B, H, _lib.stream_ptr())
    return d_c_prev


class DecSession:
    """One batch's decoder loop: detached weights, the batched embedding projection, and what the deferred
    weight-gradient GEMMs need."""

    def __init__(self, cell, E, emb_dim, steps):
        w_ih = cell.weight_ih.detach()
        self.emb_dim = emb_dim
        self.w_ctx = w_ih[:, emb_dim:]                     # [4H, D] view
        self.w_hh = cell.weight_hh.detach()                # [4H, H]
        self.b_ih = cell.bias_ih.detach() if cell.bias_ih is not None else None
        self.b_hh = cell.bias_hh.detach() if cell.bias_hh is not None else None
        self.E = E.detach() if E is not None else None     # [B, L, 4H]
        self.steps = steps
        self.dG = None
        self.ctxs = [None] * steps
        self.hprevs = [None] * steps
        self.own_e = [False] * steps                       # step fed by its own (sampled) embedding, not E[:, t]

    def grad_buffer(self, B, H4, device):
        if self.dG is None:
            self.dG = torch.zeros(B, self.steps, H4, dtype=torch.float32, device=device)
        return self.dG


class DecOpen(torch.autograd.Function):
    """Entry node: every step depends on its token, so this backward runs last and hands out dE and the
    weight gradients from the filled [B,L,4H] buffer."""

    @staticmethod
    def forward(ctx, session, E, w_ih, w_hh, b_ih, b_hh):
        ctx.session = session
        ctx.has_E = E is not None
        ctx.set_materialize_grads(False)
        return torch.zeros(1, dtype=torch.float32, device=w_ih.device)

    @staticmethod
    def backward(ctx, _g):
        ses = ctx.session
        g_E = g_wih = g_whh = g_bih = g_bhh = None
        dG = ses.dG
        if dG is not None:
            used = [t for t in range(ses.steps) if ses.ctxs[t] is not None]
            B, L, H4 = dG.shape
            if len(used) == ses.steps:
                dG2 = dG.reshape(B * L, H4)
                CTX = torch.stack(ses.ctxs, dim=1).reshape(B * L, -1)
                HP = torch.stack(ses.hprevs, dim=1).reshape(B * L, -1)
            else:                                            # loop ended early
                dG2 = dG[:, used].reshape(B * len(used), H4)
                CTX = torch.stack([ses.ctxs[t] for t in used], dim=1).reshape(B * len(used), -1)
                HP = torch.stack([ses.hprevs[t] for t in used], dim=1).reshape(B * len(used), -1)
            if ctx.needs_input_grad[2]:
                g_wih = torch.zeros(H4, ses.emb_dim + CTX.shape[1], dtype=torch.float32, device=dG.device)
                g_wih[:, ses.emb_dim:] = dG2.t().mm(CTX)
            if ctx.needs_input_grad[3]:
                g_whh = dG2.t().mm(HP)
            if ses.b_ih is not None and (ctx.needs_input_grad[4] or ctx.needs_input_grad[5]):
                g_bih = dG2.sum(0)
                g_bhh = g_bih
            if ctx.has_E and ctx.needs_input_grad[1]:
                g_E = dG
                if any(ses.own_e):
                    g_E = dG.clone()
                    for t, own in enumerate(ses.own_e):
                        if own:
                            g_E[:, t].zero_()
        ses.ctxs, ses.hprevs = [None] * ses.steps, [None] * ses.steps
        return None, g_E, g_wih, g_whh, g_bih, g_bhh


class DecStep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, session, token, t, e_t, context, h_prev, c_prev):
        ses = session
        base = e_t if e_t is not None else ses.E[:, t]
        gates = torch.addmm(base, context, ses.w_ctx.t())
        gates.addmm_(h_prev, ses.w_hh.t())
        h, c = cell_forward(gates, ses.b_ih, ses.b_hh, c_prev)
        ctx.session, ctx.t, ctx.own_e = ses, t, e_t is not None
        ctx.save_for_backward(gates, c_prev, c, context, h_prev)
        ctx.set_materialize_grads(False)
        return h, c

    @staticmethod
    def backward(ctx, d_h, d_c):
        ses, t = ctx.session, ctx.t
        act, c_prev, c, context, h_prev = ctx.saved_tensors
        dG = ses.grad_buffer(act.shape[0], act.shape[1], act.device)
        dg_t = dG[:, t]
        if d_h is not None and not d_h.is_contiguous():
            d_h = d_h.contiguous()
        if d_c is not None and not d_c.is_contiguous():
            d_c = d_c.contiguous()
        d_c_prev = cell_backward(d_h, d_c, act, c_prev, c, dg_t)
        ses.ctxs[t], ses.hprevs[t], ses.own_e[t] = context, h_prev, ctx.own_e
        d_ctx = dg_t.mm(ses.w_ctx) if ctx.needs_input_grad[4] else None
        d_hp = dg_t.mm(ses.w_hh) if ctx.needs_input_grad[5] else None
        d_e = dg_t if (ctx.own_e and ctx.needs_input_grad[3]) else None
        # the token carries no gradient: it only orders DecOpen.backward after every step
        return None, None, None, d_e, d_ctx, d_hp, (d_c_prev if ctx.needs_input_grad[6] else None)


def supported(decoder):
    """Single-layer LSTMCell without dropout: what both shipped configs use (config/*.yaml `decoder`)."""
    cell = getattr(decoder, "layer0", None)
    return (decoder.layer == 1 and isinstance(cell, torch.nn.LSTMCell) and
            (decoder.dropout.p == 0 or not decoder.training) and cell.weight_ih.is_cuda)


def step_nograd(cell, emb_dim, e_t, context, h_prev, c_prev):
    """Inference step (no autograd bookkeeping)."""
    gates = torch.addmm(e_t, context, cell.weight_ih[:, emb_dim:].t())
    gates.addmm_(h_prev, cell.weight_hh.t())
    return cell_forward(gates, cell.bias_ih, cell.bias_hh, c_prev)
