"""The teacher-forced decoder loop of Seq2Seq.forward (reference src/asr.py:84-108) as ONE autograd node whose
forward and backward are each one call into the native loop of csrc/decloop.cu.

The per-step path (asr._AttnStep + _dec.DecStep) costs ~130 us of Python/autograd time per decoder step in each
direction against ~55 / ~90 us of GPU work, so the whole loop was launch-bound (tools/phase_wall.py).  Here the
host does per batch: the batched embedding projection, one C call for all L steps, the batched output layer; and in
backward one C call plus the deferred weight-gradient GEMMs over the step stacks:

    dh      = sum_t alpha_t (x) d_ctx_t          (one bmm)          dW_phi = dzq^T [h0, h_1 .. h_{L-1}]
    dW_ctx  = dG^T ctx      dW_hh = dG^T h_prev      db = sum dG      dE = dG
    dW_l, dW_c, dw_g, db_g  = sum of the per-CTA slots the attention backward kernel accumulates

Scheduled sampling (a failed teacher-forcing coin) needs the output layer and a sample between two steps, so such
batches take the per-step path; coins are drawn exactly as the reference draws them (asr.py:96).
"""
import ctypes
import os

import torch
import torch.nn.functional as F
from torch.distributions.categorical import Categorical

from . import _attn, _gemm, _lib


# one SGEMM per step over [ctx_t | h_{t-1}] x [W_ctx | W_hh] instead of two (B2ASR_DEC_CAT=0: the two-GEMM form)
FUSED_OPERANDS = os.environ.get("B2ASR_DEC_CAT", "1") != "0"


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _mm_nt(a, b):
    """a^T b over the rows (weight-gradient form) on the tensor-core pipeline; library GEMM for what it does not take"""
    a, b = a.contiguous(), b.contiguous()
    out = _gemm.mm_nt(a, b)
    return out if out is not None else a.t().mm(b)


def _fill(io, **ptrs):
    for k, v in ptrs.items():
        setattr(io, k, _p(v))


class DecoderLoop(torch.autograd.Function):
    """(keys, h, E, h0, c0, attention weights, cell weights) -> hs [B,L,H], att [L,B,T], h_L, c_L."""

    @staticmethod
    def forward(ctx, meta, keys, h, E, h0, c0, phi, conv, lproj, gen_w, gen_b, w_ih, w_hh, b_ih, b_hh, embed_w=None):
        """meta["coins"] (optional, one bool per step): steps whose coin is False feed a SAMPLED character to the next
        step (scheduled sampling, reference asr.py:96-101).  The loop then runs range by range: after such a step
        meta["char_trans"] gives its logits, one Categorical sample is drawn -- the same draws in the same order as
        the per-step loop -- and the sample's embedding projection replaces row t+1 of the gate buffer.  embed_w is the
        embedding table; its gradient and the embedding half of w_ih's for the sampled rows are formed in backward."""
        _lib.require_cuda(keys, h, E, h0, c0, w_ih, w_hh)
        mode, lens, scale, emb = meta["mode"], meta["lens"], float(meta["scale"]), int(meta["emb_dim"])
        B, T, A = keys.shape
        D = h.shape[-1]
        L, H4 = E.shape[1], E.shape[2]
        H = H4 // 4
        dev = keys.device
        loc = mode == "loc"
        keys, h = keys.detach().contiguous(), h.detach().contiguous()
        w_ih, w_hh = w_ih.detach(), w_hh.detach().contiguous()
        assert w_ih.stride(1) == 1 and w_ih.shape[1] == emb + D
        h0, c0 = h0.detach().contiguous(), c0.detach().contiguous()
        f32 = dict(dtype=torch.float32, device=dev)
        att = torch.empty(L, B, T, **f32)
        xcat = w_cat = None
        if FUSED_OPERANDS:
            xcat = torch.empty(L, B, D + H, **f32)                         # [ctx_t | h_{t-1}]
            w_cat = torch.cat([w_ih[:, emb:], w_hh], 1).contiguous()      # [4H, D+H]
            cvec = xcat[:, :, :D]
        else:
            cvec = torch.empty(L, B, D, **f32)
        q = torch.empty(L, B, A, **f32)
        gates = E.detach().clone(memory_format=torch.contiguous_format)     # [B,L,4H] -> activations in place
        hs = torch.empty(B, L, H, **f32)
        cs = torch.empty(L, B, H, **f32)
        io = _lib.DecLoop()
        io.B, io.T, io.D, io.A, io.H, io.L = B, T, D, A, H, L
        io.mode = _attn.MODES[mode]
        io.C, io.Kw = (conv.shape[0], (conv.shape[-1] - 1) // 2) if loc else (0, 0)
        io.cluster, io.scale = 0, scale
        wts = dict(w_phi=phi, w_conv=conv, w_lproj=lproj, w_gen=gen_w, b_gen=gen_b, b_ih=b_ih, b_hh=b_hh)
        wts = {k: (v.detach().contiguous() if v is not None else None) for k, v in wts.items()}
        w_ctx = w_ih[:, emb:]
        _fill(io, keys=keys, h=h, lens=lens, w_ctx=w_ctx, w_hh=w_hh, h0=h0, c0=c0, att=att,
              ctx=None if FUSED_OPERANDS else cvec, q=q, gates=gates, hs=hs, cs=cs, xcat=xcat, w_cat=w_cat, **wts)
        io.w_ctx_ld = w_ih.stride(0)
        coins = meta.get("coins")
        samples = {}
        if coins is None or all(coins):
            _lib.call("b2asr_decoder_loop_fwd", 2 * L, _lib.lib().b2asr_decoder_loop_fwd, ctypes.byref(io), _lib.stream_ptr())
        else:
            fwd_range = _lib.lib().b2asr_decoder_loop_fwd_range
            w_emb = w_ih[:, :emb]
            t0 = 0
            for t in range(L):
                if coins[t]:
                    continue
                _lib.call("b2asr_decoder_loop_fwd_range", 2 * (t + 1 - t0), fwd_range, ctypes.byref(io), t0, t + 1, _lib.stream_ptr())
                t0 = t + 1
                tok = Categorical(F.softmax(meta["char_trans"](hs[:, t]), dim=-1)).sample()
                samples[t] = tok
                if t + 1 < L:
                    gates[:, t + 1] = F.linear(embed_w.detach()[tok], w_emb)
            if t0 < L:
                _lib.call("b2asr_decoder_loop_fwd_range", 2 * (L - t0), fwd_range, ctypes.byref(io), t0, L, _lib.stream_ptr())
        ctx.samples = {t: tok for t, tok in samples.items() if t + 1 < L}
        ctx.embed_w = embed_w
        ctx.io, ctx.meta = io, meta
        ctx.keep = (keys, h, lens, w_ih, w_hh, wts, h0, c0, att, cvec, q, gates, hs, cs, xcat, w_cat)   # what io points at
        ctx.mark_non_differentiable(att)
        ctx.set_materialize_grads(False)
        return hs, att, hs[:, L - 1].clone(), cs[L - 1].clone()

    @staticmethod
    def backward(ctx, d_hs, _d_att, d_h_last, d_c_last):
        io, meta = ctx.io, ctx.meta
        keys, h, lens, w_ih, w_hh, wts, h0, c0, att, cvec, q, gates, hs, cs, xcat, w_cat = ctx.keep
        B, T, D, A, H, L = io.B, io.T, io.D, io.A, io.H, io.L
        emb = int(meta["emb_dim"])
        loc = meta["mode"] == "loc"
        dev = keys.device
        f32 = dict(dtype=torch.float32, device=dev)
        d_hs = torch.zeros(B, L, H, **f32) if d_hs is None else d_hs.contiguous()
        d_h_last = None if d_h_last is None else d_h_last.contiguous()
        d_c_last = None if d_c_last is None else d_c_last.contiguous()
        d_gates = torch.empty(B, L, 4 * H, **f32)
        d_xcat = None
        if xcat is not None:
            d_xcat = torch.empty(L, B, D + H, **f32)                       # [d_ctx_t | recurrent d_h_{t-1}]
            d_ctx = d_xcat[:, :, :D]
        else:
            d_ctx = torch.empty(L, B, D, **f32)
        d_zq = torch.empty(L, B, A, **f32)
        d_keys = torch.zeros(B, T, A, **f32)
        pgrad = None
        if loc:
            stride = _attn.pgrad_layout(A, io.C, io.Kw)[-1]
            slots = _attn.num_slots(B, T, D, A, H, "loc", io.C, io.Kw)
            pgrad = torch.zeros(slots, stride, **f32)
            io.pgrad_stride = stride
        d_prev = torch.empty(2, B, T, **f32) if loc else None
        d_s = torch.zeros(B, H, **f32)
        d_hrec = torch.zeros(B, H, **f32)
        d_c = torch.zeros(2, B, H, **f32)
        _fill(io, d_hs=d_hs, d_h_last=d_h_last, d_c_last=d_c_last, d_gates=d_gates,
              d_ctx=None if d_xcat is not None else d_ctx, d_zq=d_zq, d_keys=d_keys, pgrad=pgrad, d_prev=d_prev, d_s=d_s,
              d_hrec=d_hrec, d_c=d_c, d_xcat=d_xcat)
        _lib.call("b2asr_decoder_loop_bwd", 2 * L, _lib.lib().b2asr_decoder_loop_bwd, ctypes.byref(io), _lib.stream_ptr())

        need = ctx.needs_input_grad
        g_keys = d_keys if need[1] else None
        g_h = torch.bmm(att.permute(1, 2, 0), d_ctx.transpose(0, 1)) if need[2] else None       # [B,T,L] x [B,L,D]
        g_E = d_gates if need[3] else None
        g_h0 = (d_s + (d_xcat[0, :, D:] if d_xcat is not None else d_hrec)) if need[4] else None
        g_c0 = d_c[0] if need[5] else None
        g_phi = g_conv = g_lproj = g_gw = g_gb = g_wih = g_whh = g_bih = g_bhh = None
        h_prev = torch.cat([h0.unsqueeze(1), hs[:, :L - 1]], dim=1)                               # [B,L,H]: s_t = h_{t-1}
        if wts["w_phi"] is not None and need[6]:
            g_phi = _mm_nt(d_zq.transpose(0, 1).reshape(B * L, A), h_prev.reshape(B * L, H))
        if loc:
            o_wl, o_wc, o_wg, o_bg, _ = _attn.pgrad_layout(A, io.C, io.Kw)
            tot = pgrad.sum(0)
            g_lproj = tot[o_wl:o_wc].view(A, io.C)
            g_conv = tot[o_wc:o_wg].view(io.C, 1, 2 * io.Kw + 1)
            g_gw = tot[o_wg:o_bg].view(1, A)
            g_gb = tot[o_bg:o_bg + 1]
        dG2 = d_gates.view(B * L, 4 * H)
        if xcat is not None and need[11] and need[12]:
            g_cat = _mm_nt(dG2, xcat.transpose(0, 1).reshape(B * L, D + H))         # [4H, D+H] = [dW_ctx | dW_hh]
            g_wih = torch.zeros(4 * H, emb + D, **f32)                 # the embedding half flows through E's own autograd
            g_wih[:, emb:] = g_cat[:, :D]
            g_whh = g_cat[:, D:]
        else:
            if need[11]:
                g_wih = torch.zeros(4 * H, emb + D, **f32)
                g_wih[:, emb:] = _mm_nt(dG2, cvec.transpose(0, 1).reshape(B * L, D))
            if need[12]:
                g_whh = _mm_nt(dG2, h_prev.reshape(B * L, H))
        if wts["b_ih"] is not None and (need[13] or need[14]):
            g_bih = dG2.sum(0)
            g_bhh = g_bih
        g_embed = None
        if ctx.samples:
            # rows of the gate buffer that held a sampled character's projection embed_w[tok] w_emb^T: their gradient
            # belongs to the embedding table and to the embedding half of w_ih, not to the teacher's row of E
            steps = sorted(ctx.samples)
            idx = torch.tensor([t + 1 for t in steps], device=dev)
            toks = torch.stack([ctx.samples[t] for t in steps], dim=1).reshape(-1)              # [B*n]
            rows = d_gates[:, idx].reshape(-1, 4 * H)                                          # [B*n, 4H]
            embed_w = ctx.embed_w.detach()
            if need[11]:
                if g_wih is None:
                    g_wih = torch.zeros(4 * H, emb + D, **f32)
                g_wih[:, :emb] += rows.t().mm(embed_w[toks])
            if need[15]:
                g_embed = torch.zeros_like(embed_w).index_add_(0, toks, rows.mm(w_ih[:, :emb]))
            if g_E is not None:
                g_E = g_E.clone()
                g_E[:, idx] = 0
        ctx.keep = None
        return (None, g_keys, g_h, g_E, g_h0, g_c0, g_phi, g_conv, g_lproj, g_gw, g_gb, g_wih, g_whh, g_bih, g_bhh, g_embed)
