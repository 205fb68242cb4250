"""Raw (no-autograd) Python entry points of the attention kernels; autograd wiring lives in asr.py."""
import torch

from . import _lib

MODES = {"dot": 0, "loc": 1}


def lens_i32(lens, device):
    if torch.is_tensor(lens):
        return lens.to(device=device, dtype=torch.int32, non_blocking=True).contiguous()
    return torch.as_tensor([int(v) for v in lens], dtype=torch.int32).to(device, non_blocking=True)


def step_forward(mode, s, w_phi, keys, h, lens, prev_att, w_conv, w_lproj, w_gen, b_gen, scale=2.0,
                 cluster=0, want_q=False, shared_memory=False, out=None):
    """One attention step.  keys/h: [B,T,A]/[B,T,D], or [1,T,*] with shared_memory=True (beam decode:
    every row of `s` attends over the same utterance).  Returns (att [B,T], ctx [B,D], q [B,A] | None)."""
    _lib.require_cuda(s, keys, h, lens)
    B, dec = s.shape
    T, A = keys.shape[-2], keys.shape[-1]
    D = h.shape[-1]
    keys = keys.contiguous()
    h = h.contiguous()
    if s.stride(1) != 1:
        s = s.contiguous()
    if out is None:
        att = torch.empty(B, T, dtype=torch.float32, device=s.device)
        ctx = torch.empty(B, D, dtype=torch.float32, device=s.device)
    else:
        att, ctx = out
    q = torch.empty(B, A, dtype=torch.float32, device=s.device) if want_q else None
    kb = 0 if shared_memory else T * A
    hb = 0 if shared_memory else T * D
    loc = mode == "loc"
    if loc:
        C, KW1 = w_conv.shape[0], w_conv.shape[-1]
        Kw = (KW1 - 1) // 2
        w_conv = w_conv.contiguous()
        w_lproj = w_lproj.contiguous()
        w_gen = w_gen.contiguous()
    else:
        C, Kw = 0, 0
    _lib.call("b2asr_attn_step_fwd", 1, _lib.lib().b2asr_attn_step_fwd,
        MODES[mode], s.data_ptr(), s.stride(0), _lib.ptr(w_phi), keys.data_ptr(), kb, h.data_ptr(), hb,
        lens.data_ptr(), _lib.ptr(prev_att), _lib.ptr(w_conv) if loc else 0, C, Kw,
        _lib.ptr(w_lproj) if loc else 0, _lib.ptr(w_gen) if loc else 0,
        _lib.ptr(b_gen) if loc else 0, float(scale), att.data_ptr(), ctx.data_ptr(), _lib.ptr(q),
        B, T, D, A, dec, int(cluster), _lib.stream_ptr())
    return att, ctx, q


def pgrad_layout(A, C, Kw):
    """Offsets of (dWl, dWc, dwg, dbg) inside one parameter-gradient slot, and the slot stride."""
    o_wl, o_wc, o_wg, o_bg = 0, A * C, A * C + C * (2 * Kw + 1), A * C + C * (2 * Kw + 1) + A
    return o_wl, o_wc, o_wg, o_bg, o_bg + 1


def num_slots(B, T, D, A, dec, mode, C, Kw, cluster=0):
    n = _lib.lib().b2asr_attn_slots(B, T, D, A, dec, MODES[mode], C, Kw, int(cluster))
    if n <= 0:
        _lib.check(n, "b2asr_attn_slots")
    return n


def step_backward(mode, w_phi, keys, h, lens, prev_att, w_conv, w_lproj, w_gen, scale, q_saved, att_saved,
                  d_ctx, d_att, d_keys, pgrad, dec, cluster=0, want_prev=True):
    """Backward of one step.  Accumulates into d_keys [B,T,A] and pgrad [slots,stride] in place; returns
    (d_prev_att [B,T] | None, d_zq [B,A], d_s [B,dec])."""
    B, T, A = keys.shape
    D = h.shape[-1]
    loc = mode == "loc"
    dev = keys.device
    d_prev = torch.empty(B, T, dtype=torch.float32, device=dev) if (loc and want_prev) else None
    d_zq = torch.empty(B, A, dtype=torch.float32, device=dev)
    d_s = torch.empty(B, dec, dtype=torch.float32, device=dev)
    if loc:
        C, Kw = w_conv.shape[0], (w_conv.shape[-1] - 1) // 2
    else:
        C, Kw = 0, 0
    d_ctx = d_ctx.contiguous()
    if d_att is not None:
        d_att = d_att.contiguous()
    _lib.call("b2asr_attn_step_bwd", 1, _lib.lib().b2asr_attn_step_bwd,
        MODES[mode], _lib.ptr(w_phi), keys.data_ptr(), T * A, h.data_ptr(), T * D, lens.data_ptr(), _lib.ptr(prev_att),
        _lib.ptr(w_conv) if loc else 0, C, Kw, _lib.ptr(w_lproj) if loc else 0, _lib.ptr(w_gen) if loc else 0,
        float(scale), q_saved.data_ptr(), att_saved.data_ptr(), d_ctx.data_ptr(), _lib.ptr(d_att),
        d_keys.data_ptr(), T * A, _lib.ptr(d_prev), d_zq.data_ptr(), d_s.data_ptr(),
        _lib.ptr(pgrad), pgrad.stride(0) if pgrad is not None else 0,
        B, T, D, A, dec, int(cluster), _lib.stream_ptr())
    return d_prev, d_zq, d_s
