"""Device-resident joint CTC / attention / RNN-LM beam search (reference src/asr.py:155-258).

The reference walks the live hypotheses one by one (batch 1), pulls the attention log-probs to the
host for `topk`, scores CTC prefixes in numpy, and keeps every hypothesis as a Python object with CPU
copies of its decoder / attention / LM / CTC state.  Here one decode step is

    attention kernel (all live hypotheses as rows, keys/features shared)   -> kernel (2b)
    LSTMCell stack, char_trans, log_softmax, RNN-LM step (library calls, batched over the beam)
    b2asr_beam_step                                                          -> kernel (3)
    row gathers of the neural state by the surviving children's parent index

with all state in fixed device buffers, so the step is shape-static and is replayed from a CUDA
graph; nothing crosses to the host until the last step, when the token history (a few KB) is read
back once and the `Hypothesis` list the reference returns is rebuilt from it.

  BeamSearch     state buffers + b2asr_beam_step launches; usable on its own with externally supplied
                 log-probs (that is how the parity tests feed oracle and kernel identical logits)
  beam_decode    the Seq2Seq.beam_decode body
"""
import math
import os

import numpy as np
import torch
import torch.nn.functional as F

from . import _attn, _lib
from .postprocess import Hypothesis

CTC_BEAM_RATIO = 1.5   # asr.py:15


class _Scalars:
    """List-like view of a 1-D CPU tensor whose items are 0-dim tensors, as the reference's Hypothesis keeps its
    tokens and scores (postprocess.py:49-50, 82-85): items are made when asked for (a beam of 20 x 160 tokens as 6 400
    eager 0-dim tensors cost 8-20 ms per utterance), slices and concatenations give plain lists, append() works."""

    def __init__(self, t):
        self._t = t
        self._extra = []

    def __len__(self):
        return int(self._t.shape[0]) + len(self._extra)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return list(self)[i]
        n = int(self._t.shape[0])
        if i < 0:
            i += n + len(self._extra)
        if i < 0 or i >= n + len(self._extra):
            raise IndexError(i)
        return self._t[i] if i < n else self._extra[i - n]

    def __iter__(self):
        for v in self._t.unbind(0):
            yield v
        for v in self._extra:
            yield v

    def append(self, v):
        self._extra.append(v)

    def __add__(self, other):
        return list(self) + list(other)

    def __eq__(self, other):
        return list(self) == list(other)

    def __repr__(self):
        return repr(list(self))


class BeamSearch:
    """Rows = decode_beam_size hypotheses of ONE utterance; `valid` marks the live ones."""

    def __init__(self, vocab, beam, max_steps, device, x_ctc=None, ctc_weight=0.0, lm_weight=0.0):
        if beam < 1 or beam > 32:
            raise ValueError('decode_beam_size must be in 1..32, got %d' % beam)
        self.V, self.n, self.max_steps, self.dev = int(vocab), int(beam), int(max_steps), device
        self.K = int(CTC_BEAM_RATIO * beam)
        if x_ctc is not None and self.K > 32:
            raise ValueError('joint CTC decoding supports int(1.5*beam) <= 32 candidates (beam <= 21), got beam %d' % beam)
        self.x_ctc = None if x_ctc is None else x_ctc.detach().to(device, torch.float32).contiguous()
        self.T = 0 if self.x_ctc is None else int(self.x_ctc.shape[0])
        # [V, T'] copy, made once per utterance: the per-step gather of the candidates' columns becomes K row reads
        self.x_ctc_t = None
        if self.x_ctc is not None:
            Tp4 = (self.T + 3) // 4 * 4
            self.x_ctc_t = torch.zeros(self.x_ctc.shape[1], Tp4, dtype=torch.float32, device=device)
            self.x_ctc_t[:, :self.T] = self.x_ctc.t()
        self.ctc_weight, self.lm_weight = float(ctc_weight), float(lm_weight)
        n, T, K, V = self.n, max(self.T, 1), max(self.K, 1), self.V
        f32 = dict(dtype=torch.float32, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        i64 = dict(dtype=torch.int64, device=device)
        # double-buffered search state
        self.valid = [torch.zeros(n, **i32) for _ in range(2)]
        self.last = [torch.zeros(n, **i64) for _ in range(2)]
        self.r = [torch.zeros(n, T, 2, **f32) for _ in range(2)]
        self.ctc_prob = [torch.zeros(n, **f32) for _ in range(2)]
        self.sum = [torch.zeros(n, **f32) for _ in range(2)]
        self.parent = torch.zeros(n, **i64)
        # scratch + history
        self.cand = torch.zeros(n, K, **i32)
        self.psi = torch.zeros(n, K, **f32)
        self.r_all = torch.zeros(n, K, T, 2, **f32)
        self.topv = torch.zeros(n, n, **f32)
        self.topi = torch.zeros(n, n, **i32)
        self.step_ctr = torch.zeros(1, **i32)
        self.ticket = torch.zeros(1, **i32)
        self.hist_parent = torch.full((self.max_steps, n), -1, **i32)
        self.hist_tok = torch.full((self.max_steps, n), -1, **i32)
        self.hist_score = torch.zeros(self.max_steps, n, **f32)
        self.fin_count = torch.zeros(1, **i32)
        self.fin_step = torch.zeros(self.max_steps * n, **i32)
        self.fin_row = torch.zeros(self.max_steps * n, **i32)
        self.fin_score = torch.zeros(self.max_steps * n, **f32)
        self.fin_sum = torch.zeros(self.max_steps * n, **f32)
        self.cur = 0
        self.steps_done = 0
        # root hypothesis (asr.py:196-199): empty prefix, ctc_prob 0, blank-only CTC state
        self.valid[0][0] = 1
        if self.x_ctc is not None:
            r0 = torch.empty(self.T, 2, **f32)
            _lib.call("b2asr_ctc_prefix_init", 1, _lib.lib().b2asr_ctc_prefix_init, self.x_ctc.data_ptr(), self.T,
                      self.V, 0, r0.data_ptr(), _lib.stream_ptr())
            self.r[0][0].copy_(r0)
        self._io = [None, None]

    def _make_io(self, src, att_logp, lm_logp, att_lse=None, lm_lse=None):
        dst = 1 - src
        io = _lib.BeamIO()
        p = _lib.ptr
        io.x_ctc = p(self.x_ctc)
        io.x_ctc_t, io.x_ctc_t_ld = p(self.x_ctc_t), (0 if self.x_ctc_t is None else self.x_ctc_t.stride(0))
        io.T, io.V, io.n, io.K, io.max_steps = self.T, self.V, self.n, self.K, self.max_steps
        io.w_att = float(np.float32(1 - self.ctc_weight))       # python double (1-w) cast to fp32, as torch does
        io.w_ctc = float(np.float32(self.ctc_weight))
        io.w_lm = float(np.float32(self.lm_weight))
        io.att_logp, io.lm_logp = p(att_logp), p(lm_logp)
        io.valid_in, io.last_in, io.r_in = p(self.valid[src]), p(self.last[src]), p(self.r[src])
        io.ctc_prob_in, io.sum_in = p(self.ctc_prob[src]), p(self.sum[src])
        io.valid_out, io.last_out, io.r_out = p(self.valid[dst]), p(self.last[dst]), p(self.r[dst])
        io.ctc_prob_out, io.sum_out, io.parent_out = p(self.ctc_prob[dst]), p(self.sum[dst]), p(self.parent)
        io.cand, io.psi, io.r_all, io.topv, io.topi = p(self.cand), p(self.psi), p(self.r_all), p(self.topv), p(self.topi)
        io.step, io.ticket = p(self.step_ctr), p(self.ticket)
        io.hist_parent, io.hist_tok, io.hist_score = p(self.hist_parent), p(self.hist_tok), p(self.hist_score)
        io.fin_count, io.fin_step, io.fin_row = p(self.fin_count), p(self.fin_step), p(self.fin_row)
        io.fin_score, io.fin_sum = p(self.fin_score), p(self.fin_sum)
        io.att_lse, io.lm_lse = p(att_lse), p(lm_lse)              # given: att_logp / lm_logp hold raw logits
        return io

    @property
    def last_tokens(self):
        """[n] int64: last token of every row's prefix (0 = <sos>), input of the next step's embeddings."""
        return self.last[self.cur]

    def launch(self, src, att_logp, lm_logp=None):
        """One b2asr_beam_step reading state buffer `src`, writing `1-src`.  att_logp / lm_logp: [n,V]
        fp32 contiguous device tensors whose ADDRESSES the call captures (CUDA-graph friendly)."""
        _lib.require_cuda(att_logp, lm_logp)
        if att_logp.shape != (self.n, self.V) or att_logp.dtype != torch.float32 or not att_logp.is_contiguous():
            raise ValueError('att_logp must be a contiguous fp32 [%d,%d] tensor' % (self.n, self.V))
        if lm_logp is not None and (lm_logp.shape != att_logp.shape or not lm_logp.is_contiguous()):
            raise ValueError('lm_logp must match att_logp')
        io = self._make_io(src, att_logp, lm_logp if self.lm_weight > 0 else None)
        self._io[src] = io                                      # keep the struct alive across the call
        import ctypes
        _lib.call("b2asr_beam_step", 1, _lib.lib().b2asr_beam_step, ctypes.byref(io), _lib.stream_ptr())

    def step(self, att_logp, lm_logp=None):
        """Eager step: launch, flip the buffers, return the parent row of every surviving child [n] int64."""
        if self.steps_done >= self.max_steps:
            raise RuntimeError('BeamSearch: more steps than max_steps=%d' % self.max_steps)
        self.launch(self.cur, att_logp, lm_logp)
        self.flip()
        return self.parent

    def flip(self):
        self.cur = 1 - self.cur
        self.steps_done += 1

    # ---- host side: rebuild what the reference returns -------------------------------------------------
    def history(self):
        """One D2H read of the search history -> dict of numpy arrays."""
        s = self.steps_done
        out = dict(parent=self.hist_parent[:s].cpu().numpy(), tok=self.hist_tok[:s].cpu().numpy(),
                   score=self.hist_score[:s].cpu().numpy(), fin_count=int(self.fin_count.item()))
        f = out['fin_count']
        out.update(fin_step=self.fin_step[:f].cpu().numpy(), fin_row=self.fin_row[:f].cpu().numpy(),
                   fin_score=self.fin_score[:f].cpu().numpy(), valid=self.valid[self.cur].cpu().numpy())
        return out

    @staticmethod
    def _path(h, step, row):
        """Tokens and scores of the hypothesis living in `row` AFTER decode step `step` (step -1 = root)."""
        toks, scs = [], []
        while step >= 0:
            toks.append(int(h['tok'][step, row]))
            scs.append(h['score'][step, row])
            row = int(h['parent'][step, row])
            step -= 1
        return toks[::-1], scs[::-1]

    @staticmethod
    def _paths(h, starts):
        """Vectorised `_path` for many hypotheses at once: starts = [(step, row)], step -1 = the root (empty path).
        One numpy gather per decode step instead of a Python loop per token."""
        m = len(starts)
        steps = np.asarray([s for s, _ in starts], dtype=np.int64)
        cur = np.asarray([r for _, r in starts], dtype=np.int64)
        L = int(steps.max()) + 1 if m else 0
        toks = np.zeros((m, max(L, 1)), dtype=np.int64)
        scs = np.zeros((m, max(L, 1)), dtype=np.float32)
        for st in range(L - 1, -1, -1):
            idx = np.nonzero(steps >= st)[0]
            r = cur[idx]
            toks[idx, st] = h['tok'][st, r]
            scs[idx, st] = h['score'][st, r]
            cur[idx] = h['parent'][st, r]
        return [(toks[i, :steps[i] + 1], scs[i, :steps[i] + 1]) for i in range(m)]

    def results(self, emb=None):
        """asr.py:244-258: finals in finalisation order, then the survivors in rank order, one stable
        sort by mean score, best `beam`.  Returns Hypothesis objects (postprocess.py:44) whose
        output_seq / output_scores are 0-dim CPU tensors, exactly what the reference's carry."""
        h = self.history()

        def hyp(toks, scs):
            # 0-dim CPU tensors like the reference's, made on demand from ONE tensor per hypothesis (_Scalars)
            seq = _Scalars(torch.from_numpy(np.ascontiguousarray(toks, dtype=np.int64)))
            sc = _Scalars(torch.from_numpy(np.ascontiguousarray(scs, dtype=np.float32)))
            return Hypothesis(None, emb, output_seq=seq, output_scores=sc)

        nfin = h['fin_count'] if self.n > 1 else min(h['fin_count'], 1)        # beam 1 returns at the first <eos>, asr.py:246-247
        starts = [(int(h['fin_step'][i]) - 1, int(h['fin_row'][i])) for i in range(nfin)]
        last = self.steps_done - 1
        live = [] if (self.n == 1 and nfin) else [row for row in range(self.n) if h['valid'][row]]
        starts += [(last, row) for row in live]
        paths = self._paths(h, starts)
        outs = []

        def mean_score(scs):
            # Hypothesis.avgScore (postprocess.py:71-74) is Python's sum() over 0-dim fp32 tensors -- a strictly
            # left-to-right fp32 sum -- divided by the count in fp32; np.cumsum is the same sequential sum
            return float(np.cumsum(scs, dtype=np.float32)[-1] / np.float32(len(scs)))

        for i in range(nfin):
            toks, scs = paths[i]
            scs = np.append(scs, np.float32(h['fin_score'][i]))
            outs.append((mean_score(scs), hyp(np.append(toks, 1), scs)))
        if self.n == 1 and nfin:
            return [outs[0][1]]
        for toks, scs in paths[nfin:]:
            outs.append((mean_score(scs), hyp(toks, scs)))
        outs.sort(key=lambda o: o[0], reverse=True)                           # stable, like the reference's sort
        return [o[1] for o in outs[:self.n]]


def _pad4(n):
    return (n + 3) // 4 * 4


def _padded_cols(w, cols):
    """[R, C] -> contiguous [R, cols] with zero columns appended (16-byte aligned rows for the tile kernels)."""
    if w.shape[1] == cols:
        return w.detach().contiguous()
    out = torch.zeros(w.shape[0], cols, dtype=w.dtype, device=w.device)
    out[:, :w.shape[1]] = w.detach()
    return out


class _DecodePack:
    """Per-model constants of the native decode step (csrc/decode.cu), rebuilt when a parameter changes:
    the embedding half of the first Speller / LM layer folded into a [V, 4H] table (+ both biases), and every other
    weight matrix with its inner dimension padded to a multiple of 4 and [x1 | x2] operands side by side."""

    def __init__(self, model, lm):
        dec = model.decoder
        self.version = _DecodePack.stamp(model, lm)
        H = dec.dim
        H4 = _pad4(H)
        D = None
        layers = [getattr(dec, 'layer' + str(i)) for i in range(dec.layer)]
        emb_w = model.embed.weight.detach()
        E = emb_w.shape[1]
        l0 = layers[0]
        D = l0.weight_ih.shape[1] - E
        self.H, self.H4, self.D, self.D4 = H, H4, D, _pad4(D)
        bias0 = 0
        if l0.bias:
            bias0 = (l0.bias_ih + l0.bias_hh).detach()
        self.dec_table = (emb_w @ l0.weight_ih.detach()[:, :E].t() + bias0).contiguous()          # [V, 4H]
        self.dec_w = [torch.cat([_padded_cols(l0.weight_ih.detach()[:, E:], self.D4),
                                 _padded_cols(l0.weight_hh, H4)], 1).contiguous()]
        self.dec_b = [(None, None)]
        for l in layers[1:]:
            self.dec_w.append(torch.cat([_padded_cols(l.weight_ih, H4), _padded_cols(l.weight_hh, H4)], 1).contiguous())
            self.dec_b.append((l.bias_ih.detach().contiguous(), l.bias_hh.detach().contiguous()) if l.bias else (None, None))
        ct = model.char_trans
        self.char_w = _padded_cols(ct.weight, H4)
        self.char_b = None if ct.bias is None else ct.bias.detach().contiguous()
        self.lm = None
        if lm is not None:
            r = lm.rnn
            HL = lm.h_dim
            HL4 = _pad4(HL)
            b0 = (r.bias_ih_l0 + r.bias_hh_l0).detach() if r.bias else 0
            table = (lm.emb.weight.detach() @ r.weight_ih_l0.detach().t() + b0).contiguous()       # [V, 4HL]
            ws = [_padded_cols(r.weight_hh_l0, HL4)]
            bs = [(None, None)]
            for l in range(1, r.num_layers):
                ws.append(torch.cat([_padded_cols(getattr(r, 'weight_ih_l%d' % l), HL4),
                                     _padded_cols(getattr(r, 'weight_hh_l%d' % l), HL4)], 1).contiguous())
                bs.append((getattr(r, 'bias_ih_l%d' % l).detach().contiguous(),
                           getattr(r, 'bias_hh_l%d' % l).detach().contiguous()) if r.bias else (None, None))
            self.lm = dict(HL=HL, HL4=HL4, table=table, w=ws, b=bs, out_w=_padded_cols(lm.out.weight, HL4),
                           out_b=None if lm.out.bias is None else lm.out.bias.detach().contiguous(),
                           layers=r.num_layers)

    @staticmethod
    def stamp(model, lm):
        ps = list(model.decoder.parameters()) + list(model.embed.parameters()) + list(model.char_trans.parameters())
        if lm is not None:
            ps += list(lm.parameters())
        return tuple((p.data_ptr(), p._version) for p in ps)

    @staticmethod
    def supported(model, lm):
        dec = model.decoder
        layers = [getattr(dec, 'layer' + str(i)) for i in range(dec.layer)]
        if not all(isinstance(l, torch.nn.LSTMCell) for l in layers) or len(layers) > 4:
            return False
        if lm is not None and (not isinstance(lm.rnn, torch.nn.LSTM) or lm.rnn.num_layers > 4 or lm.rnn.bidirectional
                               or getattr(lm.rnn, 'proj_size', 0) != 0):
            return False
        return True


def _decode_pack(model, lm):
    pack = getattr(model, '_b2asr_decode_pack', None)
    if pack is None or pack.version != _DecodePack.stamp(model, lm):
        pack = _DecodePack(model, lm)
        object.__setattr__(model, '_b2asr_decode_pack', pack)
    return pack


def _native_decode(model, lm, bs, h, keys, enc_len0, decode_step):
    """All decode steps through b2asr_decode_run: two per-parity programs over double-buffered state; rows read
    their parent's state through the survivor map the beam step wrote (bs.parent), so nothing is gathered."""
    import ctypes
    att = model.attention
    pk = _decode_pack(model, lm)
    dev = h.device
    n, V = bs.n, bs.V
    Tp, D = h.shape[1], h.shape[2]
    A = keys.shape[-1]
    loc = att.mode == 'loc'
    pr = att._params()
    f32 = dict(dtype=torch.float32, device=dev)
    nl = len(pk.dec_w)
    lens_dev = torch.full((n,), enc_len0, dtype=torch.int32, device=dev)
    dec_h = [[torch.zeros(n, pk.H4, **f32) for _ in range(2)] for _ in range(nl)]
    dec_c = [[torch.zeros(n, pk.H4, **f32) for _ in range(2)] for _ in range(nl)]
    att_buf = [torch.zeros(n, Tp, **f32) for _ in range(2)]
    if loc:
        for b in att_buf:
            b[:, :enc_len0] = 1.0 / enc_len0                                   # asr.py:444-449
    ctx = torch.zeros(n, pk.D4, **f32)
    nct = _lib.lib().b2asr_linear_lse_ctas(V)
    char_logits = torch.empty(n, V, **f32)
    char_lse = torch.empty(n, **f32)
    char_part = torch.empty(n, nct, 2, **f32)
    tickets = torch.zeros(2, dtype=torch.int32, device=dev)
    keep = [lens_dev, dec_h, dec_c, att_buf, ctx, char_logits, char_lse, char_part, tickets]
    if pk.lm is not None:
        L = pk.lm
        lm_h = [[torch.zeros(n, L['HL4'], **f32) for _ in range(2)] for _ in range(L['layers'])]
        lm_c = [[torch.zeros(n, L['HL4'], **f32) for _ in range(2)] for _ in range(L['layers'])]
        lm_logits = torch.empty(n, V, **f32)
        lm_lse = torch.empty(n, **f32)
        lm_part = torch.empty(n, nct, 2, **f32)
        keep += [lm_h, lm_c, lm_logits, lm_lse, lm_part]
    p = _lib.ptr
    C, Kw = (pr['conv'].shape[0], (pr['conv'].shape[-1] - 1) // 2) if loc else (0, 0)
    conv = pr['conv'].contiguous() if loc else None
    lproj = pr['lproj'].contiguous() if loc else None
    gen_w = pr['gen_w'].contiguous() if loc else None
    keep += [conv, lproj, gen_w]

    def cell(x1, K1, x2, W, base, base_row, b, c_prev, c_new, h_new, H, H4, after, mapped=True):
        cs = _lib.CellStep()
        cs.B, cs.H, cs.K1, cs.K2 = n, H, K1, H4
        cs.x1, cs.x1_ld = p(x1), (x1.stride(0) if x1 is not None else 0)
        cs.x2, cs.x2_ld = p(x2), x2.stride(0)
        cs.map = p(bs.parent) if mapped else 0
        cs.W, cs.w_ld = p(W), W.stride(0)
        cs.base, cs.base_ld, cs.base_row = p(base), (base.stride(0) if base is not None else 0), p(base_row)
        cs.b1, cs.b2 = p(b[0]), p(b[1])
        cs.c_prev, cs.c, cs.c_ld = p(c_prev), p(c_new), H4
        cs.h, cs.h_ld = p(h_new), H4
        cs.after_linear = after
        return cs

    progs = (_lib.DecodeProg * 2)()
    ios = []
    for src in (0, 1):
        dst = 1 - src
        pg = progs[src]
        a = pg.attn
        # cluster hint: with the RNN-LM chain running next to it on the second stream the attention launch is better
        # off with half the CTAs (B2ASR_ATT_FWD_CS A/B at M4: 4 CTAs per row 218 us per step, the automatic 8: 228 us)
        hint = 4 if (pk.lm is not None and Tp >= 128 and 8 * n > 148) else 0
        a.mode, a.C, a.Kw, a.B, a.T, a.D, a.A, a.dec, a.cluster = _attn.MODES[att.mode], C, Kw, n, Tp, D, A, pk.H, hint
        a.scale = 2.0
        a.s, a.s_stride = p(dec_h[0][src]), pk.H4
        a.Wphi, a.keys, a.k_bstride, a.h, a.h_bstride = p(pr['phi']), p(keys), 0, p(h), 0
        a.lens, a.prev_att, a.row_map = p(lens_dev), (p(att_buf[src]) if loc else 0), p(bs.parent)
        a.Wc, a.Wl, a.wg, a.bg = p(conv), p(lproj), p(gen_w), (p(pr['gen_b']) if loc else 0)
        a.att, a.ctx, a.ctx_stride = p(att_buf[dst]), p(ctx), pk.D4
        last = bs.last[src]
        cells = []
        for l in range(nl):                                                   # Speller (asr.py:352-357)
            if l == 0:
                cells.append(cell(ctx, pk.D4, dec_h[0][src], pk.dec_w[0], pk.dec_table, last, pk.dec_b[0],
                                  dec_c[0][src], dec_c[0][dst], dec_h[0][dst], pk.H, pk.H4, 0))
            else:
                cells.append(cell(dec_h[l - 1][dst], pk.H4, dec_h[l][src], pk.dec_w[l], None, None, pk.dec_b[l],
                                  dec_c[l][src], dec_c[l][dst], dec_h[l][dst], pk.H, pk.H4, 0))
        lin = pg.linear[0]                                                    # char_trans + lse (asr.py:210)
        lin.B, lin.K, lin.V = n, pk.H4, V
        lin.x, lin.x_ld, lin.W, lin.w_ld, lin.bias = p(dec_h[nl - 1][dst]), pk.H4, p(pk.char_w), pk.H4, p(pk.char_b)
        lin.logits, lin.out_ld, lin.lse, lin.part, lin.ticket = p(char_logits), V, p(char_lse), p(char_part), tickets.data_ptr()
        pg.n_linear = 1
        lmlog = lmlse = None
        if pk.lm is not None:                                                 # RNN-LM step (rnnlm.py:19-25)
            L = pk.lm
            for l in range(L['layers']):
                if l == 0:
                    cells.append(cell(None, 0, lm_h[0][src], L['w'][0], L['table'], last, L['b'][0],
                                      lm_c[0][src], lm_c[0][dst], lm_h[0][dst], L['HL'], L['HL4'], 1))
                else:
                    cells.append(cell(lm_h[l - 1][dst], L['HL4'], lm_h[l][src], L['w'][l], None, None, L['b'][l],
                                      lm_c[l][src], lm_c[l][dst], lm_h[l][dst], L['HL'], L['HL4'], 1))
            lin = pg.linear[1]
            lin.B, lin.K, lin.V = n, L['HL4'], V
            lin.x, lin.x_ld, lin.W, lin.w_ld, lin.bias = p(lm_h[-1][dst]), L['HL4'], p(L['out_w']), L['HL4'], p(L['out_b'])
            lin.logits, lin.out_ld, lin.lse, lin.part, lin.ticket = p(lm_logits), V, p(lm_lse), p(lm_part), tickets.data_ptr() + 4
            pg.n_linear = 2
            lmlog, lmlse = lm_logits, lm_lse
        pg.n_cells = len(cells)
        for i, cs in enumerate(cells):
            pg.cells[i] = cs
        io = bs._make_io(src, char_logits, lmlog, char_lse, lmlse)
        ios.append(io)
        pg.beam = ctypes.pointer(io)
    per_step = 1 + len(cells) + progs[0].n_linear + 1
    _lib.call("b2asr_decode_run", per_step * decode_step, _lib.lib().b2asr_decode_run, progs, 0, int(decode_step),
              _lib.stream_ptr())
    bs.cur = decode_step & 1
    bs.steps_done = decode_step
    bs._keep = (keep, progs, ios, pk)                                          # alive until the results are read


def beam_decode_many(model, utterances, decode_beam_size, streams=4):
    """Several utterances in flight on one GPU.  utterances: iterable of (audio_feature [1,T,D], decode_step, state_len);
    returns their Hypothesis lists in order.  One decode is a latency-bound chain that keeps a fraction of the SMs busy
    (the beam step runs `beam` CTAs), so up to `streams` utterances are enqueued on CUDA streams of their own -- the
    reference's answer to the same observation is `--njobs` worker processes (README.md:123).  Each utterance's search
    is the one beam_decode runs; utterances whose model configuration the native loop does not take are decoded one
    after the other."""
    utterances = list(utterances)
    if not utterances:
        return []
    dev = utterances[0][0].device
    pool = [torch.cuda.Stream(dev) for _ in range(max(1, min(int(streams), 8)))]
    main = torch.cuda.current_stream(dev)
    results = [None] * len(utterances)
    pending = {}                                                     # stream slot -> (index, BeamSearch)

    def finish(slot):
        i, bs = pending.pop(slot)
        with torch.cuda.stream(pool[slot]):
            results[i] = bs.results(emb=model.embed)

    for i, (x, decode_step, state_len) in enumerate(utterances):
        slot = i % len(pool)
        if slot in pending:
            finish(slot)                                             # at most len(pool) utterances hold buffers
        pool[slot].wait_stream(main)
        with torch.cuda.stream(pool[slot]):
            x.record_stream(pool[slot])
            out = beam_decode(model, x, decode_step, state_len, decode_beam_size, _defer=True)
        if isinstance(out, BeamSearch):
            pending[slot] = (i, out)
        else:
            results[i] = out                                         # decoded synchronously (per-step path) or empty
    for slot in sorted(pending):
        finish(slot)
    for st in pool:
        main.wait_stream(st)
    from . import _enc
    _enc.check_status(dev)
    return results


def beam_decode(model, audio_feature, decode_step, state_len, decode_beam_size, use_graph=None, _defer=False):
    """Seq2Seq.beam_decode (asr.py:155-258) for one utterance; returns the top `decode_beam_size`
    Hypothesis objects sorted by mean score."""
    assert audio_feature.shape[0] == 1
    assert model.training is False
    if use_graph is None:
        # measured on B200 (profiles/): the step is GPU-bound, replay saves nothing over eager launches and
        # capture + instantiation costs 0.1-0.5 s per utterance shape, so graphs are opt-in
        use_graph = os.environ.get('B2ASR_DECODE_GRAPH', '0') == '1'
    if not audio_feature.is_cuda:
        raise RuntimeError('beam_decode runs on the sm_100a kernels only (no CPU path); move the model and input to a CUDA device')
    n = int(decode_beam_size)
    model.decode_beam_size = n
    dev = audio_feature.device
    with torch.no_grad():
        h, enc_len = model.encoder(audio_feature, state_len)
        if decode_step == 0:
            decode_step = int(enc_len[0])
        if not model.joint_att:
            return []                                            # asr.py:186 guards the whole search
        x_ctc = None
        if model.joint_ctc:
            x_ctc = F.log_softmax(model.ctc_layer(h), dim=-1)[0]                # asr.py:180
        lm_w = float(getattr(model, 'decode_lm_weight', 0) or 0)
        lm = getattr(model, 'rnn_lm', None) if lm_w > 0 else None
        ctc_w = float(model.ctc_weight) if model.joint_ctc else 0.0
        bs = BeamSearch(model.char_dim, n, decode_step, dev, x_ctc=x_ctc, ctc_weight=ctc_w, lm_weight=lm_w)

        att = model.attention
        dec = model.decoder
        loc = att.mode == 'loc'
        h = h.contiguous()
        keys = (att.project_keys(h) if att.proj else h).contiguous()
        Tp, D = h.shape[1], h.shape[2]
        from . import _enc
        if _DecodePack.supported(model, lm) and os.environ.get('B2ASR_DECODE_NATIVE', '1') != '0' and decode_step > 0:
            # every step's kernels enqueued by ONE C call (csrc/decode.cu); the per-step path below stays for cells
            # the tile kernels do not take (GRU) and for A/B runs
            _native_decode(model, lm, bs, h, keys, int(enc_len[0]), int(decode_step))
            if _defer:
                return bs                                        # beam_decode_many reads the results later
            out = bs.results(emb=model.embed)
            _enc.check_status(dev)
            return out
        p = att._params()
        lens_dev = torch.full((n,), int(enc_len[0]), dtype=torch.int32, device=dev)
        layers = [getattr(dec, 'layer' + str(i)) for i in range(dec.layer)]
        dec_h = [torch.zeros(n, dec.dim, device=dev) for _ in layers]
        dec_c = [torch.zeros(n, dec.dim, device=dev) for _ in layers]
        prev_att = torch.zeros(n, Tp, device=dev)
        if loc:
            prev_att[:, :int(enc_len[0])] = 1.0 / int(enc_len[0])             # asr.py:444-449
        att_buf = torch.empty(n, Tp, device=dev)
        ctx_buf = torch.empty(n, D, device=dev)
        if lm is not None:
            nl, hd = lm.rnn.num_layers, lm.h_dim
            lm_h = torch.zeros(nl, n, hd, device=dev)
            lm_c = torch.zeros(nl, n, hd, device=dev)

        def one_step(src):
            last = bs.last[src]
            _attn.step_forward(att.mode, dec_h[0], p['phi'], keys, h, lens_dev, prev_att if loc else None,
                               p['conv'], p['lproj'], p['gen_w'], p['gen_b'], 2.0, shared_memory=True,
                               out=(att_buf, ctx_buf))
            x = torch.cat([model.embed(last), ctx_buf], dim=-1)               # asr.py:208
            new_h, new_c = [], []
            hh, cc = layers[0](x, (dec_h[0], dec_c[0]))
            new_h.append(hh); new_c.append(cc)
            for l in range(1, len(layers)):
                hh, cc = layers[l](new_h[l - 1], (dec_h[l], dec_c[l]))
                new_h.append(hh); new_c.append(cc)
            att_logp = F.log_softmax(model.char_trans(new_h[-1]), dim=-1)     # asr.py:210
            lm_logp = None
            if lm is not None:
                (nh, nc), logit = lm.step(last, (lm_h, lm_c))                 # asr.py:232-235, whole beam at once
                lm_logp = F.log_softmax(logit, dim=-1)
            bs.launch(src, att_logp.contiguous(), lm_logp)
            par = bs.parent
            for l in range(len(layers)):
                torch.index_select(new_h[l], 0, par, out=dec_h[l])
                torch.index_select(new_c[l], 0, par, out=dec_c[l])
            if loc:
                torch.index_select(att_buf, 0, par, out=prev_att)
            if lm is not None:
                torch.index_select(nh, 1, par, out=lm_h)
                torch.index_select(nc, 1, par, out=lm_c)

        eager = min(decode_step, 2)
        for t in range(eager):
            one_step(t & 1)
            bs.flip()
        if decode_step > eager:
            if use_graph and decode_step - eager >= 4:
                graphs = []
                for src in (0, 1):
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        one_step(src)
                    graphs.append(g)
                for t in range(eager, decode_step):
                    graphs[t & 1].replay()
                    _lib.LAUNCHES += 2
                    bs.flip()
            else:
                for t in range(eager, decode_step):
                    one_step(t & 1)
                    bs.flip()
        return bs.results(emb=model.embed)
