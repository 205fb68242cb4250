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
        self.r_all = torch.zeros(n, T, K, 2, **f32)
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

    def _make_io(self, src, att_logp, lm_logp):
        dst = 1 - src
        io = _lib.BeamIO()
        p = _lib.ptr
        io.x_ctc = p(self.x_ctc)
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

    def results(self, emb=None):
        """asr.py:244-258: finals in finalisation order, then the survivors in rank order, one stable
        sort by mean score, best `beam`.  Returns Hypothesis objects (postprocess.py:44) whose
        output_seq / output_scores are 0-dim CPU tensors, exactly what the reference's carry."""
        h = self.history()

        def hyp(toks, scs):
            return Hypothesis(None, emb, output_seq=[torch.tensor(t) for t in toks],
                              output_scores=[torch.tensor(float(v), dtype=torch.float32) for v in scs])

        outs = []
        for i in range(h['fin_count']):
            toks, scs = self._path(h, int(h['fin_step'][i]) - 1, int(h['fin_row'][i]))
            outs.append(hyp(toks + [1], scs + [h['fin_score'][i]]))
            if self.n == 1:
                return outs[:1]                                 # asr.py:246-247
        last = self.steps_done - 1
        for row in range(self.n):
            if h['valid'][row]:
                outs.append(hyp(*self._path(h, last, row)))
        outs.sort(key=lambda o: o.avgScore(), reverse=True)
        return outs[:self.n]


def beam_decode(model, audio_feature, decode_step, state_len, decode_beam_size, use_graph=None):
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
