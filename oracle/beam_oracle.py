"""CPU oracle for kernel (3), part 2: the joint CTC / attention / RNN-LM beam search.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Restates the search loop of `Seq2Seq.beam_decode` (reference src/asr.py:201-258)
and the hypothesis bookkeeping of `Hypothesis` (src/postprocess.py:44-119) over
plain records, with the neural parts abstracted behind `step_fn` so the same
per-hypothesis log-probabilities can be fed to the CUDA beam step:

    step_fn(entry) -> (att_logp [V] float32 tensor,      # log_softmax(char_trans(dec_out)), asr.py:210
                       lm_logp  [V] float32 tensor|None, # log_softmax(lm_out),              asr.py:235
                       payload)                          # opaque decoder/attention/LM state for children

Arithmetic that decides which tokens survive, all reproduced in float32 in the
reference's order of operations:
  * CTC candidates = top int(1.5*beam) of att_logp (asr.py:15,160,215-216);
  * ctc_char = psi - parent.ctc_prob scattered into a -1e6 vector (asr.py:222-227);
  * fused = (1-w)*att + w*ctc_vec, then fused[0] = -1e7 (asr.py:228-229; only when
    joint CTC is on), then fused += lm_w * lm_logp (asr.py:232-235) -- separate
    float32 multiply and add roundings, no FMA;
  * top-`beam` of fused (asr.py:238); <eos>(=1) in the top-k finalises the PARENT
    in place with that score, every other token spawns a child
    (postprocess.py:76-106);
  * children ranked by mean score = left-to-right float32 sum / count
    (postprocess.py:71-74), stable descending sort, parent order then rank order
    breaking ties (asr.py:251-252); beam==1 returns at the first <eos> (asr.py:246-247);
  * after the last step survivors are appended to the finals and the whole list is
    sorted once more (asr.py:255-258).
"""
from dataclasses import dataclass, field
from typing import Any, List, Optional

import numpy as np
import torch

from . import prefix_oracle as po

CTC_BEAM_RATIO = 1.5      # asr.py:15
EOS = 1
NEG_CTC = -1000000.0      # asr.py:225
NEG_SOS = -10000000.0     # asr.py:229


@dataclass
class Entry:
    tokens: List[int] = field(default_factory=list)
    scores: List[torch.Tensor] = field(default_factory=list)   # 0-dim float32 CPU tensors
    ctc_state: Optional[np.ndarray] = None
    ctc_prob: Any = 0
    payload: Any = None

    def mean_score(self):
        return sum(self.scores) / len(self.scores)             # postprocess.py:71-74


def expand(entry, att_logp, lm_logp, payload, beam, x_ctc, ctc_weight, lm_weight):
    """One (parent hypothesis, decode step) expansion: asr.py:210-242 + postprocess.py:76-106."""
    cur = att_logp.detach().to(torch.float32).cpu().clone().view(1, -1)
    psi = states = cands = None
    if x_ctc is not None:
        k = int(CTC_BEAM_RATIO * beam)
        cands = [int(c) for c in cur.topk(k)[1][0].numpy()]
        psi, states = po.extend_prefix(x_ctc, entry.tokens, entry.ctc_state, cands)
        ctc_char = torch.FloatTensor(psi - entry.ctc_prob)
        vec = torch.zeros_like(cur).fill_(NEG_CTC)
        for i, c in enumerate(cands):
            vec[0, c] = ctc_char[i]
        cur = (1 - ctc_weight) * cur + ctc_weight * vec
        cur[0, 0] = NEG_SOS
    if lm_weight > 0:
        cur += lm_weight * lm_logp.detach().to(torch.float32).cpu().view(1, -1)
    topv, topi = cur.topk(beam)
    children, eos_score = [], None
    for rank in range(beam):
        tok = int(topi[0, rank])
        if tok == EOS:
            eos_score = topv[0, rank].clone()
            continue
        st = pr = None
        if states is not None:
            j = cands.index(tok)
            st, pr = states[j], psi[j]
        children.append(Entry(entry.tokens + [tok], entry.scores + [topv[0, rank].clone()], st, pr, payload))
    finished = None
    if eos_score is not None:
        entry.tokens = entry.tokens + [EOS]
        entry.scores = entry.scores + [eos_score]
        finished = entry
    return finished, children, dict(cands=cands, psi=psi, topv=topv[0].clone(), topi=topi[0].clone())


def joint_beam_search(step_fn, root_payload, decode_steps, beam, x_ctc=None, ctc_weight=0.0,
                      lm_weight=0.0, trace=None):
    """asr.py:183-258.  x_ctc: [T,V] float32 CTC log-posteriors or None (joint CTC off)."""
    root = Entry(payload=root_payload)
    if x_ctc is not None:
        x_ctc = np.asarray(x_ctc, dtype=np.float32)
        root.ctc_state = po.blank_only_state(x_ctc)
    finals, live = [], [root]
    for step in range(decode_steps):
        pool = []
        for parent in live:
            att_logp, lm_logp, payload = step_fn(parent)
            done, kids, dbg = expand(parent, att_logp, lm_logp, payload, beam, x_ctc, ctc_weight, lm_weight)
            if trace is not None:
                trace.append((step, list(parent.tokens), dbg))
            if done is not None:
                finals.append(done)
                if beam == 1:
                    return finals
            pool.extend(kids)
        pool.sort(key=lambda e: e.mean_score(), reverse=True)
        live = pool[:beam]
    finals += live
    finals.sort(key=lambda e: e.mean_score(), reverse=True)
    return finals[:beam]
