"""Host-side mirror of the reference's src/postprocess.py (index->token mapping, metrics, the beam
hypothesis record).  Pure host code, same names and return conventions.

  Mapper        postprocess.py:7-41     Hypothesis   postprocess.py:44-119
  cal_acc       postprocess.py:121-133  cal_cer      postprocess.py:135-146
  draw_att      postprocess.py:149-155  collapse_phn postprocess.py:158-169   trim_eos postprocess.py:171-177

`editdistance` is absent from this image; `edit_distance` (same `eval(a, b)` contract) runs the
Levenshtein DP in the library's host C++ (csrc/host.cu).  The device-resident beam search (beam.py) does not build Hypothesis objects while it
searches; it rebuilds them from the device history after the last step, with the same fields.
"""
import os
import pickle

import numpy as np
import torch

from . import _lib


def edit_distance(a, b):
    """Levenshtein distance between two sequences of hashable items -- `editdistance.eval(a, b)` as eval.py:17-18 and
    cal_cer (postprocess.py:135-146) call it.  The items are numbered and the DP runs in the C++ of csrc/host.cu."""
    import ctypes
    ids = {}
    ia = [ids.setdefault(x, len(ids)) for x in a]
    ib = [ids.setdefault(x, len(ids)) for x in b]
    arr_a = (ctypes.c_int32 * max(len(ia), 1))(*ia)
    arr_b = (ctypes.c_int32 * max(len(ib), 1))(*ib)
    d = _lib.lib().b2asr_edit_distance(ctypes.cast(arr_a, ctypes.c_void_p), len(ia), ctypes.cast(arr_b, ctypes.c_void_p), len(ib))
    if d < 0:
        raise ValueError("b2asr_edit_distance: bad argument")
    return int(d)


_SPECIAL = ('<sos>', '<eos>')
_UNITS = (('subword', lambda keys, n: any('▁' in k for k in keys)),      # unit detection, postprocess.py:14-22:
          ('phone', lambda keys, n: any('#' in k for k in keys)),         # sentencepiece marker, TIMIT's 'h#',
          ('char', lambda keys, n: n < 50),                               # a small table, else words
          ('word', lambda keys, n: True))


class Mapper():
    """index -> token table of a corpus directory (`<data_path>/mapping.pkl`, token -> index with <sos>=0, <eos>=1;
    reference postprocess.py:7-41).  Attributes kept: mapping, r_mapping, unit."""

    def __init__(self, file_path):
        with open(os.path.join(file_path, 'mapping.pkl'), 'rb') as fp:
            self.mapping = pickle.load(fp)
        self.r_mapping = dict((idx, tok) for tok, idx in self.mapping.items())
        keys = list(self.mapping.keys())
        self.unit = next(name for name, test in _UNITS if test(keys, len(keys)))

    def get_dim(self):
        return len(self.mapping)

    def translate(self, seq, return_string=False):
        """Tokens of `seq` up to and including the first <eos>; with return_string the text the unit implies
        (subwords glued with '▁' -> ' ', words / folded phones space-separated, characters concatenated)."""
        tokens = [self.r_mapping[i] for i in trim_eos(seq)]
        if not return_string:
            return tokens
        if self.unit == 'phone':
            tokens = collapse_phn(tokens)
        text = (' ' if self.unit in ('word', 'phone') else '').join(tokens)
        for sp in _SPECIAL:
            text = text.replace(sp, '')
        if self.unit == 'subword':
            text = text.replace('▁', ' ')
        return text.lstrip() if self.unit in ('subword', 'word') else text


class Hypothesis:
    '''Hypothesis for beam search decoding: label/score history plus the decoder, CTC, LM and attention
    state the next expansion needs (postprocess.py:44-119).'''

    def __init__(self, decoder_state, emb, output_seq=[], output_scores=[],
                 lm_state=None, ctc_state=None, ctc_prob=0.0, att_map=None):
        assert len(output_seq) == len(output_scores)
        self.decoder_state = decoder_state
        self.att_map = att_map
        self.lm_state = lm_state
        self.output_seq = output_seq
        self.output_scores = output_scores
        self.ctc_state = ctc_state
        self.ctc_prob = ctc_prob
        self.emb = emb

    def avgScore(self):
        '''Mean log-probability of the hypothesis: left-to-right sum / count (postprocess.py:71-74)'''
        assert len(self.output_scores) != 0
        return sum(self.output_scores) / len(self.output_scores)

    def addTopk(self, topi, topv, decoder_state, att_map=None,
                lm_state=None, ctc_state=None, ctc_prob=0.0, ctc_candidates=[]):
        """Expand by the picks (topi, topv of shape [1, k]): <eos> (= 1) finalises THIS hypothesis in place, every
        other token becomes a child that copies the history and, with joint CTC, takes the prefix state of its
        candidate slot (postprocess.py:76-106).  Returns (self if finalised else None, children).
        Host path kept for API compatibility; the device beam search does not call it (beam.py)."""
        children, eos_score = [], None
        for tok, score in zip(topi[0], topv[0]):
            t = int(tok.item())
            if t == 1:
                eos_score = score.cpu()
                continue
            state_c = prob_c = None
            if ctc_state is not None:
                slot = ctc_candidates.index(t)
                state_c, prob_c = ctc_state[slot, :, :], ctc_prob[slot]
            children.append(Hypothesis(decoder_state, self.emb,
                                       output_seq=list(self.output_seq) + [tok.cpu()],
                                       output_scores=list(self.output_scores) + [score.cpu()],
                                       lm_state=lm_state, ctc_state=state_c, ctc_prob=prob_c, att_map=att_map))
        if eos_score is None:
            return None, children
        self.output_seq.append(torch.tensor(1))
        self.output_scores.append(eos_score)
        return self, children

    @property
    def outIndex(self):
        return [i.item() for i in self.output_seq]

    @property
    def last_char_idx(self):
        idx = self.output_seq[-1] if len(self.output_seq) != 0 else 0
        return torch.LongTensor([[idx]])

    @property
    def last_char(self):
        idx = self.output_seq[-1] if len(self.output_seq) != 0 else 0
        return self.emb(torch.LongTensor([idx]).to(next(self.emb.parameters()).device))


def cal_acc(pred, label, hyp=None):
    '''Mean per-utterance token accuracy up to the first padding label (postprocess.py:121-133),
    computed on the device; one scalar crosses to the host.  `hyp` = the row argmax of `pred` when the caller
    already has it (the CE kernel returns it), saving the pass over the logits.'''
    if hyp is None:
        hyp = pred.detach().argmax(dim=-1)
    label = label.to(hyp.device)
    live = torch.cumprod((label != 0).to(torch.int64), dim=-1) > 0       # stop at the first 0, as the loop does
    n = min(hyp.shape[1], label.shape[1])
    correct = ((hyp[:, :n] == label[:, :n]) & live[:, :n]).sum(-1).to(torch.float64)
    total = live[:, :n].sum(-1).to(torch.float64)
    return float((correct / total).mean())


def cal_cer(pred, label, mapper, get_sentence=False, argmax=True):
    """Mean word-level error rate of the greedy transcripts against the references (postprocess.py:135-146): both
    sides go through `mapper.translate(..., return_string=True)`, are split on blanks, and the Levenshtein distance is
    divided by the reference's length.  get_sentence=True returns the two lists of strings instead."""
    ids = pred.detach().argmax(dim=-1).cpu() if argmax else pred
    hyp_txt = [mapper.translate(row, return_string=True) for row in ids]
    ref_txt = [mapper.translate(row, return_string=True) for row in label.cpu()]
    if get_sentence:
        return hyp_txt, ref_txt
    rates = []
    for h, r in zip(hyp_txt, ref_txt):
        ref_words = r.split(' ')
        rates.append(float(edit_distance(h.split(' '), ref_words)) / len(ref_words))
    return sum(rates) / len(rates)


def draw_att(att_list, hyp_txt):
    """[3, steps, T'] image tensors of the first head's alignments, cut at each utterance's <eos> step
    (postprocess.py:149-155); `hyp_txt` are the decoder logits the lengths are read from."""
    lengths = [len(trim_eos(row)) for row in hyp_txt.detach().argmax(dim=-1).cpu()]
    images = []
    for att, n in zip(att_list[0], lengths):
        grey = att.detach().cpu()[:n]
        images.append(grey.unsqueeze(0).expand(3, -1, -1).contiguous())
    return images


_PHN_REDUCE = {
    "b": "b", "bcl": "h#", "d": "d", "dcl": "h#", "g": "g", "gcl": "h#", "p": "p", "pcl": "h#", "t": "t", "tcl": "h#",
    "k": "k", "kcl": "h#", "dx": "dx", "q": "q", "jh": "jh", "ch": "ch", "s": "s", "sh": "sh", "z": "z", "zh": "sh",
    "f": "f", "th": "th", "v": "v", "dh": "dh", "m": "m", "n": "n", "ng": "ng", "em": "m", "en": "n", "eng": "ng",
    "nx": "n", "l": "l", "r": "r", "w": "w", "y": "y", "hh": "hh", "hv": "hh", "el": "l", "iy": "iy", "ih": "ih",
    "eh": "eh", "ey": "ey", "ae": "ae", "aa": "aa", "aw": "aw", "ay": "ay", "ah": "ah", "ao": "aa", "oy": "oy",
    "ow": "ow", "uh": "uh", "uw": "uw", "ux": "uw", "er": "er", "ax": "ah", "ix": "ih", "axr": "er", "ax-h": "ah",
    "pau": "h#", "epi": "h#", "h#": "h#", "<sos>": "<sos>", "<unk>": "<unk>", "<eos>": "<eos>"}


def collapse_phn(seq):
    '''TIMIT 61 -> 39 phone folding (postprocess.py:158-169).'''
    return [_PHN_REDUCE[c] for c in seq]


def trim_eos(seqence):
    """ints of `seqence` up to and including the first <eos> (= 1) (postprocess.py:171-177)."""
    out = []
    for tok in seqence:
        out.append(int(tok))
        if out[-1] == 1:
            break
    return out
