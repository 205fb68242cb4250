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


class Mapper():
    '''Mapper for index2token'''

    def __init__(self, file_path):
        with open(os.path.join(file_path, 'mapping.pkl'), 'rb') as fp:
            self.mapping = pickle.load(fp)
        self.r_mapping = {v: k for k, v in self.mapping.items()}
        symbols = ''.join(list(self.mapping.keys()))
        if '▁' in symbols:
            self.unit = 'subword'
        elif '#' in symbols:
            self.unit = 'phone'
        elif len(self.mapping) < 50:
            self.unit = 'char'
        else:
            self.unit = 'word'

    def get_dim(self):
        return len(self.mapping)

    def translate(self, seq, return_string=False):
        new_seq = [self.r_mapping[c] for c in trim_eos(seq)]
        if return_string:
            if self.unit == 'subword':
                new_seq = ''.join(new_seq).replace('<sos>', '').replace('<eos>', '').replace('▁', ' ').lstrip()
            elif self.unit == 'word':
                new_seq = ' '.join(new_seq).replace('<sos>', '').replace('<eos>', '').lstrip()
            elif self.unit == 'phone':
                new_seq = ' '.join(collapse_phn(new_seq)).replace('<sos>', '').replace('<eos>', '')
            elif self.unit == 'char':
                new_seq = ''.join(new_seq).replace('<sos>', '').replace('<eos>', '')
        return new_seq


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
        '''Expand with the given top-k: <eos>(=1) finalises THIS hypothesis in place, every other token
        spawns a child (postprocess.py:76-106).  Returns (self or None, children).'''
        new_hypothesis = []
        term_score = None
        ctc_s, ctc_p = None, None
        for i in range(len(topi[0])):
            if topi[0][i].item() == 1:
                term_score = topv[0][i].cpu()
                continue
            idxes = self.output_seq[:]
            scores = self.output_scores[:]
            idxes.append(topi[0][i].cpu())
            scores.append(topv[0][i].cpu())
            if ctc_state is not None:
                idx = ctc_candidates.index(topi[0][i].item())
                ctc_s = ctc_state[idx, :, :]
                ctc_p = ctc_prob[idx]
            new_hypothesis.append(Hypothesis(decoder_state, self.emb, output_seq=idxes, output_scores=scores,
                                             lm_state=lm_state, ctc_state=ctc_s, ctc_prob=ctc_p, att_map=att_map))
        if term_score is not None:
            self.output_seq.append(torch.tensor(1))
            self.output_scores.append(term_score)
            return self, new_hypothesis
        return None, new_hypothesis

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
    if argmax:
        pred = np.argmax(pred.cpu().detach(), axis=-1)
    label = label.cpu()
    pred = [mapper.translate(p, return_string=True) for p in pred]
    label = [mapper.translate(l, return_string=True) for l in label]
    if get_sentence:
        return pred, label
    eds = [float(edit_distance(p.split(' '), l.split(' '))) / len(l.split(' ')) for p, l in zip(pred, label)]
    return sum(eds) / len(eds)


def draw_att(att_list, hyp_txt):
    '''Only the first attention head is drawn (postprocess.py:149-155).'''
    attmaps = []
    for att, hyp in zip(att_list[0], np.argmax(hyp_txt.cpu().detach(), axis=-1)):
        att_len = len(trim_eos(hyp))
        att = att.detach().cpu()
        attmaps.append(torch.stack([att, att, att], dim=0)[:, :att_len, :])
    return attmaps


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
    new_pred = []
    for char in seqence:
        new_pred.append(int(char))
        if char == 1:
            break
    return new_pred
