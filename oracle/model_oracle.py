"""CPU oracle for the whole hybrid train step.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.  bench.py's cpu_baseline leg
and `--impl reference` arm time this file on the host cores (kind "port"): /root/reference does not
exist on the GPU box, so the reference's own Python cannot travel.

Plain-PyTorch restatement, driven by a reference-compatible state_dict, of
  * Listener / RNNLayer / VGGExtractor forward      reference src/asr.py:311-317, 475-501, 543-558
  * the teacher-forced / greedy decoder loop        src/asr.py:72-112
  * the hybrid loss of the train step               src/solver.py:144-163
  * backward, clip, NaN guard, optimizer step       src/solver.py:177-182
Attention steps come from oracle/attention_oracle.py, CTC from torch.nn.CTCLoss exactly as the
reference calls it.  Pinned against the real reference classes in tests/test_oracle_pins.py.
"""
import math
import random

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.distributions.categorical import Categorical
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

from . import attention_oracle as ao


class OracleModel(nn.Module):
    """Holds stock torch modules under the reference's parameter names; forward is `run()` below."""

    def __init__(self, input_dim, vocab, cfg):
        super().__init__()
        enc, att, dec = cfg['encoder'], cfg['attention'], cfg['decoder']
        self.cfg = cfg
        self.vgg = 'VGG' in enc['enc_type']
        self.bidir = 'BiRNN' in enc['enc_type']
        self.rates = [int(v) for v in enc['sample_rate'].split('_')]
        self.style = enc['sample_style']
        dims = [int(v) for v in enc['dim'].split('_')]
        self.encoder = nn.Module()
        d = input_dim
        if self.vgg:
            ch, fq = (d // 13, 13) if d % 13 == 0 else (d // 40, 40)
            self.in_ch, self.freq = ch, fq
            v = nn.Module()
            v.conv1, v.conv2 = nn.Conv2d(ch, 64, 3, padding=1), nn.Conv2d(64, 64, 3, padding=1)
            v.conv3, v.conv4 = nn.Conv2d(64, 128, 3, padding=1), nn.Conv2d(128, 128, 3, padding=1)
            self.encoder.vgg_extractor = v
            d = (fq // 4) * 128
        for l, (h, r) in enumerate(zip(dims, self.rates)):
            layer = nn.Module()
            layer.layer = getattr(nn, enc['rnn_cell'].upper())(d, h, bidirectional=self.bidir, batch_first=True)
            setattr(self.encoder, 'layer%d' % l, layer)
            d = h * (2 if self.bidir else 1) * (r if self.style == 'concat' else 1)
            setattr(self.encoder, 'proj%d' % l, nn.Linear(d, d))
        self.enc_dim = d
        w = cfg['optimizer']['joint_ctc']
        self.joint_ctc, self.joint_att, self.ctc_weight = w > 0, w < 1, w
        if self.joint_att:
            self.mode = att['att_mode'].lower()
            a = nn.Module()
            a.phi = nn.Linear(dec['dim'], att['dim'], bias=False)
            a.psi = nn.Linear(d, att['dim'])
            if self.mode == 'loc':
                a.loc_conv = nn.Conv1d(1, 10, 201, padding=100, bias=False)
                a.loc_proj = nn.Linear(10, att['dim'], bias=False)
                a.gen_energy = nn.Linear(att['dim'], 1)
            self.attention = a
            self.decoder = nn.Module()
            self.n_dec = dec['layer']
            for i in range(self.n_dec):
                setattr(self.decoder, 'layer%d' % i, getattr(nn, dec['rnn_cell'])(d + dec['dim'] if i == 0 else dec['dim'], dec['dim']))
            self.dec_dim = dec['dim']
            self.embed = nn.Embedding(vocab, dec['dim'])
            self.char_trans = nn.Linear(dec['dim'], vocab)
        if self.joint_ctc:
            self.ctc_layer = nn.Linear(d, vocab)

    # ---- encoder (asr.py:311-317) ---------------------------------------------------------------
    def encode(self, x, lens):
        if self.vgg:
            v = self.encoder.vgg_extractor
            lens = [n // 4 for n in lens]
            x = x[:, :x.shape[1] - x.shape[1] % 4]
            b, t, _ = x.shape
            x = x.reshape(b, t, self.in_ch, self.freq).transpose(1, 2)
            x = F.max_pool2d(F.relu(v.conv2(F.relu(v.conv1(x)))), 2, 2)
            x = F.max_pool2d(F.relu(v.conv4(F.relu(v.conv3(x)))), 2, 2)
            x = x.transpose(1, 2).contiguous().view(b, x.shape[2], -1)
        for l, r in enumerate(self.rates):
            packed = pack_padded_sequence(x, lens, batch_first=True)
            out, _ = getattr(self.encoder, 'layer%d' % l).layer(packed)
            x, ln = pad_packed_sequence(out, batch_first=True)
            lens = ln.tolist()
            if r > 1:
                if self.style == 'drop':
                    x = x[:, ::r]
                else:
                    b, t, f = x.shape
                    x = x[:, :t - t % r].contiguous().view(b, t // r, f * r)
                lens = [int(n / r) for n in lens]
            x = torch.tanh(getattr(self.encoder, 'proj%d' % l)(x))
        return x, lens

    # ---- decoder loop (asr.py:72-112) -------------------------------------------------------------
    def run(self, x, decode_step, tf_rate=0.0, teacher=None, state_len=None):
        h, lens = self.encode(x, state_len)
        ctc_out = self.ctc_layer(h) if self.joint_ctc else None
        att_out = maps = None
        if self.joint_att:
            a = self.attention
            B = x.shape[0]
            w = dict(phi=a.phi.weight)
            if self.mode == 'loc':
                w.update(loc_conv=a.loc_conv.weight, loc_proj=a.loc_proj.weight, gen_w=a.gen_energy.weight, gen_b=a.gen_energy.bias)
            keys = ao.project_keys(h, a.psi.weight, a.psi.bias)
            prev = ao.uniform_alignment(lens, h.shape[1], h.dtype)
            hs = [torch.zeros(B, self.dec_dim, dtype=h.dtype) for _ in range(self.n_dec)]
            cs = [torch.zeros(B, self.dec_dim, dtype=h.dtype) for _ in range(self.n_dec)]
            emb_t = self.embed(teacher) if teacher is not None else None
            last = self.embed(torch.zeros(B, dtype=torch.long))
            outs, maps = [], []
            for t in range(decode_step):
                if self.mode == 'loc':
                    alpha, ctx = ao.loc_step(hs[0], h, keys, lens, prev, w)
                    prev = alpha
                else:
                    alpha, ctx = ao.dot_step(hs[0], h, keys, lens, w)
                hs[0], cs[0] = self.decoder.layer0(torch.cat([last, ctx], -1), (hs[0], cs[0]))
                for i in range(1, self.n_dec):
                    hs[i], cs[i] = getattr(self.decoder, 'layer%d' % i)(hs[i - 1], (hs[i], cs[i]))
                logit = self.char_trans(hs[-1])
                if emb_t is not None:
                    if random.random() <= tf_rate:
                        last = emb_t[:, t + 1]
                    else:
                        last = self.embed(Categorical(F.softmax(logit, -1)).sample())
                else:
                    last = self.embed(logit.argmax(-1))
                outs.append(logit)
                maps.append(alpha.detach())
            att_out = torch.stack(outs, 1)
            maps = [torch.stack(maps, 1)]
        return ctc_out, lens, att_out, maps


def hybrid_loss(ctc_pred, att_pred, y, state_len, ctc_weight):
    """solver.py:144-163: per-utterance-normalised CE on the attention branch + CTCLoss 'mean'."""
    ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
    label = y[:, 1:ans_len + 1].contiguous()
    att_loss = ctc_loss = 0
    if ctc_weight < 1:
        b, t, c = att_pred.shape
        ce = F.cross_entropy(att_pred.reshape(b * t, c), label.reshape(-1), ignore_index=0, reduction='none')
        att_loss = torch.mean(ce.view(b, t).sum(-1) / torch.sum(y != 0, dim=-1).to(att_pred.dtype))
    if ctc_weight > 0:
        tl = torch.sum(y != 0, dim=-1)
        ctc_loss = nn.CTCLoss(blank=0, reduction='mean')(F.log_softmax(ctc_pred.transpose(0, 1), dim=-1), label,
                                                         torch.LongTensor(state_len), tl)
    return (1 - ctc_weight) * att_loss + ctc_weight * ctc_loss, att_loss, ctc_loss


def infer_lengths(x):
    """solver.py:134-136: valid frames = frames whose feature sum is non-zero."""
    return [int(v) for v in (x.sum(-1) != 0).sum(-1)]


def train_step(model, opt, x, y, tf_rate=1.0, grad_clip=5.0):
    """Body of Trainer.exec for one batch (solver.py:132-182) on CPU; returns (loss, grad_norm)."""
    state_len = infer_lengths(x)
    ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
    opt.zero_grad()
    ctc_pred, enc_len, att_pred, _ = model.run(x, ans_len, tf_rate=tf_rate, teacher=y, state_len=state_len)
    loss, _, _ = hybrid_loss(ctc_pred, att_pred, y, enc_len, model.ctc_weight)
    loss.backward()
    gn = torch.nn.utils.clip_grad_norm_(model.parameters(), grad_clip)
    if not math.isnan(float(gn)):
        opt.step()
    return loss.detach(), gn
