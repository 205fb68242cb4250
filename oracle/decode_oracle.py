"""CPU oracle for the full joint beam decode: glues model_oracle,
attention_oracle, prefix_oracle and beam_oracle together the way `Seq2Seq.beam_decode`
(reference src/asr.py:155-258) and `RNN_LM.forward` (src/rnnlm.py:19-25) do.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import attention_oracle as ao
from . import beam_oracle as bo


class OracleLM(nn.Module):
    """Embedding -> LSTM(layers) -> Linear under the reference's parameter names (rnnlm.py:10-17)."""

    def __init__(self, emb_dim, h_dim, out_dim, layers=1):
        super().__init__()
        self.emb = nn.Embedding(out_dim, emb_dim)
        self.rnn = nn.LSTM(emb_dim, h_dim, num_layers=layers, batch_first=True)
        self.out = nn.Linear(h_dim, out_dim)

    def step(self, token, hidden):
        """One token for one hypothesis (rnnlm.py:19-25 with lens=[1]); returns (hidden, logits [V])."""
        e = self.emb(torch.tensor([[int(token)]]))
        o, hidden = self.rnn(e, hidden)
        return hidden, self.out(o)[0, 0]


def beam_decode(model, lm, x, decode_step, state_len, beam, ctc_weight, lm_weight, trace=None):
    """model: model_oracle.OracleModel (eval), lm: OracleLM or None.  Returns list[beam_oracle.Entry]."""
    with torch.no_grad():
        h, lens = model.encode(x, state_len)
        if decode_step == 0:
            decode_step = int(lens[0])
        x_ctc = None
        if ctc_weight > 0:
            x_ctc = F.log_softmax(model.ctc_layer(h), dim=-1)[0].numpy()          # asr.py:180
        a = model.attention
        w = dict(phi=a.phi.weight)
        if model.mode == 'loc':
            w.update(loc_conv=a.loc_conv.weight, loc_proj=a.loc_proj.weight, gen_w=a.gen_energy.weight, gen_b=a.gen_energy.bias)
        keys = ao.project_keys(h, a.psi.weight, a.psi.bias)
        zeros = [torch.zeros(1, model.dec_dim) for _ in range(model.n_dec)]
        root = dict(hs=zeros, cs=list(zeros), prev=None, lm=None)

        def step_fn(entry):
            st = entry.payload
            hs, cs = list(st['hs']), list(st['cs'])
            if model.mode == 'loc':
                prev = st['prev'] if st['prev'] is not None else ao.uniform_alignment(lens, h.shape[1], h.dtype)
                alpha, ctx = ao.loc_step(hs[0], h, keys, lens, prev, w)
            else:
                alpha, ctx = ao.dot_step(hs[0], h, keys, lens, w)
            last = entry.tokens[-1] if entry.tokens else 0
            emb = model.embed(torch.tensor([last]))
            hs[0], cs[0] = model.decoder.layer0(torch.cat([emb, ctx], -1), (hs[0], cs[0]))
            for i in range(1, model.n_dec):
                hs[i], cs[i] = getattr(model.decoder, 'layer%d' % i)(hs[i - 1], (hs[i], cs[i]))
            att_logp = F.log_softmax(model.char_trans(hs[-1]), dim=-1)[0]          # asr.py:210
            lm_logp, lm_hidden = None, None
            if lm_weight > 0:
                lm_hidden, logits = lm.step(last, st['lm'])                       # asr.py:232-235
                lm_logp = F.log_softmax(logits, dim=-1)
            return att_logp, lm_logp, dict(hs=hs, cs=cs, prev=alpha if model.mode == 'loc' else None, lm=lm_hidden)

        return bo.joint_beam_search(step_fn, root, decode_step, beam, x_ctc, ctc_weight, lm_weight, trace=trace)
