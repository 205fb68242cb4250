"""Host-side mirror of the reference's src/rnnlm.py:9-25 (RNN language model used for shallow fusion in
the beam search, asr.py:232-235).  Same constructor, parameter names (emb / rnn / out) and
`forward(x, lens, hidden) -> (hidden, logits)`; the arithmetic stays cuDNN/cuBLAS.
`step()` is the beam-batched one-token path: all live hypotheses in ONE LSTM call instead of the
reference's batch-1 call per hypothesis, so the LM weights are read once per decode step."""
import torch
import torch.nn as nn

from . import fp32_library_math


class RNN_LM(nn.Module):
    def __init__(self, emb_dim, h_dim, out_dim, layers=1, rnn='LSTM', dropout_rate=0.0):
        super().__init__()
        fp32_library_math()
        self.h_dim = h_dim
        self.emb = nn.Embedding(out_dim, emb_dim)
        self.drop1 = nn.Dropout(dropout_rate)
        self.drop2 = nn.Dropout(dropout_rate)
        self.rnn = getattr(nn, rnn.upper())(emb_dim, h_dim, num_layers=layers, dropout=dropout_rate, batch_first=True)
        self.out = nn.Linear(h_dim, out_dim)

    def forward(self, x, lens, hidden=None):
        embedded = self.drop2(self.emb(x))
        packed = nn.utils.rnn.pack_padded_sequence(embedded, lens, batch_first=True)
        outputs, hidden = self.rnn(packed, hidden)
        outputs, _ = nn.utils.rnn.pad_packed_sequence(outputs, batch_first=True)
        return hidden, self.out(self.drop2(outputs))

    def step(self, tokens, hidden):
        """tokens [n] int64 (one per hypothesis), hidden (h, c) each [layers, n, h_dim] -> (hidden, logits [n, V])."""
        out, hidden = self.rnn(self.emb(tokens).unsqueeze(1), hidden)
        return hidden, self.out(out[:, 0, :])
