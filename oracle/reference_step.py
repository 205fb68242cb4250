"""The reference's hybrid CTC/attention train step, driven through the reference's OWN code.

MEASUREMENT INFRASTRUCTURE ONLY (bench.py's `--impl reference` arm, its `cpu_baseline` leg and the
`stock_torch_gpu` leg).  Nothing in the product package imports this.

`Trainer.exec` (reference src/solver.py:122-198) is a loop over a DataLoader with the step's body inline, so the
body cannot be called on its own; `ReferenceStep.step` issues the same statements in the same order --
solver.py:134-139 (length inference, zero_grad), :140 (`Seq2Seq.forward`, the reference's class, unmodified),
:144-163 (`CrossEntropyLoss(ignore_index=0, reduction='none')` normalised per utterance, `torch.nn.CTCLoss` on
`F.log_softmax(ctc_pred.transpose(0,1))` with CPU input lengths), :177-182 (`backward`, `clip_grad_norm_(.., 5)`,
NaN guard, `Adadelta(lr, eps=1e-8).step()`) -- on the model the reference's `Seq2Seq` constructor builds from the
same config dict.  With `kind == "reference"` every module on the path is the reference's (imported by file path
from /root/reference or from the copy under oracle/_ref/); if neither is present the oracle port
(oracle/model_oracle.py, pinned against the reference's goldens) takes its place and `kind == "port"`.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from . import reference_loader

GRAD_CLIP = 5   # solver.py:21


class ReferenceStep:
    def __init__(self, d_in, vocab, cfg, device="cpu", seed=0):
        self.device = torch.device(device)
        self.w = cfg['optimizer']['joint_ctc']
        torch.manual_seed(seed)
        if reference_loader.available():
            ref = reference_loader.load()
            self.kind = "reference"
            self.root = ref.root
            self.model = ref.asr.Seq2Seq(torch.zeros(1, 8, d_in), vocab, cfg).to(self.device)
            self._forward = self.model                                       # Seq2Seq.forward, asr.py:58
        else:
            from . import model_oracle as mo
            self.kind = "port"
            self.root = None
            self.model = mo.OracleModel(d_in, vocab, cfg).to(self.device)
            self._forward = self.model.run
        self.model.train()
        self.seq_loss = torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none').to(self.device)    # solver.py:90
        self.ctc_loss = torch.nn.CTCLoss(blank=0, reduction='mean')                                    # solver.py:93
        opt = cfg['optimizer']
        self.opt = getattr(torch.optim, opt['type'])(self.model.parameters(), lr=opt['learning_rate'], eps=1e-8)

    def step(self, x, y, tf_rate=1.0):
        dev = self.device
        x = x.to(device=dev, dtype=torch.float32)
        y = y.to(device=dev, dtype=torch.long)
        state_len = np.sum(np.sum(x.cpu().data.numpy(), axis=-1) != 0, axis=-1)                        # solver.py:134
        state_len = [int(sl) for sl in state_len]
        ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
        self.opt.zero_grad()
        ctc_pred, state_len, att_pred, _ = self._forward(x, ans_len, tf_rate=tf_rate, teacher=y, state_len=state_len)
        label = y[:, 1:ans_len + 1].contiguous()
        ctc_loss = att_loss = 0
        if self.w < 1:
            b, t, c = att_pred.shape
            att_loss = self.seq_loss(att_pred.view(b * t, c), label.view(-1))
            att_loss = torch.sum(att_loss.view(b, t), dim=-1) / torch.sum(y != 0, dim=-1).to(device=dev, dtype=torch.float32)
            att_loss = torch.mean(att_loss)
        if self.w > 0:
            target_len = torch.sum(y != 0, dim=-1)
            ctc_loss = self.ctc_loss(F.log_softmax(ctc_pred.transpose(0, 1), dim=-1), label,
                                     torch.LongTensor(state_len), target_len)
        loss = (1 - self.w) * att_loss + self.w * ctc_loss
        loss.backward()
        grad_norm = torch.nn.utils.clip_grad_norm_(self.model.parameters(), GRAD_CLIP)
        if not math.isnan(grad_norm):
            self.opt.step()
        return float(loss)
