"""Host-side mirror of the reference's src/solver.py for the hybrid CTC/attention path.

  train_step / TrainStep   body of Trainer.exec for one batch          solver.py:132-182
  hybrid_loss              CE (attention branch) + CTC, weighted       solver.py:144-163
  infer_lengths            valid frames = frames with non-zero sum     solver.py:134-135
  allreduce_gradients      (new) NCCL average of the flat gradient before clip / NaN guard, so every
                           rank takes the same branch at solver.py:178-182

The CTC term runs on the sm_100a kernel with the log-softmax of solver.py:160 fused in (logits in,
gradient w.r.t. logits out, no transpose materialised); the decoder loop runs one fused attention
kernel per step (asr.py).  There is no CPU path: inputs must live on a CUDA device.
"""
import math

import torch
import torch.nn.functional as F

from . import ctc as _ctc

GRAD_CLIP = 5            # solver.py:20
VAL_STEP = 30            # solver.py:18
TRAIN_WER_STEP = 250     # solver.py:19


def infer_lengths(x):
    """solver.py:134-135 without the full-tensor D2H round trip: B integers cross to the host
    (pack_padded_sequence wants host lengths), the frame sums stay on the device."""
    return [int(v) for v in (x.sum(dim=-1) != 0).sum(dim=-1).tolist()]


def hybrid_loss(ctc_pred, att_pred, y, enc_len, ctc_weight, ans_len=None):
    """(1-w)*att_loss + w*ctc_loss exactly as solver.py:144-163 forms them.

    ctc_pred [B,T',V] raw logits (Seq2Seq.ctc_layer), att_pred [B,L,V] logits, y [B,>=L+1] int64 with
    y[:,0]=<sos>=0, <eos>=1 appended, right padded with 0; enc_len list[int].
    Returns (loss, att_loss, ctc_loss)."""
    tgt_len = torch.sum(y != 0, dim=-1)
    if ans_len is None:
        ans_len = int(torch.max(tgt_len))
    label = y[:, 1:ans_len + 1].contiguous()
    att_loss = ctc_loss = 0
    if ctc_weight < 1:
        b, t, c = att_pred.shape
        ce = F.cross_entropy(att_pred.reshape(b * t, c), label.reshape(-1), ignore_index=0, reduction='none')
        att_loss = torch.mean(ce.view(b, t).sum(dim=-1) / tgt_len.to(torch.float32))
    if ctc_weight > 0:
        # solver.py:160 -- F.log_softmax(ctc_pred.transpose(0,1)) + CTCLoss(blank=0,'mean'); fused here
        ctc_loss = _ctc.ctc_loss(ctc_pred, label, enc_len, tgt_len, blank=0, reduction='mean',
                                 from_logits=True, batch_first=True, grad_prescale=ctc_weight)
    return (1 - ctc_weight) * att_loss + ctc_weight * ctc_loss, att_loss, ctc_loss


class FlatGradients:
    """All parameter gradients as views into ONE flat fp32 buffer, so the data-parallel exchange is a
    single NCCL all-reduce over NVLink (98 MB for the libri-shaped model) and zero_grad is one memset."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)
        o = 0
        for p in self.params:
            p.grad = self.flat[o:o + p.numel()].view_as(p)
            o += p.numel()

    def zero(self):
        self.flat.zero_()

    def allreduce_mean(self, group=None):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            self.flat.mul_(1.0 / dist.get_world_size(group))

    def clip_(self, max_norm):
        """clip_grad_norm_ (solver.py:178) on the flat buffer: same 2-norm, one reduction."""
        norm = torch.linalg.vector_norm(self.flat)
        coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
        self.flat.mul_(coef)
        return norm


class TrainStep:
    """One hybrid CTC/attention training step (solver.py:132-182) bound to a model and an optimizer.

    step(x, y, tf_rate) -> dict(loss, att_loss, ctc_loss, grad_norm) of 0-dim DEVICE tensors, plus
    'stepped' (bool, host): the NaN guard of solver.py:179-182 needs the norm on the host, which is the
    step's one unavoidable D2H read besides the B encoder lengths."""

    def __init__(self, model, optimizer, ctc_weight, grad_clip=GRAD_CLIP, group=None):
        self.model, self.opt, self.w, self.clip, self.group = model, optimizer, ctc_weight, grad_clip, group
        self.grads = FlatGradients(model.parameters())

    def step(self, x, y, tf_rate=1.0):
        if not x.is_cuda:
            raise RuntimeError('train step runs on the sm_100a kernels only (no CPU path)')
        state_len = infer_lengths(x)
        tgt_len = torch.sum(y != 0, dim=-1)
        ans_len = int(torch.max(tgt_len))
        self.grads.zero()                                           # asr_opt.zero_grad(), solver.py:139
        ctc_pred, enc_len, att_pred, _ = self.model(x, ans_len, tf_rate=tf_rate, teacher=y, state_len=state_len)
        loss, att_loss, ctc_loss = hybrid_loss(ctc_pred, att_pred, y, enc_len, self.w, ans_len)
        loss.backward()
        self.grads.allreduce_mean(self.group)
        grad_norm = self.grads.clip_(self.clip)
        stepped = not math.isnan(float(grad_norm))                  # solver.py:179
        if stepped:
            self.opt.step()
        return dict(loss=loss.detach(), att_loss=att_loss if not torch.is_tensor(att_loss) else att_loss.detach(),
                    ctc_loss=ctc_loss if not torch.is_tensor(ctc_loss) else ctc_loss.detach(),
                    grad_norm=grad_norm, stepped=stepped, att_pred=att_pred, ctc_pred=ctc_pred, enc_len=enc_len)


def make_optimizer(model, opt_cfg):
    """solver.py:101-106 (apex absent in this image -> the torch.optim branch)."""
    cls = getattr(torch.optim, opt_cfg['type'])
    return cls(model.parameters(), lr=opt_cfg['learning_rate'], eps=1e-8)
