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

from . import ce as _ce
from . import ctc as _ctc

GRAD_CLIP = 5            # solver.py:20
VAL_STEP = 30            # solver.py:18
TRAIN_WER_STEP = 250     # solver.py:19


def infer_lengths(x):
    """solver.py:134-135 without the full-tensor D2H round trip: B integers cross to the host
    (pack_padded_sequence wants host lengths), the frame sums stay on the device."""
    return [int(v) for v in (x.sum(dim=-1) != 0).sum(dim=-1).tolist()]


def hybrid_loss(ctc_pred, att_pred, y, enc_len, ctc_weight, ans_len=None, aux=None):
    """(1-w)*att_loss + w*ctc_loss exactly as solver.py:144-163 forms them.

    ctc_pred [B,T',V] raw logits (Seq2Seq.ctc_layer), att_pred [B,L,V] logits, y [B,>=L+1] int64 with
    y[:,0]=<sos>=0, <eos>=1 appended, right padded with 0; enc_len list[int].
    Returns (loss, att_loss, ctc_loss); a dict passed as `aux` receives 'att_argmax' [B,L], the row argmax of
    att_pred that the CE kernel finds on its way (what cal_acc / cal_cer take of the same logits)."""
    tgt_len = torch.sum(y != 0, dim=-1)
    if ans_len is None:
        ans_len = int(torch.max(tgt_len))
    label = y[:, 1:ans_len + 1].contiguous()
    att_loss = ctc_loss = 0
    if ctc_weight < 1:
        # solver.py:149-155 -- CE(ignore_index=0,'none'), per-utterance sum / length, batch mean; one launch
        att_loss, _, argmax = _ce.seq_loss_mean(att_pred, label, tgt_len, ignore_index=0,
                                                grad_prescale=1 - ctc_weight, return_aux=True)
        if aux is not None:
            aux['att_argmax'] = argmax
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
        self.offsets, n = [], 0
        for p in self.params:                       # every tensor starts on a 16-byte boundary (float4 passes)
            self.offsets.append(n)
            n += (p.numel() + 3) // 4 * 4
        dev = self.params[0].device
        self.flat = torch.zeros(n, dtype=torch.float32, device=dev)
        self.timeline = None                    # list -> allreduce_sum appends (start, end) CUDA events
        self.rebind()

    def rebind(self):
        """(Re)attach every parameter's .grad as a view into the flat buffer."""
        for p, o in zip(self.params, self.offsets):
            p.grad = self.flat[o:o + p.numel()].view_as(p)

    def zero(self):
        self.flat.zero_()

    def allreduce_sum(self, group=None):
        """-> world size; the flat buffer then holds the SUM over ranks (the 1/world goes into the clip coefficient)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            rec = self.timeline
            if rec is not None:                 # bench.py's data-parallel breakdown: events around the collective
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            if rec is not None:
                e1.record()
                rec.append((e0, e1))
            return dist.get_world_size(group)
        return 1

    def allreduce_mean(self, group=None):
        world = self.allreduce_sum(group)
        if world > 1:
            self.flat.mul_(1.0 / world)

    def clip_(self, max_norm):
        """clip_grad_norm_ (solver.py:178) on the flat buffer: same 2-norm, one reduction."""
        norm = torch.linalg.vector_norm(self.flat)
        coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
        self.flat.mul_(coef)
        return norm


class DevicePrefetcher:
    """Iterates host batches (tuples of tensors) as device tensors; the host->device copies of batch i+1 are issued on a
    side stream before batch i is handed out, so they run under step i (asynchronous when the host tensors are pinned).
    dtypes: optional per-tensor target dtypes (converted on the device, like `.to(device, dtype)`)."""

    def __init__(self, batches, device, dtypes=None):
        self.batches, self.device, self.dtypes = batches, torch.device(device), dtypes

    def _load(self, batch, stream):
        dtypes = self.dtypes or [None] * len(batch)
        if stream is None:
            return tuple(t.to(self.device, dtype=d) if d is not None else t.to(self.device) for t, d in zip(batch, dtypes)), None
        with torch.cuda.stream(stream):
            out = tuple(t.to(self.device, dtype=d, non_blocking=True) if d is not None else t.to(self.device, non_blocking=True)
                        for t, d in zip(batch, dtypes))
            ev = torch.cuda.Event()
            ev.record(stream)
        return out, ev

    def __iter__(self):
        stream = torch.cuda.Stream(self.device) if self.device.type == 'cuda' else None
        it = iter(self.batches)
        try:
            nxt = self._load(next(it), stream)
        except StopIteration:
            return
        while nxt is not None:
            cur, ev = nxt
            try:
                nxt = self._load(next(it), stream)          # batch i+1 starts moving before batch i is used
            except StopIteration:
                nxt = None
            if ev is not None:
                main = torch.cuda.current_stream(self.device)
                main.wait_event(ev)
                for t in cur:
                    t.record_stream(main)
            yield cur


class _Stepped:
    """Truth value of "the optimizer stepped" (reference solver.py:179: `math.isnan(grad_norm)` -> no step), read from
    the device only when somebody asks."""

    def __init__(self, owner, before):
        self._owner, self._before, self._value = owner, before, None

    def __bool__(self):
        if self._value is None:
            self._value = self._owner.check() == self._before
        return self._value


class TrainStep:
    """One hybrid CTC/attention training step (solver.py:132-182) bound to a model and an optimizer.

    step(x, y, tf_rate) -> dict(loss, att_loss, ctc_loss, grad_norm) of 0-dim DEVICE tensors plus 'stepped'.
    With the one-launch Adadelta the clip and the NaN guard of solver.py:178-182 run inside the optimizer kernel
    (`b2asr_adadelta_step_clipped`), so the step reads nothing back: 'stepped' is an object whose truth value fetches
    the device-side skip counter on demand, and `check()` -- called by Trainer.exec where it reads the loss back
    anyway -- also raises if a Listener recurrence launch timed out.  Other optimizers keep the host-side guard."""

    def __init__(self, model, optimizer, ctc_weight, grad_clip=GRAD_CLIP, group=None):
        self.model, self.opt, self.w, self.clip, self.group = model, optimizer, ctc_weight, grad_clip, group
        self.grads = FlatGradients(model.parameters())
        if hasattr(optimizer, 'bind') and self.grads.flat.is_cuda:
            optimizer.bind(self.grads)                              # one-launch optimizer step (_optim.py)
        self.skipped = torch.zeros(1, dtype=torch.int32, device=self.grads.flat.device)
        self._skipped_seen = 0

    def check(self):
        """One small device read: -> number of NaN-skipped steps so far; raises if an encoder kernel timed out."""
        from . import _enc
        dev = self.grads.flat.device
        word = torch.cat([self.skipped, _enc.status_word(dev)[:1]]).tolist()
        if word[1] != 0:
            _enc.status_word(dev).zero_()
            raise RuntimeError("b2asr_lstm_seq: a persistent launch timed out waiting for a peer CTA; the step's "
                               "outputs are undefined")
        self._skipped_seen = int(word[0])
        return self._skipped_seen

    def step(self, x, y, tf_rate=1.0):
        from . import _enc
        if not x.is_cuda:
            raise RuntimeError('train step runs on the sm_100a kernels only (no CPU path)')
        state_len = infer_lengths(x)
        tgt_len = torch.sum(y != 0, dim=-1)
        ans_len = int(torch.max(tgt_len))
        self.grads.zero()                                           # asr_opt.zero_grad(), solver.py:139
        ctc_pred, enc_len, att_pred, _ = self.model(x, ans_len, tf_rate=tf_rate, teacher=y, state_len=state_len)
        aux = {}
        loss, att_loss, ctc_loss = hybrid_loss(ctc_pred, att_pred, y, enc_len, self.w, ans_len, aux)
        with _enc.deferred_weight_grads():
            loss.backward()
        if getattr(self.opt, 'can_clip_on_device', None) and self.opt.can_clip_on_device():
            world = self.grads.allreduce_sum(self.group)
            raw_norm = torch.linalg.vector_norm(self.grads.flat).reshape(1)
            self.opt.step_clipped(raw_norm, 1.0 / world, float(self.clip), self.skipped)
            grad_norm = raw_norm[0] / world if world > 1 else raw_norm[0]
            stepped = _Stepped(self, self._skipped_seen)
        else:
            self.grads.allreduce_mean(self.group)
            grad_norm = self.grads.clip_(self.clip)
            stepped = not math.isnan(float(grad_norm))              # solver.py:179
            if stepped:
                self.opt.step()
        return dict(loss=loss.detach(), att_loss=att_loss if not torch.is_tensor(att_loss) else att_loss.detach(),
                    ctc_loss=ctc_loss if not torch.is_tensor(ctc_loss) else ctc_loss.detach(),
                    grad_norm=grad_norm, stepped=stepped, att_pred=att_pred, ctc_pred=ctc_pred, enc_len=enc_len,
                    att_argmax=aux.get('att_argmax'))


def make_optimizer(model, opt_cfg):
    """solver.py:101-106 (apex absent in this image -> the torch.optim branch)."""
    if (opt_cfg['type'] == 'Adadelta' and next(model.parameters()).is_cuda
            and os.environ.get('B2ASR_OPT', 'fused') != 'torch'):
        from . import _optim
        return _optim.Adadelta(model.parameters(), lr=opt_cfg['learning_rate'], eps=1e-8)
    cls = getattr(torch.optim, opt_cfg['type'])
    return cls(model.parameters(), lr=opt_cfg['learning_rate'], eps=1e-8)


# ====================================================================================================
# Solver / Trainer / Tester / RNNLM_Trainer: the classes main.py instantiates (solver.py:23-537)
# ====================================================================================================
import itertools
import os

import numpy as np

from .postprocess import Mapper, cal_acc, cal_cer, draw_att

CLM_MIN_SEQ_LEN = 5      # solver.py:21 (CLM itself is out of scope)


def _summary_writer(logdir):
    """tensorboardX is absent from this image; torch.utils.tensorboard has the same add_* calls.  Falls back to a
    no-op writer when tensorboard itself is missing."""
    try:
        from torch.utils.tensorboard import SummaryWriter
        return SummaryWriter(logdir)
    except Exception:
        class _Null:
            def __getattr__(self, name):
                return lambda *a, **k: None
        return _Null()


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class Solver():
    ''' Super class Solver for all kinds of tasks (solver.py:23-49)'''

    def __init__(self, config, paras):
        self.config = config
        self.paras = paras
        if not (self.paras.gpu and torch.cuda.is_available()):
            raise RuntimeError('this build runs the hybrid CTC/attention path on sm_100a kernels only; a CUDA '
                               'device is required (--cpu is not supported: there is no CPU fallback)')
        self.device = torch.device('cuda', torch.cuda.current_device())
        self.exp_name = paras.name
        if self.exp_name is None:
            self.exp_name = '_'.join([paras.config.split('/')[-1].replace('.yaml', ''), 'sd' + str(paras.seed)])
        os.makedirs(paras.ckpdir, exist_ok=True)
        self.ckpdir = os.path.join(paras.ckpdir, self.exp_name)
        os.makedirs(self.ckpdir, exist_ok=True)
        self.mapper = Mapper(config['solver']['data_path'])

    def verbose(self, msg):
        if self.paras.verbose and _dist()[0] == 0:
            print('[INFO]', msg)

    def progress(self, msg):
        if self.paras.verbose and _dist()[0] == 0:
            print(msg + '                              ', end='\r')


class Trainer(Solver):
    ''' Handler for complete training progress (solver.py:52-291)'''

    def __init__(self, config, paras):
        super(Trainer, self).__init__(config, paras)
        self.logdir = os.path.join(paras.logdir, self.exp_name)
        self.log = _summary_writer(self.logdir)
        self.valid_step = config['solver']['dev_step']
        self.best_val_ed = 2.0
        self.step = 0
        self.max_step = config['solver']['total_steps']
        self.tf_start = config['solver']['tf_start']
        self.tf_end = config['solver']['tf_end']
        self.apex = config['solver']['apex']
        self.apply_clm = config['clm']['enable']
        if self.apply_clm:
            raise NotImplementedError('CLM adversarial training (src/clm.py) is outside the accelerated path; '
                                      'set clm.enable: False as every shipped config does')

    def load_data(self):
        from .dataset import LoadDataset
        self.verbose('Loading data from ' + self.config['solver']['data_path'])
        self.train_set = LoadDataset('train', text_only=False, use_gpu=self.paras.gpu, **self.config['solver'])
        self.dev_set = LoadDataset('dev', text_only=False, use_gpu=self.paras.gpu, **self.config['solver'])
        for self.sample_x, _ in self.train_set:
            break
        if len(self.sample_x.shape) == 4:
            self.sample_x = self.sample_x[0]

    def set_model(self):
        from .asr import Seq2Seq
        self.verbose('Init ASR model. Note: validation is done through greedy decoding w/ attention decoder.')
        opt_cfg = self.config['asr_model']['optimizer']
        self.asr_model = Seq2Seq(self.sample_x, self.mapper.get_dim(), self.config['asr_model']).to(self.device)
        if 'VGG' in self.config['asr_model']['encoder']['enc_type']:
            self.verbose('VCC Extractor in Encoder is enabled, time subsample rate = 4.')
        self.seq_loss = _ce.CrossEntropyLoss(ignore_index=0, reduction='none').to(self.device)
        self.ctc_loss = _ctc.CTCLoss(blank=0, reduction='mean')                    # solver.py:93, on the sm_100a kernel
        self.ctc_weight = opt_cfg['joint_ctc']
        if self.paras.load:
            raise NotImplementedError                                              # solver.py:97-98
        self.asr_opt = make_optimizer(self.asr_model, opt_cfg)                     # solver.py:101-106
        self.train_step = TrainStep(self.asr_model, self.asr_opt, self.ctc_weight)

    def write_log(self, val_name, val_dict):
        '''TensorBoard routing by name (solver.py:201-208)'''
        if _dist()[0] != 0:
            return
        if 'att' in val_name:
            self.log.add_image(val_name, val_dict, self.step)
        elif 'txt' in val_name or 'hyp' in val_name:
            self.log.add_text(val_name, val_dict, self.step)
        else:
            self.log.add_scalars(val_name, val_dict, self.step)

    def exec(self):
        ''' Training End-to-end ASR system (solver.py:118-198)'''
        self.verbose('Training set total ' + str(len(self.train_set)) + ' batches.')
        # data parallel (new): every rank walks the SAME shuffled order (main.py seeds all ranks
        # alike) and takes batch i when i mod world == rank; a trailing partial group is dropped so that all ranks
        # enter the same number of gradient all-reduces
        rank, world = _dist()
        usable = (len(self.train_set) // world) * world
        # everything built so far (model, data set, configuration) lives for the whole run: moved out of the garbage
        # collector's generations, its periodic full passes stay short -- with the loss read back every step a long
        # host pause is GPU idle time
        import gc
        gc.collect()
        gc.freeze()
        while self.step < self.max_step:
            def mine():
                for i, (x, y) in enumerate(self.train_set):
                    if i >= usable:
                        break
                    if i % world != rank:
                        continue
                    assert len(x.shape) == 4, 'Bucketing should cause acoustic feature to have shape 1xBxTxD'
                    assert len(y.shape) == 3, 'Bucketing should cause label have to shape 1xBxT'
                    yield x.squeeze(0), y.squeeze(0)

            # the next batch's host->device copies run under the current step (reference solver.py:139-140 moves each
            # batch synchronously at the top of its step)
            for x, y in DevicePrefetcher(mine(), self.device, dtypes=(torch.float32, torch.long)):
                self.progress('Training step - ' + str(self.step))
                tf_rate = self.tf_start - self.step * (self.tf_start - self.tf_end) / self.max_step
                out = self.train_step.step(x, y, tf_rate)
                if not out['stepped']:
                    self.verbose('Error : grad norm is NaN @ step ' + str(self.step))
                loss_log = {'train_full': out['loss']}
                if self.ctc_weight < 1:
                    loss_log['train_att'] = out['att_loss']
                if self.ctc_weight > 0:
                    loss_log['train_ctc'] = out['ctc_loss']
                self.write_log('loss', loss_log)
                label = y[:, 1:out['att_pred'].shape[1] + 1] if out['att_pred'] is not None else None
                if self.ctc_weight < 1:
                    self.write_log('acc', {'train': cal_acc(out['att_pred'], label, hyp=out['att_argmax'])})
                if self.step % TRAIN_WER_STEP == 0 and self.ctc_weight < 1:
                    self.write_log('error rate', {'train': cal_cer(out['att_pred'], label, mapper=self.mapper)})
                if self.step % self.valid_step == 0:
                    self.train_step.grads.zero()
                    self.valid()
                self.step += 1
                if self.step > self.max_step:
                    break

    def valid(self):
        '''Greedy attention decoding on the dev set, losses, best-model checkpoint (solver.py:211-291)'''
        self.asr_model.eval()
        val_ctc = val_att = val_acc = val_cer = 0.0
        val_len = 0
        all_pred, all_true = [], []
        att_pred = att_maps = label = None
        with torch.no_grad():
            for cur_b, (x, y) in enumerate(self.dev_set):
                self.progress(' '.join(['Valid step -', str(self.step), '(', str(cur_b), '/', str(len(self.dev_set)), ')']))
                if len(x.shape) == 4:
                    x = x.squeeze(0)
                if len(y.shape) == 3:
                    y = y.squeeze(0)
                x = x.to(device=self.device, dtype=torch.float32)
                y = y.to(device=self.device, dtype=torch.long)
                state_len = infer_lengths(x)
                ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
                ctc_pred, state_len, att_pred, att_maps = self.asr_model(x, ans_len + VAL_STEP, state_len=state_len)
                label = y[:, 1:ans_len + 1].contiguous()
                n = int(x.shape[0])
                if self.ctc_weight < 1:
                    sl = self.seq_loss(att_pred[:, :ans_len, :].contiguous().view(-1, att_pred.shape[-1]), label.view(-1))
                    sl = torch.sum(sl.view(n, -1), dim=-1) / torch.sum(y != 0, dim=-1).to(torch.float32)
                    val_att += torch.mean(sl).detach() * n
                    t1, t2 = cal_cer(att_pred, label, mapper=self.mapper, get_sentence=True)
                    all_pred += t1
                    all_true += t2
                    val_acc += cal_acc(att_pred, label) * n
                    val_cer += cal_cer(att_pred, label, mapper=self.mapper) * n
                if self.ctc_weight > 0:
                    target_len = torch.sum(y != 0, dim=-1)
                    cl = self.ctc_loss(F.log_softmax(ctc_pred.transpose(0, 1), dim=-1), label,
                                       torch.LongTensor(state_len), target_len)
                    val_ctc += cl.detach() * n
                val_len += n
        val_loss = (1 - self.ctc_weight) * val_att + self.ctc_weight * val_ctc
        loss_log = {}
        for k, v in zip(['dev_full', 'dev_ctc', 'dev_att'], [val_loss, val_ctc, val_att]):
            if v > 0.0:
                loss_log[k] = v / val_len
        self.write_log('loss', loss_log)
        if self.ctc_weight < 1:
            val_hyp, val_txt = cal_cer(att_pred, label, mapper=self.mapper, get_sentence=True)
            val_attmap = draw_att(att_maps, att_pred)
            self.write_log('error rate', {'dev': val_cer / val_len})
            self.write_log('acc', {'dev': val_acc / val_len})
            for idx, attmap in enumerate(val_attmap):
                self.write_log('att_' + str(idx), attmap)
                self.write_log('hyp_' + str(idx), val_hyp[idx])
                self.write_log('txt_' + str(idx), val_txt[idx])
            if val_cer / val_len < self.best_val_ed and _dist()[0] == 0:
                self.best_val_ed = val_cer / val_len
                self.verbose('Best val er       : {:.4f}       @ step {}'.format(self.best_val_ed, self.step))
                self.asr_model.clear_att()
                for p in self.asr_model.parameters():                              # the flat-gradient views do not pickle
                    p.grad = None
                torch.save(self.asr_model, os.path.join(self.ckpdir, 'asr'))      # solver.py:283, whole module
                self.train_step.grads.rebind()
                with open(os.path.join(self.ckpdir, 'best_hyp.txt'), 'w') as f:
                    for t1, t2 in zip(all_pred, all_true):
                        f.write(t1 + ',' + t2 + '\n')
        self.asr_model.train()


class Tester(Solver):
    ''' Handler for complete inference progress (solver.py:294-441).

    The reference clones the model per utterance and fans utterances out to joblib worker processes
    (solver.py:349,380).  Here the beam search is device-resident (beam.py), so one process per GPU decodes its
    share of the utterances in place: utterance i goes to rank i mod world (no collective); every rank keeps its
    results in memory and rank 0 writes the files in utterance order, so the output is byte-identical whatever
    the number of GPUs (the reference's file order is nondeterministic for njobs > 1).'''

    def __init__(self, config, paras):
        super(Tester, self).__init__(config, paras)
        self.verbose('During beam decoding, batch size is set to 1; utterances are sharded across GPUs, --njobs utterances in flight per GPU.')
        self.njobs = self.paras.njobs
        self.decode_step_ratio = config['solver']['max_decode_step_ratio']
        self.decode_file = "_".join(['decode', 'beam', str(self.config['solver']['decode_beam_size']),
                                     'len', str(self.config['solver']['max_decode_step_ratio'])])

    def load_data(self):
        from .dataset import LoadDataset
        self.verbose('Loading testing data ' + str(self.config['solver']['test_set'])
                     + ' from ' + self.config['solver']['data_path'])
        self.test_set = LoadDataset('test', text_only=False, use_gpu=self.paras.gpu, **self.config['solver'])
        self.dev_set = LoadDataset('dev', text_only=False, use_gpu=self.paras.gpu, **self.config['solver'])

    def set_model(self):
        ''' Load saved ASR (solver.py:311-340)'''
        cfg = self.config['solver']
        self.verbose('Load ASR model from ' + os.path.join(self.ckpdir))
        self.asr_model = torch.load(os.path.join(self.ckpdir, 'asr'), weights_only=False, map_location=self.device)
        trained_with_ctc = bool(getattr(self.asr_model, 'joint_ctc', False))
        self.asr_model.joint_ctc = cfg['decode_ctc_weight'] > 0
        if cfg['decode_ctc_weight'] > 0:
            assert trained_with_ctc, "The ASR was not trained with CTC"
            self.verbose('Joint CTC decoding is enabled with weight = ' + str(cfg['decode_ctc_weight']))
            self.decode_file += '_ctc{:}'.format(cfg['decode_ctc_weight'])
            self.asr_model.ctc_weight = cfg['decode_ctc_weight']
        self.decode_lm = cfg['decode_lm_weight'] > 0
        setattr(self.asr_model, 'decode_lm_weight', cfg['decode_lm_weight'])
        if self.decode_lm:
            assert os.path.exists(cfg['decode_lm_path']), 'Please specify RNNLM path.'
            self.asr_model.load_lm(**cfg)
            self.asr_model.rnn_lm = self.asr_model.rnn_lm.to(self.device)
            self.verbose('Joint RNNLM decoding is enabled with weight = ' + str(cfg['decode_lm_weight']))
            self.verbose('Loading RNNLM from ' + cfg['decode_lm_path'])
            self.decode_file += '_lm{:}'.format(cfg['decode_lm_weight'])
        self.asr_model.eval()
        self.asr_model.clear_att()
        self.asr_model = self.asr_model.to(self.device)
        self.verbose('Checking models performance on dev set ' + str(cfg['dev_set']) + '...')
        self.valid()

    def exec(self):
        '''Beam-search decode of the test set (solver.py:342-354)'''
        rank, world = _dist()
        self.decode_beam_size = self.config['solver']['decode_beam_size']
        self.verbose('Start decoding with beam search, beam size = ' + str(self.decode_beam_size))
        self.verbose('Number of utts to decode : {}, sharded over {} GPU(s).'.format(len(self.test_set), world))
        mine = []
        journal = None
        if world > 1:                       # rank-local journal, appended per utterance: an abort loses nothing decoded
            import json
            journal = open(os.path.join(self.ckpdir, self.decode_file + '.rank%d.jsonl' % rank), 'w')
        def one(i, x, ref):
            try:
                return self.beam_decode(x, ref, write=False)
            except ValueError as e:         # a shape the kernels refuse (e.g. an over-long utterance): report, go on
                self.verbose('Utterance {} skipped: {}'.format(i, e))
                return (self.mapper.translate(ref, return_string=True), [''])

        def emit(i, item):
            if world == 1:
                self._write(*item)          # appended utterance by utterance, like the reference (solver.py:360-367)
            else:
                mine.append((i, item))
                journal.write(json.dumps([i, item[0], item[1]]) + '\n')
                journal.flush()

        # --njobs N (the reference's worker processes, README.md:123) = up to N utterances in flight on this GPU,
        # each on a CUDA stream of its own (beam.beam_decode_many); results are written in corpus order either way
        group_size = max(1, min(int(getattr(self, 'njobs', 1) or 1), 8))
        group = []

        def flush():
            if not group:
                return
            items = None
            if len(group) > 1:
                try:
                    from . import beam as _beam
                    utts = []
                    for _, x, _ in group:
                        xd = x.to(device=self.device, dtype=torch.float32)
                        lens = infer_lengths(xd)
                        utts.append((xd, int(np.ceil(lens[0] * self.decode_step_ratio)), lens))
                    with torch.no_grad():
                        outs = _beam.beam_decode_many(self.asr_model, utts, self.decode_beam_size, streams=len(group))
                    items = [(self.mapper.translate(ref, return_string=True),
                              [self.mapper.translate(h.outIndex, return_string=True) for h in hyps])
                             for (_, _, ref), hyps in zip(group, outs)]
                except ValueError:
                    items = None            # one of them is refused: decode the group one by one to find and skip it
            if items is None:
                items = [one(i, x, ref) for i, x, ref in group]
            for (i, _, _), item in zip(group, items):
                emit(i, item)
            del group[:]

        for i, (x, y) in enumerate(self.test_set):
            if i % world != rank:
                continue
            group.append((i, x[0], y[0].tolist()[0]))
            if len(group) >= group_size:
                flush()
        flush()
        if world > 1:
            import torch.distributed as dist
            journal.close()
            gathered = [None] * world
            dist.all_gather_object(gathered, mine)                 # host-side gather of strings, not a GPU collective
            if rank == 0:
                for _, (gt, hyps) in sorted((item for part in gathered for item in part), key=lambda t: t[0]):
                    self._write(gt, hyps)
            dist.barrier()
            os.remove(os.path.join(self.ckpdir, self.decode_file + '.rank%d.jsonl' % rank))
        self.verbose('Decode done, best results at {}.'.format(str(os.path.join(self.ckpdir, self.decode_file + '.txt'))))
        self.verbose('Top {} results at {}.'.format(self.decode_beam_size,
                                                    str(os.path.join(self.ckpdir, self.decode_file + '_nbest.txt'))))

    def _write(self, gt, hyp_texts):
        with open(os.path.join(self.ckpdir, self.decode_file + '.txt'), 'a') as f:
            f.write(gt + '\t' + hyp_texts[0] + '\n')
        with open(os.path.join(self.ckpdir, self.decode_file + '_nbest.txt'), 'a') as f:
            for h in hyp_texts:
                f.write(gt + '\t' + h + '\n')

    def write_hyp(self, hyps, y):
        '''Record decoding results: "gt<TAB>hyp" lines, best and n-best files (solver.py:356-367)'''
        gt = self.mapper.translate(y, return_string=True)
        self._write(gt, [self.mapper.translate(h.outIndex, return_string=True) for h in hyps])

    def beam_decode(self, x, y, write=True):
        '''One utterance (solver.py:370-387); x [1,T,D] features, y list[int] reference labels'''
        x = x.to(device=self.device, dtype=torch.float32)
        state_len = infer_lengths(x)
        with torch.no_grad():
            max_decode_step = int(np.ceil(state_len[0] * self.decode_step_ratio))
            hyps = self.asr_model.beam_decode(x, max_decode_step, state_len, self.decode_beam_size)
        if write:
            self.write_hyp(hyps, y)
            return 1
        return (self.mapper.translate(y, return_string=True),
                [self.mapper.translate(h.outIndex, return_string=True) for h in hyps])

    def valid(self):
        '''Greedy attention (and CTC) decoding of the dev set into dev_att_decode.txt / dev_ctc_decode.txt
        (solver.py:390-441)'''
        val_cer, val_len = 0.0, 0
        all_pred, all_true, ctc_results = [], [], []
        with torch.no_grad():
            for cur_b, (x, y) in enumerate(self.dev_set):
                self.progress(' '.join(['Valid step - (', str(cur_b), '/', str(len(self.dev_set)), ')']))
                if len(x.shape) == 4:
                    x = x.squeeze(0)
                if len(y.shape) == 3:
                    y = y.squeeze(0)
                x = x.to(device=self.device, dtype=torch.float32)
                y = y.to(device=self.device, dtype=torch.long)
                state_len = infer_lengths(x)
                ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
                ctc_pred, state_len, att_pred, att_maps = self.asr_model(x, ans_len + VAL_STEP, state_len=state_len)
                ctc_results.append(torch.argmax(ctc_pred, dim=-1).cpu() if ctc_pred is not None else None)
                label = y[:, 1:ans_len + 1].contiguous()
                t1, t2 = cal_cer(att_pred, label, mapper=self.mapper, get_sentence=True)
                all_pred += t1
                all_true += t2
                val_cer += cal_cer(att_pred, label, mapper=self.mapper) * int(x.shape[0])
                val_len += int(x.shape[0])
        self.verbose('Validation Error Rate of Current model : {:.4f}      '.format(val_cer / val_len))
        self.verbose('See {} for validation results.'.format(os.path.join(self.ckpdir, 'dev_att_decode.txt')))
        if _dist()[0] != 0:
            return
        with open(os.path.join(self.ckpdir, 'dev_att_decode.txt'), 'w') as f:
            for hyp, gt in zip(all_pred, all_true):
                f.write(gt.lstrip() + '\t' + hyp.lstrip() + '\n')
        if ctc_results[0] is not None:
            ctc_text = []
            for pred in [i for ins in ctc_results for i in ins]:
                p = [i for i in pred.tolist() if i != 0]
                p = [k for k, g in itertools.groupby(p)]
                ctc_text.append(self.mapper.translate(p, return_string=True))
            self.verbose('Also, see {} for CTC validation results.'.format(os.path.join(self.ckpdir, 'dev_ctc_decode.txt')))
            with open(os.path.join(self.ckpdir, 'dev_ctc_decode.txt'), 'w') as f:
                for hyp, gt in zip(ctc_text, all_true):
                    f.write(gt.lstrip() + '\t' + hyp.lstrip() + '\n')


class RNNLM_Trainer(Solver):
    ''' Trainer for RNN-LM only (solver.py:444-537).  Stock cuDNN/cuBLAS training loop, kept so that
    `main.py --rnnlm` keeps working; LM training is not part of the accelerated path.'''

    def __init__(self, config, paras):
        super(RNNLM_Trainer, self).__init__(config, paras)
        self.logdir = os.path.join(paras.logdir, self.exp_name)
        self.log = _summary_writer(self.logdir)
        self.valid_step = config['solver']['dev_step']
        self.best_dev_ppx = 1000
        self.step = 0
        self.max_step = config['solver']['total_steps']
        self.apex = config['solver']['apex']

    def load_data(self):
        from .dataset import LoadDataset
        self.verbose('Loading text data from ' + self.config['solver']['data_path'])
        self.train_set = LoadDataset('train', text_only=True, use_gpu=self.paras.gpu, **self.config['solver'])
        self.dev_set = LoadDataset('dev', text_only=True, use_gpu=self.paras.gpu, **self.config['solver'])

    def set_model(self):
        from .rnnlm import RNN_LM
        self.verbose('Init RNNLM model.')
        self.rnnlm = RNN_LM(out_dim=self.mapper.get_dim(), **self.config['rnn_lm']['model_para']).to(self.device)
        if self.paras.load:
            raise NotImplementedError
        self.rnnlm_opt = make_optimizer(self.rnnlm, self.config['rnn_lm']['optimizer'])

    def _loss(self, y):
        if len(y.shape) == 3:
            y = y.squeeze(0)
        y = y.to(device=self.device, dtype=torch.long)
        ans_len = torch.sum(y != 0, dim=-1).cpu()
        _, prob = self.rnnlm(y[:, :-1], ans_len.clamp(max=y.shape[1] - 1), None)
        return F.cross_entropy(prob.reshape(-1, prob.shape[-1]), y[:, 1:prob.shape[1] + 1].contiguous().view(-1),
                               ignore_index=0), y.shape[0]

    def exec(self):
        self.verbose('RNN-LM Training set total ' + str(len(self.train_set)) + ' batches.')
        while self.step < self.max_step:
            for y in self.train_set:
                self.progress('Training step - ' + str(self.step))
                self.rnnlm_opt.zero_grad()
                loss, _ = self._loss(y)
                loss.backward()
                self.rnnlm_opt.step()
                self.log.add_scalars('perplexity', {'train': torch.exp(loss.detach()).item()}, self.step)
                self.step += 1
                if self.step % self.valid_step == 0:
                    self.valid()
                if self.step > self.max_step:
                    break

    def valid(self):
        self.rnnlm.eval()
        total, n = 0.0, 0
        with torch.no_grad():
            for cur_b, y in enumerate(self.dev_set):
                self.progress(' '.join(['Valid step -', str(self.step), '(', str(cur_b), '/', str(len(self.dev_set)), ')']))
                loss, b = self._loss(y)
                total += float(loss) * b
                n += b
        dev_ppx = math.exp(total / max(n, 1))
        self.log.add_scalars('perplexity', {'dev': dev_ppx}, self.step)
        if dev_ppx < self.best_dev_ppx:
            self.best_dev_ppx = dev_ppx
            self.verbose('Best val ppx      : {:.4f}       @ step {}'.format(self.best_dev_ppx, self.step))
            torch.save(self.rnnlm, os.path.join(self.ckpdir, 'rnnlm'))
        self.rnnlm.train()
