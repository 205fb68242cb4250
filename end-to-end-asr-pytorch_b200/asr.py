"""Host-side mirror of the reference's src/asr.py, with the hot path on the sm_100a kernels.

Same class names, constructor arguments, parameter names (state_dicts interchange with the
reference) and return conventions:

  Seq2Seq(example_input, output_dim, model_para)               asr.py:18-48
    .forward(audio_feature, decode_step, tf_rate, teacher, state_len)   asr.py:58-112
    .beam_decode(audio_feature, decode_step, state_len, decode_beam_size)  asr.py:155-258
    .load_lm / .clear_att                                       asr.py:50-56
  Listener / RNNLayer / VGGExtractor                            asr.py:266-317, 466-558 (LSTM recurrence:
                                                                csrc/enc.cu persistent kernels; VGG: cuDNN)
  Speller                                                       asr.py:320-357 (nn.LSTMCell)
  Attention                                                     asr.py:370-462 -> kernels (2a)/(2b)

What runs where: the VGG convs and every GEMM stay library calls (cuDNN/cuBLAS), as SURVEY.md section 2
scopes them; the Listener's LSTM recurrences run as one persistent launch per layer and direction pair (_enc.py).  Attention.forward is one fused kernel
launch per decoder step (forward and backward), the key projection is its own kernel, and the
joint CTC/attention/LM beam step runs device-resident (beam.py).  There is no CPU path through
Attention or beam_decode: they raise if the tensors are not on a CUDA device.
"""
import math
import os
import random

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.distributions.categorical import Categorical
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

from . import _attn, _dec, _decloop, _enc, _gemm, _keys, fp32_library_math

CTC_BEAM_RATIO = 1.5   # asr.py:15
# B2ASR_ENC=cudnn keeps the Listener's LSTMs on the library's per-frame path (A/B measurements only)
ENCODER_NATIVE = os.environ.get("B2ASR_ENC", "native") != "cudnn"


# --------------------------------------------------------------------------------------------
# Encoder (library ops; API kept)
# --------------------------------------------------------------------------------------------
class VGGExtractor(nn.Module):
    """4x time-subsampling conv front-end (asr.py:507-558)."""

    def __init__(self, example_input):
        super().__init__()
        d = example_input.shape[-1]
        if d % 13 == 0:
            self.in_channel, self.freq_dim = d // 13, 13
        elif d % 40 == 0:
            self.in_channel, self.freq_dim = d // 40, 40
        else:
            raise ValueError('Acoustic feature dimension for VGG should be 13/26/39(MFCC) or 40/80/120(Fbank) but got %d' % d)
        self.out_dim = (self.freq_dim // 4) * 128
        self.conv1 = nn.Conv2d(self.in_channel, 64, 3, stride=1, padding=1)
        self.conv2 = nn.Conv2d(64, 64, 3, stride=1, padding=1)
        self.pool1 = nn.MaxPool2d(2, stride=2)
        self.conv3 = nn.Conv2d(64, 128, 3, stride=1, padding=1)
        self.conv4 = nn.Conv2d(128, 128, 3, stride=1, padding=1)
        self.pool2 = nn.MaxPool2d(2, stride=2)

    def forward(self, feature, xlen):
        xlen = [n // 4 for n in xlen]
        keep = feature.shape[1] - feature.shape[1] % 4
        feature = feature[:, :keep, :]
        bs, ts, _ = feature.shape
        x = feature.reshape(bs, ts, self.in_channel, self.freq_dim).transpose(1, 2)
        x = self.pool1(F.relu(self.conv2(F.relu(self.conv1(x)))))
        x = self.pool2(F.relu(self.conv4(F.relu(self.conv3(x)))))
        x = x.transpose(1, 2).contiguous()
        return x.view(x.shape[0], x.shape[1], self.out_dim), xlen


class RNNLayer(nn.Module):
    """One (bi)RNN layer + time subsampling by dropping or concatenating frames (asr.py:466-501)."""

    def __init__(self, in_dim, out_dim, sample_rate, sample_style='drop', layers=1, rnn_cell='LSTM',
                 dropout_rate=0.0, bidir=True):
        super().__init__()
        self.sample_style = sample_style
        self.sample_rate = sample_rate
        self.layer = getattr(nn, rnn_cell.upper())(in_dim, out_dim, bidirectional=bidir, num_layers=layers,
                                                   dropout=dropout_rate, batch_first=True)

    def forward(self, input_x, state=None, state_len=None, pack_input=False):
        if pack_input:
            assert state_len is not None, "Please specify seq len for pack_padded_sequence."
        if pack_input and state is None and ENCODER_NATIVE and _enc.supported(self.layer, input_x):
            # persistent recurrence kernels (csrc/enc.cu) instead of the library's per-frame launches; the final
            # (h_n, c_n) pair is not materialised -- Listener.forward (asr.py:313) discards it
            state_len = [int(n) for n in state_len]
            output, hidden = _enc.lstm_layer(self.layer, input_x, state_len), None
        else:
            if pack_input:
                input_x = pack_padded_sequence(input_x, state_len, batch_first=True)
            output, hidden = self.layer(input_x, state)
            if pack_input:
                output, lens = pad_packed_sequence(output, batch_first=True)
                state_len = lens.tolist()
        if self.sample_rate > 1:
            bs, ts, fd = output.shape
            if self.sample_style == 'drop':
                output = output[:, ::self.sample_rate, :]
            elif self.sample_style == 'concat':
                keep = ts - ts % self.sample_rate
                output = output[:, :keep, :].contiguous().view(bs, keep // self.sample_rate, fd * self.sample_rate)
            else:
                raise ValueError('Unsupported Sample Style: ' + self.sample_style)
            if state_len is not None:
                state_len = [int(n / self.sample_rate) for n in state_len]
        if state_len is not None:
            return output, hidden, state_len
        return output, hidden


class Listener(nn.Module):
    """(VGG +) stacked (pyramidal) RNN encoder with a tanh(Linear) after every layer (asr.py:266-317)."""

    def __init__(self, example_input, enc_type, sample_rate, sample_style, dim, dropout, rnn_cell):
        super().__init__()
        input_dim = example_input.shape[-1]
        self.enc_type = enc_type
        self.vgg = False
        self.dims = [int(v) for v in dim.split('_')]
        self.sample_rate = [int(v) for v in sample_rate.split('_')]
        self.dropout = [float(v) for v in dropout.split('_')]
        self.sample_style = sample_style
        assert len(self.sample_rate) == len(self.dropout), 'Number of layer mismatch'
        assert len(self.dropout) == len(self.dims), 'Number of layer mismatch'
        self.num_layers = len(self.sample_rate)
        assert self.num_layers >= 1, 'Listener should have at least 1 layer'
        if 'VGG' in enc_type:
            self.vgg = True
            self.vgg_extractor = VGGExtractor(example_input)
            input_dim = self.vgg_extractor.out_dim
        if 'RNN' not in enc_type:
            raise ValueError('Unsupported Encoder Type: ' + enc_type)
        bidir = 'BiRNN' in enc_type
        for l, (out_dim, sr, drop) in enumerate(zip(self.dims, self.sample_rate, self.dropout)):
            setattr(self, 'layer' + str(l), RNNLayer(input_dim, out_dim, sr, rnn_cell=rnn_cell, dropout_rate=drop,
                                                     bidir=bidir, sample_style=sample_style))
            width = out_dim * (2 if bidir else 1) * (sr if sample_style == 'concat' else 1)
            setattr(self, 'proj' + str(l), nn.Linear(width, width))
            input_dim = width

    def forward(self, input_x, enc_len):
        if self.vgg:
            input_x, enc_len = self.vgg_extractor(input_x, enc_len)
        for l in range(self.num_layers):
            input_x, _, enc_len = getattr(self, 'layer' + str(l))(input_x, state_len=enc_len, pack_input=True)
            proj = getattr(self, 'proj' + str(l))
            if input_x.is_cuda and input_x.dtype == torch.float32 and ENCODER_NATIVE:
                input_x = _keys.project(input_x, proj.weight, proj.bias)      # tanh(Linear) on the tcgen05 pipeline
            else:
                input_x = torch.tanh(proj(input_x))
        return input_x, enc_len


# --------------------------------------------------------------------------------------------
# Decoder RNN (library LSTMCell; API kept, state stays on the device)
# --------------------------------------------------------------------------------------------
class Speller(nn.Module):
    """Stack of recurrent cells stepped by hand (asr.py:320-357)."""

    def __init__(self, input_dim, dim, layer, rnn_cell, dropout):
        super().__init__()
        assert "Cell" in rnn_cell, 'Please use Recurrent Cell instead of layer in decoder'
        self.layer = layer
        self.dim = dim
        self.dropout = nn.Dropout(p=dropout)
        self.layer0 = getattr(nn, rnn_cell)(input_dim, dim)
        for i in range(1, layer):
            setattr(self, 'layer' + str(i), getattr(nn, rnn_cell)(dim, dim))
        self.state_list = []
        self.cell_list = []

    def init_rnn(self, context):
        z = torch.zeros(context.shape[0], self.dim, device=context.device)
        self.state_list = [z] * self.layer
        self.cell_list = [z] * self.layer

    @property
    def hidden_state(self):
        # the reference returns CPU clones (asr.py:342-344); kept for API compatibility only --
        # the device-resident beam search never goes through this property
        return [s.clone().detach().cpu() for s in self.state_list], [c.clone().detach().cpu() for c in self.cell_list]

    @hidden_state.setter
    def hidden_state(self, state):
        device = self.state_list[0].device
        self.state_list = [s.to(device) for s in state[0]]
        self.cell_list = [c.to(device) for c in state[1]]

    def forward(self, input_context):
        self.state_list[0], self.cell_list[0] = self.layer0(self.dropout(input_context),
                                                            (self.state_list[0], self.cell_list[0]))
        for l in range(1, self.layer):
            self.state_list[l], self.cell_list[l] = getattr(self, 'layer' + str(l))(
                self.state_list[l - 1], (self.dropout(self.state_list[l]), self.cell_list[l]))
        return self.state_list[-1]


# --------------------------------------------------------------------------------------------
# Attention (kernel group 2)
# --------------------------------------------------------------------------------------------
class _AttnSession:
    """Everything the per-step backward kernels accumulate into for one batch (one reset_enc_mem()
    to the next): dK, the per-CTA parameter-gradient slots and the stacks for the deferred GEMMs."""

    def __init__(self, mode, keys, h, lens_dev, params, scale, dec_dim):
        self.mode, self.keys, self.h, self.lens, self.p, self.scale, self.dec = mode, keys, h, lens_dev, params, scale, dec_dim
        self.d_keys = None
        self.pgrad = None
        self.atts, self.dctxs, self.dzqs, self.states = [], [], [], []

    def buffers(self):
        if self.d_keys is None:
            B, T, A = self.keys.shape
            self.d_keys = torch.zeros_like(self.keys)
            if self.mode == 'loc':
                C, Kw = self.p['conv'].shape[0], (self.p['conv'].shape[-1] - 1) // 2
                stride = _attn.pgrad_layout(A, C, Kw)[-1]
                slots = _attn.num_slots(B, T, self.h.shape[-1], A, self.dec, 'loc', C, Kw)
                self.pgrad = torch.zeros(slots, stride, dtype=torch.float32, device=self.keys.device)
        return self.d_keys, self.pgrad


class _AttnOpen(torch.autograd.Function):
    """Entry node of a session: hands out a token every step depends on, so that its backward runs
    after ALL step backwards and can return the accumulated / deferred gradients in one go."""

    @staticmethod
    def forward(ctx, session, keys, h, phi, conv, lproj, gen_w, gen_b):
        ctx.session = session
        ctx.set_materialize_grads(False)
        return torch.zeros(1, dtype=torch.float32, device=keys.device)

    @staticmethod
    def backward(ctx, _g):
        ses = ctx.session
        d_keys, pgrad = ses.buffers()
        g_keys = d_keys if ctx.needs_input_grad[1] else None
        g_h = g_phi = g_conv = g_lproj = g_gw = g_gb = None
        if ses.atts:
            att = torch.stack(ses.atts, dim=2)                     # [B,T,L]
            dctx = torch.stack(ses.dctxs, dim=1)                   # [B,L,D]
            if ctx.needs_input_grad[2]:
                g_h = torch.bmm(att, dctx)                         # dh = sum_steps alpha (x) d_ctx
            if ses.p['phi'] is not None and ctx.needs_input_grad[3]:
                dzq = torch.cat(ses.dzqs, dim=0)                   # [L*B,A]
                st = torch.cat(ses.states, dim=0)                  # [L*B,dec]
                g_phi = dzq.t().mm(st)
        if ses.mode == 'loc':
            A = ses.keys.shape[-1]
            C, KW1 = ses.p['conv'].shape[0], ses.p['conv'].shape[-1]
            o_wl, o_wc, o_wg, o_bg, _ = _attn.pgrad_layout(A, C, (KW1 - 1) // 2)
            tot = pgrad.sum(0)
            g_lproj = tot[o_wl:o_wc].view(A, C)
            g_conv = tot[o_wc:o_wg].view(C, 1, KW1)
            g_gw = tot[o_wg:o_bg].view(1, A)
            g_gb = tot[o_bg:o_bg + 1]
        ses.atts, ses.dctxs, ses.dzqs, ses.states = [], [], [], []
        return None, g_keys, g_h, g_phi, g_conv, g_lproj, g_gw, g_gb


class _AttnStep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, session, token, s, prev_att):
        ses = session
        p = ses.p
        att, cvec, q = _attn.step_forward(ses.mode, s, p['phi'], ses.keys, ses.h, ses.lens, prev_att,
                                          p['conv'], p['lproj'], p['gen_w'], p['gen_b'], ses.scale, want_q=True)
        ctx.session = ses
        ctx.has_prev = prev_att is not None
        ctx.save_for_backward(s, prev_att, att, q if q is not None else s)
        ctx.set_materialize_grads(False)
        return att, cvec

    @staticmethod
    def backward(ctx, d_att, d_ctx):
        ses = ctx.session
        s, prev_att, att, q = ctx.saved_tensors
        p = ses.p
        d_keys, pgrad = ses.buffers()
        if d_ctx is None:
            d_ctx = torch.zeros(att.shape[0], ses.h.shape[-1], dtype=torch.float32, device=att.device)
        d_prev, d_zq, d_s = _attn.step_backward(
            ses.mode, p['phi'], ses.keys, ses.h, ses.lens, prev_att if ctx.has_prev else None,
            p['conv'], p['lproj'], p['gen_w'], ses.scale, q, att, d_ctx, d_att, d_keys, pgrad, s.shape[1],
            want_prev=ctx.has_prev and ctx.needs_input_grad[3])
        ses.atts.append(att)
        ses.dctxs.append(d_ctx)
        ses.dzqs.append(d_zq)
        ses.states.append(s)
        return None, None, d_s, d_prev      # the token only orders _AttnOpen.backward after every step


class Attention(nn.Module):
    """Dot / location-aware attention of the reference (asr.py:370-462); one kernel launch per call.

    Input : decoder state [B, dec], listener feature [B, T, D], state_len list[int]
    Output: ([attention score [B, T]], context [B, D])
    """

    def __init__(self, in_dim, dec_dim, att_mode, dim, proj, num_head):
        super().__init__()
        self.mode = att_mode.lower()
        self.num_head = num_head
        self.softmax = nn.Softmax(dim=-1)
        self.proj = proj
        if proj:
            self.proj_dim = dim
            self.phi = nn.Linear(dec_dim, dim * num_head, bias=False)
            self.psi = nn.Linear(in_dim, dim)
        if self.num_head > 1:
            # the multi-head branch of the reference dereferences an undefined attribute
            # (asr.py:436, "UNTESTED" in the shipped configs); it is out of scope here
            raise ValueError('num_head > 1 is not supported (broken upstream, asr.py:434-442)')
        if self.mode == 'loc':
            assert self.proj, "Location-awared attetion requires proj==True"
            assert self.num_head == 1
            C, K = 10, 100                                   # asr.py:395-396
            self.prev_att = None
            self.loc_conv = nn.Conv1d(1, C, kernel_size=2 * K + 1, padding=K, bias=False)
            self.loc_proj = nn.Linear(C, dim, bias=False)
            self.gen_energy = nn.Linear(dim, 1)
        elif self.mode != 'dot':
            raise ValueError('Unsupported Attention Mode: ' + self.mode)
        self.comp_listener_feature = None
        self.state_mask = None
        self._session = None
        self._token = None
        self._lens_dev = None

    def __setstate__(self, state):
        """Whole-module checkpoints written by the UPSTREAM code (torch.save(self.asr_model), reference
        src/solver.py:283, loaded at :314) unpickle into this class through the `src` shim package; they lack the
        per-batch caches this mirror adds, and may hold a cached mask / key projection of the last training batch."""
        super().__setstate__(state)
        for name in ('comp_listener_feature', 'state_mask', '_session', '_token', '_lens_dev'):
            self.__dict__[name] = None
        if self.mode == 'loc':
            self.__dict__['prev_att'] = None

    def reset_enc_mem(self):
        self.comp_listener_feature = None
        self.state_mask = None
        self._session = None
        self._token = None
        self._lens_dev = None
        if self.mode == 'loc':
            self.prev_att = None

    def _params(self):
        loc = self.mode == 'loc'
        return dict(phi=self.phi.weight if self.proj else None,
                    conv=self.loc_conv.weight if loc else None,
                    lproj=self.loc_proj.weight if loc else None,
                    gen_w=self.gen_energy.weight if loc else None,
                    gen_b=self.gen_energy.bias if loc else None)

    def project_keys(self, listener_feature):
        """K = tanh(psi(h)) once per batch (asr.py:419)."""
        from . import _keys
        return _keys.project(listener_feature, self.psi.weight, self.psi.bias)

    def prepare(self, listener_feature, state_len):
        """First call after reset_enc_mem(): padding mask and cached key projection (asr.py:412-419)."""
        if self.comp_listener_feature is None:
            dev = listener_feature.device
            self._lens_dev = _attn.lens_i32(state_len, dev)
            ar = torch.arange(listener_feature.shape[1], device=dev)
            self.state_mask = ar[None, :] >= self._lens_dev[:, None]            # asr.py:415-418
            self.comp_listener_feature = self.project_keys(listener_feature) if self.proj else listener_feature
        return self.comp_listener_feature

    def forward(self, decoder_state, listener_feature, state_len, scale=2.0):
        if not decoder_state.is_cuda:
            raise RuntimeError('Attention runs on the sm_100a kernels only (no CPU path); move the model to a CUDA device')
        keys = self.prepare(listener_feature, state_len)
        p = self._params()
        prev = None
        if self.mode == 'loc' and self.prev_att is not None:
            prev = self.prev_att.reshape(self.prev_att.shape[0], -1)
            if prev.device != decoder_state.device:
                prev = prev.to(decoder_state.device)
        need_grad = torch.is_grad_enabled() and (decoder_state.requires_grad or keys.requires_grad or
                                                 listener_feature.requires_grad or
                                                 any(v is not None and v.requires_grad for v in p.values()))
        if need_grad:
            if self._session is None:
                det = {k: (v.detach() if v is not None else None) for k, v in p.items()}
                self._session = _AttnSession(self.mode, keys.detach().contiguous(), listener_feature.detach().contiguous(),
                                             self._lens_dev, det, float(scale), decoder_state.shape[1])
                self._token = _AttnOpen.apply(self._session, keys, listener_feature, p['phi'], p['conv'], p['lproj'],
                                              p['gen_w'], p['gen_b'])
            att, context = _AttnStep.apply(self._session, self._token, decoder_state, prev)
        else:
            att, context, _ = _attn.step_forward(self.mode, decoder_state.detach(), p['phi'], keys.detach(),
                                                 listener_feature.detach(), self._lens_dev, prev, p['conv'], p['lproj'],
                                                 p['gen_w'], p['gen_b'], float(scale))
        if self.mode == 'loc':
            self.prev_att = att.unsqueeze(1)                                     # asr.py:456
        return [att], context


# --------------------------------------------------------------------------------------------
# Seq2Seq
# --------------------------------------------------------------------------------------------
class Seq2Seq(nn.Module):
    """Listener + (Attention, Speller) + optional CTC head (asr.py:18-258)."""

    def __init__(self, example_input, output_dim, model_para):
        super().__init__()
        fp32_library_math()
        enc = model_para['encoder']
        enc_out_dim = int(enc['dim'].split('_')[-1]) * (2 if 'Bi' in enc['enc_type'] else 1) \
            * (int(enc['sample_rate'].split('_')[-1]) if enc['sample_style'] == 'concat' else 1)
        w = model_para['optimizer']['joint_ctc']
        self.joint_ctc = w > 0
        self.joint_att = w < 1
        self.encoder = Listener(example_input, **enc)
        if self.joint_att:
            self.dec_dim = model_para['decoder']['dim']
            self.attention = Attention(enc_out_dim, self.dec_dim, **model_para['attention'])
            self.decoder = Speller(enc_out_dim + self.dec_dim, **model_para['decoder'])
            self.embed = nn.Embedding(output_dim, self.dec_dim)
            self.char_dim = output_dim
            self.char_trans = nn.Linear(self.dec_dim, self.char_dim)
        if self.joint_ctc:
            self.ctc_weight = w
            self.ctc_layer = nn.Linear(enc_out_dim, output_dim)
        self.init_parameters()

    # -- reference API -------------------------------------------------------------------------
    def load_lm(self, decode_lm_weight, decode_lm_path, **kwargs):
        self.rnn_lm = torch.load(decode_lm_path, weights_only=False)
        self.rnn_lm.eval()

    def clear_att(self):
        self.attention.reset_enc_mem()

    def init_parameters(self):
        """LeCun-normal weights, zero biases, N(0,1) embedding, decoder forget-gate bias 1 (asr.py:114-153)."""
        with torch.no_grad():
            for prm in self.parameters():
                if prm.dim() == 1:
                    prm.zero_()
                elif prm.dim() in (2, 3, 4):
                    fan_in = prm[0].numel()
                    prm.normal_(0, 1.0 / math.sqrt(fan_in))
                else:
                    raise NotImplementedError
            if self.joint_att:
                self.embed.weight.normal_(0, 1)
                for i in range(self.decoder.layer):
                    bias = getattr(self.decoder, 'layer' + str(i)).bias_ih
                    n = bias.numel()
                    bias[n // 4:n // 2].fill_(1.0)

    def forward(self, audio_feature, decode_step, tf_rate=0.0, teacher=None, state_len=None):
        bs = audio_feature.shape[0]
        encode_feature, encode_len = self.encoder(audio_feature, state_len)
        ctc_output = att_output = att_maps = None
        if self.joint_ctc:
            ctc_output = _gemm.linear(self.ctc_layer, encode_feature)
        if self.joint_att:
            if teacher is not None:
                teacher = self.embed(teacher)
            self.decoder.init_rnn(encode_feature)
            self.attention.reset_enc_mem()
            if _dec.supported(self.decoder) and encode_feature.is_cuda:
                att_output, att_maps = self._decode_loop_fused(encode_feature, encode_len, decode_step, tf_rate, teacher)
                return ctc_output, encode_len, att_output, att_maps
            last_char = self.embed(torch.zeros(bs, dtype=torch.long, device=encode_feature.device))
            chars, maps = [], []
            for t in range(decode_step):
                score, context = self.attention(self.decoder.state_list[0], encode_feature, encode_len)
                dec_out = self.decoder(torch.cat([last_char, context], dim=-1))
                cur_char = self.char_trans(dec_out)
                if teacher is not None:
                    if random.random() <= tf_rate:                     # one coin per step (asr.py:96)
                        last_char = teacher[:, t + 1, :]
                    else:
                        last_char = self.embed(Categorical(F.softmax(cur_char, dim=-1)).sample())
                else:
                    last_char = self.embed(torch.argmax(cur_char, dim=-1))
                chars.append(cur_char)
                maps.append(score[0].detach())
            att_output = torch.stack(chars, dim=1)
            # the reference copies every step's map to the host (asr.py:107); here the stack stays on
            # the device and is copied only if somebody asks (draw_att does .cpu() itself)
            att_maps = [torch.stack(maps, dim=1)]
        return ctc_output, encode_len, att_output, att_maps

    def _decode_loop_fused(self, encode_feature, encode_len, decode_step, tf_rate, teacher):
        """The decoder loop of forward() (asr.py:84-108) for a single-layer LSTMCell Speller, regrouped as
        _dec.py describes: batched embedding projection before the loop, 3 launches per step around the
        attention kernel, batched char_trans and weight gradients after it.  Same arithmetic, same coins."""
        bs = encode_feature.shape[0]
        dev = encode_feature.device
        cell = self.decoder.layer0
        emb_dim = self.embed.embedding_dim
        w_emb = cell.weight_ih[:, :emb_dim]
        sos = self.embed(torch.zeros(bs, dtype=torch.long, device=dev))
        need_grad = torch.is_grad_enabled() and (encode_feature.requires_grad or cell.weight_ih.requires_grad)
        E = None
        if teacher is not None:
            x_emb = torch.cat([sos.unsqueeze(1), teacher[:, 1:decode_step, :]], dim=1) if decode_step > 1 else sos.unsqueeze(1)
            E = F.linear(x_emb, w_emb)                                  # [B, L, 4H]: every teacher-forced step at once
        h, c = self.decoder.state_list[0], self.decoder.cell_list[0]
        # the reference draws one coin per step inside the loop (asr.py:96); nothing else touches Python's
        # generator in between, so drawing the L coins up front yields the same sequence
        coins = [random.random() <= tf_rate for _ in range(decode_step)] if teacher is not None else None
        if (coins is not None and decode_step > 0 and self.attention.comp_listener_feature is None
                and os.environ.get("B2ASR_DECLOOP", "1") != "0"
                and (all(coins) or os.environ.get("B2ASR_DECLOOP_SAMPLING", "1") != "0")):
            return self._decode_loop_native(encode_feature, encode_len, E, h, c, None if all(coins) else coins)
        if need_grad:
            ses = _dec.DecSession(cell, E, emb_dim, decode_step)
            token = _dec.DecOpen.apply(ses, E, cell.weight_ih, cell.weight_hh, cell.bias_ih, cell.bias_hh)
        outs, chars, maps = [], [None] * decode_step, []
        own = None if teacher is not None else F.linear(sos, w_emb)     # embedding projection this step must use
        for t in range(decode_step):
            score, context = self.attention(h, encode_feature, encode_len)
            if need_grad:
                h, c = _dec.DecStep.apply(ses, token, t, own, context, h, c)
            else:
                h, c = _dec.step_nograd(cell, emb_dim, own if own is not None else E[:, t], context, h, c)
            self.decoder.state_list[0], self.decoder.cell_list[0] = h, c
            outs.append(h)
            own = None
            if teacher is not None:
                if not coins[t]:
                    chars[t] = self.char_trans(h)
                    own = F.linear(self.embed(Categorical(F.softmax(chars[t], dim=-1)).sample()), w_emb)
            else:
                chars[t] = self.char_trans(h)
                own = F.linear(self.embed(torch.argmax(chars[t], dim=-1)), w_emb)
            maps.append(score[0].detach())
        todo = [t for t in range(decode_step) if chars[t] is None]
        if len(todo) == decode_step:
            att_output = _gemm.linear(self.char_trans, torch.stack(outs, dim=1))      # one GEMM for every step
        else:
            if todo:
                late = self.char_trans(torch.stack([outs[t] for t in todo], dim=1)).unbind(1)
                for t, v in zip(todo, late):
                    chars[t] = v
            att_output = torch.stack(chars, dim=1)
        return att_output, [torch.stack(maps, dim=1)]

    def _decode_loop_native(self, encode_feature, encode_len, E, h0, c0, coins=None):
        """The whole loop and its backward are one native call each (_decloop.py, csrc/decloop.cu) when every step is
        teacher-forced; with scheduled sampling (coins: False = the next step gets a sampled character) the forward
        runs range by range between the sampled steps.  Leaves the module state (decoder state, prev_att) as the
        per-step loop would."""
        att_mod, cell = self.attention, self.decoder.layer0
        keys = att_mod.prepare(encode_feature, encode_len)
        p = att_mod._params()
        meta = dict(mode=att_mod.mode, lens=att_mod._lens_dev, scale=2.0, emb_dim=self.embed.embedding_dim,
                    coins=coins, char_trans=self.char_trans)
        hs, att, h_last, c_last = _decloop.DecoderLoop.apply(
            meta, keys, encode_feature, E, h0, c0, p['phi'], p['conv'], p['lproj'], p['gen_w'], p['gen_b'],
            cell.weight_ih, cell.weight_hh, cell.bias_ih, cell.bias_hh, self.embed.weight if coins is not None else None)
        self.decoder.state_list[0], self.decoder.cell_list[0] = h_last, c_last
        if att_mod.mode == 'loc':
            att_mod.prev_att = att[-1].unsqueeze(1)                              # asr.py:456
        return _gemm.linear(self.char_trans, hs), [att.transpose(0, 1)]

    def beam_decode(self, audio_feature, decode_step, state_len, decode_beam_size):
        """Joint CTC/attention/RNN-LM beam search, device-resident (asr.py:155-258); returns the top
        `decode_beam_size` Hypothesis objects sorted by mean score."""
        from . import beam
        return beam.beam_decode(self, audio_feature, decode_step, state_len, decode_beam_size)
