"""Generate tests/golden/*.pt by running the UNMODIFIED reference (/root/reference) in-process.

Run in the build container only (the reference is not on the GPU box):  python tools/make_golden.py
Every fixture stores the seeded inputs, the reference's state_dicts and the reference's outputs, so
both the CPU oracle (tests/test_oracle_golden.py) and the CUDA path (tests/test_*_gpu.py) can be
checked against what the reference itself computed.
"""
import os
import random
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import reference_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

TINY_CFG = {
    'optimizer': {'type': 'Adadelta', 'learning_rate': 1.0, 'joint_ctc': 0.5},
    'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2_1', 'sample_style': 'drop', 'dim': '16_16_16',
                'dropout': '0_0_0', 'rnn_cell': 'LSTM'},
    'attention': {'att_mode': 'loc', 'dim': 12, 'proj': True, 'num_head': 1},
    'decoder': {'dim': 16, 'layer': 1, 'dropout': 0.0, 'rnn_cell': 'LSTMCell'},
}


def seed_all(s):
    random.seed(s)
    np.random.seed(s)
    torch.manual_seed(s)


def make_batch(B, T, D, V, Lmax, seed):
    g = torch.Generator().manual_seed(seed)
    lens = sorted([int(v) for v in torch.randint(T // 2, T + 1, (B,), generator=g)], reverse=True)
    lens[0] = T
    x = torch.randn(B, T, D, generator=g)
    for b, n in enumerate(lens):
        x[b, n:] = 0
    y = torch.zeros(B, Lmax + 2, dtype=torch.long)
    for b in range(B):
        n = int(torch.randint(2, Lmax + 1, (1,), generator=g))
        toks = torch.randint(2, V, (n,), generator=g)
        for i in range(1, n):                       # no adjacent repeats: keeps the CTC alignment feasible
            if toks[i] == toks[i - 1]:
                toks[i] = 2 + (int(toks[i]) - 1) % (V - 2)
        y[b, 1:n + 1] = toks
        y[b, n + 1] = 1
    return x, y, lens


def attention_fixture(ref, mode):
    seed_all(3)
    B, T, D, dec, A = 3, 37, 24, 16, 12
    att = ref.asr.Attention(D, dec, mode, A, True, 1)
    g = torch.Generator().manual_seed(4)
    h = torch.randn(B, T, D, generator=g)
    lens = [37, 30, 19]
    states = [torch.randn(B, dec, generator=g) for _ in range(3)]
    outs = []
    att.reset_enc_mem()
    with torch.no_grad():
        for s in states:
            score, ctx = att(s, h, lens)
            outs.append((score[0].clone(), ctx.clone()))
        keys = att.comp_listener_feature.clone()
    return dict(mode=mode, state_dict=att.state_dict(), h=h, lens=lens, states=states, keys=keys,
                att=[o[0] for o in outs], ctx=[o[1] for o in outs], dims=(D, dec, A))


def prefix_fixture(ref):
    seed_all(5)
    T, V = 25, 9
    x = F.log_softmax(torch.randn(1, T, V) * 2, dim=-1)
    sc = ref.ctc.CTCPrefixScore(x)
    r0 = sc.init_state()
    calls = []
    prefix, r_prev = [], r0
    for step, cands in enumerate([[2, 5, 0, 7], [5, 3, 1, 8, 2], [3, 3, 6], [6, 4, 2, 1]]):
        if step == 2:
            cands = [prefix[-1], 4, 6]          # repeated-label case (ctc.py:88-91)
        psi, r = sc.cheap_compute(prefix, r_prev, cands)
        calls.append(dict(prefix=list(prefix), r_prev=r_prev.copy(), cands=list(cands), psi=psi.copy(), r=r.copy()))
        prefix = prefix + [cands[0]]
        r_prev = r[0]
    psi_full, r_full = sc.full_compute([2], calls[1]['r_prev'])
    return dict(x=x[0].clone(), init=r0.copy(), calls=calls, full=dict(prefix=[2], r_prev=calls[1]['r_prev'],
                                                                       psi=psi_full.copy(), r=np.ascontiguousarray(r_full)))


def model_fixture(ref):
    """Tiny reference Seq2Seq: teacher-forced forward + hybrid loss + gradients (solver.py:132-177)."""
    seed_all(7)
    V, D = 11, 13
    x, y, lens = make_batch(4, 48, D, V, 4, seed=8)
    model = ref.asr.Seq2Seq(x, V, TINY_CFG)
    model.train()
    ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
    ctc_pred, enc_len, att_pred, att_maps = model(x, ans_len, tf_rate=1.0, teacher=y, state_len=lens)
    label = y[:, 1:ans_len + 1].contiguous()
    b, t, c = att_pred.shape
    seq_loss = torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none')
    att_loss = seq_loss(att_pred.view(b * t, c), label.view(-1))
    att_loss = torch.mean(torch.sum(att_loss.view(b, t), dim=-1) / torch.sum(y != 0, dim=-1).float())
    tl = torch.sum(y != 0, dim=-1)
    ctc_loss = torch.nn.CTCLoss(blank=0, reduction='mean')(F.log_softmax(ctc_pred.transpose(0, 1), dim=-1), label,
                                                           torch.LongTensor(enc_len), tl)
    loss = 0.5 * att_loss + 0.5 * ctc_loss
    loss.backward()
    grads = {k: p.grad.clone() for k, p in model.named_parameters()}
    return dict(cfg=TINY_CFG, vocab=V, x=x, y=y, lens=lens, state_dict={k: v.clone() for k, v in model.state_dict().items()},
                ctc_pred=ctc_pred.detach(), att_pred=att_pred.detach(), att_maps=att_maps[0].detach(), enc_len=enc_len,
                att_loss=att_loss.detach(), ctc_loss=ctc_loss.detach(), loss=loss.detach(), grads=grads)


def beam_fixture(ref, ctc_w, lm_w, beam, seed):
    """Reference Seq2Seq.beam_decode on a tiny random model (+ RNN-LM), asr.py:155-258."""
    seed_all(seed)
    V, D = 11, 13
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(1, 40, D, generator=g)
    model = ref.asr.Seq2Seq(x, V, TINY_CFG)
    with torch.no_grad():                       # sharpen the heads so hypotheses differ in score
        model.char_trans.weight.mul_(3.0)
        model.ctc_layer.weight.mul_(3.0)
    model.eval()
    model.joint_ctc = ctc_w > 0
    model.ctc_weight = ctc_w
    model.decode_lm_weight = lm_w
    lm = ref.rnnlm.RNN_LM(8, 8, V, layers=2)
    lm.eval()
    model.rnn_lm = lm
    with torch.no_grad():
        hyps = model.beam_decode(x, 7, [40], beam)
    res = [dict(tokens=h.outIndex, avg=float(h.avgScore()), scores=[float(s) for s in h.output_scores]) for h in hyps]
    return dict(cfg=TINY_CFG, vocab=V, x=x, state_len=[40], decode_step=7, beam=beam, ctc_weight=ctc_w, lm_weight=lm_w,
                state_dict={k: v.clone() for k, v in model.state_dict().items() if not k.startswith('rnn_lm.')},
                lm_state_dict={k: v.clone() for k, v in lm.state_dict().items()}, lm_dims=(8, 8, V, 2), hyps=res)


def seqloss_fixture(ref):
    """The attention branch's loss exactly as the reference forms it (solver.py:90,149-155) plus the reference's
    own cal_acc (postprocess.py:121-133) on the same logits; two vocabulary sizes, padded labels, one tie row."""
    seed_all(11)
    out = {}
    for name, (B, L, V) in {'small': (4, 9, 31), 'wide': (3, 6, 1301)}.items():
        att_pred = (torch.randn(B, L, V) * 3).requires_grad_(True)
        lens = [L] + [int(v) for v in torch.randint(1, L + 1, (B - 1,))]
        y = torch.zeros(B, L + 1, dtype=torch.long)
        for b, n in enumerate(lens):
            y[b, 1:n] = torch.randint(2, V, (n - 1,))
            y[b, n] = 1                                                  # <eos>
        with torch.no_grad():
            hit = torch.rand(B, L) < 0.5                                 # make about half the steps "correct"
            for b in range(B):
                for t in range(L):
                    if hit[b, t]:
                        att_pred[b, t, y[b, t + 1]] += 20.0
            att_pred[0, 1, 5] = att_pred[0, 1, 17] = att_pred[0, 1].max() + 1    # tie: argmax takes the first
        ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
        label = y[:, 1:ans_len + 1].contiguous()
        seq_loss = torch.nn.CrossEntropyLoss(ignore_index=0, reduction='none')      # solver.py:90
        b, t, c = att_pred.shape
        rows = seq_loss(att_pred.view(b * t, c), label.view(-1))                    # solver.py:151
        att_loss = torch.sum(rows.view(b, t), dim=-1) / torch.sum(y != 0, dim=-1).to(dtype=torch.float32)
        att_loss = torch.mean(att_loss)                                             # solver.py:154
        w = 0.5
        ((1 - w) * att_loss).backward()                                             # solver.py:163,177
        out[name] = dict(att_pred=att_pred.detach().clone(), y=y, label=label, rows=rows.detach().view(b, t),
                         att_loss=att_loss.detach(), grad=att_pred.grad.clone(), upstream=1 - w,
                         acc=float(ref.postprocess.cal_acc(att_pred, label)),
                         argmax=torch.as_tensor(np.argmax(att_pred.detach().numpy(), axis=-1)))
    return out


def main():
    ref = reference_loader.load()
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1 and sys.argv[1] == 'seqloss':      # add this fixture without regenerating the others
        torch.save(seqloss_fixture(ref), os.path.join(OUT, 'seqloss.pt'))
        return
    torch.save(attention_fixture(ref, 'loc'), os.path.join(OUT, 'attention_loc.pt'))
    torch.save(attention_fixture(ref, 'dot'), os.path.join(OUT, 'attention_dot.pt'))
    torch.save(prefix_fixture(ref), os.path.join(OUT, 'prefix.pt'))
    torch.save(model_fixture(ref), os.path.join(OUT, 'trainstep.pt'))
    torch.save(beam_fixture(ref, 0.5, 0.5, 4, 21), os.path.join(OUT, 'beam_joint.pt'))
    torch.save(beam_fixture(ref, 0.0, 0.0, 3, 22), os.path.join(OUT, 'beam_att_only.pt'))
    torch.save(beam_fixture(ref, 0.3, 0.0, 5, 23), os.path.join(OUT, 'beam_ctc.pt'))
    torch.save(seqloss_fixture(ref), os.path.join(OUT, 'seqloss.pt'))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == '__main__':
    main()
