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


MID_CFG = {
    'optimizer': {'type': 'Adadelta', 'learning_rate': 1.0, 'joint_ctc': 0.5},
    'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2', 'sample_style': 'drop', 'dim': '320_320',
                'dropout': '0_0', 'rnn_cell': 'LSTM'},
    'attention': {'att_mode': 'loc', 'dim': 300, 'proj': True, 'num_head': 1},
    'decoder': {'dim': 320, 'layer': 1, 'dropout': 0.0, 'rnn_cell': 'LSTMCell'},
}
MID_SEED = 31


def midsize_fixture(ref, att_mode):
    """The reference's teacher-forced forward + hybrid loss + backward (asr.py:58-112, solver.py:149-163,177) at the
    layer widths the shipped configs use (attention 300, decoder 320, D = 640 encoder features, T' = 100 frames, 30
    decoder steps), so that the GPU test exercises the real tiling of decloop.cu / deccell.cu / attn_*.cu against the
    reference rather than against the per-step GPU path.  Weights are oracle/seeded.reseed_parameters(MID_SEED);
    large gradients are stored as oracle/seeded.summarise_gradient samples."""
    from oracle import seeded
    import copy
    seed_all(13)
    cfg = copy.deepcopy(MID_CFG)
    cfg['attention']['att_mode'] = att_mode
    B, T, D, V = 4, 400, 40, 31
    g = torch.Generator().manual_seed(14)
    lens = [400, 372, 316, 248]
    x = torch.randn(B, T, D, generator=g)
    for b, n in enumerate(lens):
        x[b, n:] = 0
    label_lens = [29, 26, 21, 14]                  # + <eos> -> 30 decoder steps
    y = torch.zeros(B, max(label_lens) + 2, dtype=torch.long)
    for b, n in enumerate(label_lens):
        toks = torch.randint(2, V, (n,), generator=g)
        y[b, 1:n + 1] = toks
        y[b, n + 1] = 1
    model = ref.asr.Seq2Seq(x, V, cfg)
    seeded.reseed_parameters(model, MID_SEED)
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
    grads = {k: seeded.summarise_gradient(k, p.grad, MID_SEED) for k, p in model.named_parameters()}
    return dict(cfg=cfg, vocab=V, seed=MID_SEED, x=x, y=y, lens=lens,
                param_shapes={k: tuple(p.shape) for k, p in model.named_parameters()},
                ctc_pred=ctc_pred.detach(), att_pred=att_pred.detach(), att_maps=att_maps[0].detach(), enc_len=enc_len,
                att_loss=att_loss.detach(), ctc_loss=ctc_loss.detach(), loss=loss.detach(), grads=grads)


FIXTURES = {
    'attention_loc.pt': lambda ref: attention_fixture(ref, 'loc'),
    'attention_dot.pt': lambda ref: attention_fixture(ref, 'dot'),
    'prefix.pt': prefix_fixture,
    'trainstep.pt': model_fixture,
    'beam_joint.pt': lambda ref: beam_fixture(ref, 0.5, 0.5, 4, 21),
    'beam_att_only.pt': lambda ref: beam_fixture(ref, 0.0, 0.0, 3, 22),
    'beam_ctc.pt': lambda ref: beam_fixture(ref, 0.3, 0.0, 5, 23),
    'seqloss.pt': seqloss_fixture,
    'midsize_loc.pt': lambda ref: midsize_fixture(ref, 'loc'),
    'midsize_dot.pt': lambda ref: midsize_fixture(ref, 'dot'),
}


def deep_equal(a, b):
    if isinstance(a, dict):
        return isinstance(b, dict) and a.keys() == b.keys() and all(deep_equal(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)):
        return type(a) is type(b) and len(a) == len(b) and all(deep_equal(u, v) for u, v in zip(a, b))
    if torch.is_tensor(a):
        return torch.is_tensor(b) and a.dtype == b.dtype and a.shape == b.shape and \
            bool((a.reshape(-1).view(torch.uint8) == b.reshape(-1).view(torch.uint8)).all() if a.numel() else True)
    if isinstance(a, np.ndarray):
        return isinstance(b, np.ndarray) and a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes()
    return a == b


def check():
    """Regenerate every fixture from the reference and compare it bit for bit with the committed file."""
    ref = reference_loader.load()
    assert ref.asr.__file__.startswith(reference_loader.UPSTREAM_ROOT + os.sep), ref.asr.__file__
    bad = []
    for name, fn in FIXTURES.items():
        want = torch.load(os.path.join(OUT, name), weights_only=False)
        same = deep_equal(fn(ref), want)
        print("%-20s %s" % (name, "bit-identical" if same else "DIFFERS"))
        if not same:
            bad.append(name)
    return bad


def main():
    if len(sys.argv) > 1 and sys.argv[1] == '--check':
        sys.exit(1 if check() else 0)
    if len(sys.argv) > 1:                                    # python tools/make_golden.py midsize_loc.pt ...
        ref = reference_loader.load()
        for name in sys.argv[1:]:
            torch.save(FIXTURES[name](ref), os.path.join(OUT, name))
            print(name, os.path.getsize(os.path.join(OUT, name)))
        return
    main_all()


def main_all():
    ref = reference_loader.load()
    assert ref.asr.__file__.startswith(reference_loader.UPSTREAM_ROOT + os.sep), ref.asr.__file__
    os.makedirs(OUT, exist_ok=True)
    for name, fn in FIXTURES.items():
        torch.save(fn(ref), os.path.join(OUT, name))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == '__main__':
    main()
