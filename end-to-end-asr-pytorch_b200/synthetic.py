"""Synthetic workloads of SURVEY.md section 8d (shapes of the reference's shipped configs; there is no
corpus on the box).  Pure host code: returns CPU tensors; callers move them.

  M1  config/timit_example.yaml-shaped : 39-dim MFCC, V=63 phones, BiRNN 3x256 '2_2_1' concat, batch 16
  M2  config/libri_example.yaml-shaped : 120-dim fbank+delta, V=31 chars, pBLSTM 5x320 '2_2_1_1_1' drop,
                                         loc attention 300, decoder 320, joint_ctc 0.5, batch 32, T<=1600
  M2S the shipped libri_example.yaml verbatim (VGGBiRNN, 80-dim, V=5000 subwords)
  M4  decode: M2S model + RNN_LM(650,650,2 layers), beam 20, ctc 0.5, lm 0.5
"""
import copy

import torch

M1_CFG = {
    'optimizer': {'type': 'Adadelta', 'learning_rate': 1.0, 'joint_ctc': 0.5},
    'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2_1', 'sample_style': 'concat', 'dim': '256_256_256',
                'dropout': '0_0_0', 'rnn_cell': 'LSTM'},
    'attention': {'att_mode': 'loc', 'dim': 256, 'proj': True, 'num_head': 1},
    'decoder': {'dim': 256, 'layer': 1, 'dropout': 0.0, 'rnn_cell': 'LSTMCell'},
}
M2_CFG = {
    'optimizer': {'type': 'Adadelta', 'learning_rate': 1.0, 'joint_ctc': 0.5},
    'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2_1_1_1', 'sample_style': 'drop',
                'dim': '320_320_320_320_320', 'dropout': '0_0_0_0_0', 'rnn_cell': 'LSTM'},
    'attention': {'att_mode': 'loc', 'dim': 300, 'proj': True, 'num_head': 1},
    'decoder': {'dim': 320, 'layer': 1, 'dropout': 0.0, 'rnn_cell': 'LSTMCell'},
}
M2S_CFG = copy.deepcopy(M2_CFG)
M2S_CFG['encoder'].update(enc_type='VGGBiRNN', sample_rate='1_1_1_1_1')

WORKLOADS = {
    # name: (model config, feature dim, vocab, batch, T range, label range as fractions of T')
    'timit_m1': dict(cfg=M1_CFG, d_in=39, vocab=63, batch=16, t_lo=200, t_hi=780, sub=4, l_lo=1 / 8, l_hi=1 / 3),
    'libri_m2': dict(cfg=M2_CFG, d_in=120, vocab=31, batch=32, t_lo=1000, t_hi=1600, sub=4, l_lo=1 / 4, l_hi=1 / 2),
    'libri_m2s': dict(cfg=M2S_CFG, d_in=80, vocab=5000, batch=32, t_lo=1000, t_hi=1600, sub=4, l_lo=30, l_hi=100),
}


def make_batch(workload, seed, batch=None, shape_seed=None):
    """-> x [B,T,D] fp32 (padded frames exactly 0, no valid frame sums to 0), y [B,Lmax+2] int64
    (y[:,0]=<sos>=0, tokens in 2..V-1, <eos>=1, right padded 0), frame lengths (sorted descending).
    shape_seed: draw the frame and label LENGTHS from that seed and only the content from `seed` -- data-parallel ranks
    then hold batches of identical shape (fixed work per GPU: weak scaling) with different content."""
    w = WORKLOADS[workload]
    B = batch or w['batch']
    if shape_seed is not None and shape_seed != seed:
        _, y0, lens = make_batch(workload, shape_seed, batch)
        labs = [int(v) for v in (y0 > 1).sum(-1)]
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(B, w['t_hi'], w['d_in'], generator=g)
        for b, n in enumerate(lens):
            x[b, n:] = 0
        y = torch.zeros_like(y0)
        for b, n in enumerate(labs):
            y[b, 1:n + 1] = torch.randint(2, w['vocab'], (n,), generator=g)
            y[b, n + 1] = 1
        return x, y, lens
    g = torch.Generator().manual_seed(seed)
    T = w['t_hi']
    lens = sorted([int(v) for v in torch.randint(w['t_lo'], T + 1, (B,), generator=g)], reverse=True)
    lens[0] = T
    x = torch.randn(B, T, w['d_in'], generator=g)
    for b, n in enumerate(lens):
        x[b, n:] = 0
    labs = []
    for n in lens:
        tp = n // w['sub']
        lo = int(w['l_lo'] * tp) if w['l_lo'] < 1 else int(w['l_lo'])
        hi = int(w['l_hi'] * tp) if w['l_hi'] <= 1 else int(w['l_hi'])
        lo, hi = max(1, lo), max(1, hi)
        labs.append(int(torch.randint(lo, hi + 1, (1,), generator=g)))
    Lmax = max(labs)
    y = torch.zeros(B, Lmax + 2, dtype=torch.long)
    for b, n in enumerate(labs):
        y[b, 1:n + 1] = torch.randint(2, w['vocab'], (n,), generator=g)
        y[b, n + 1] = 1
    return x, y, lens


def encoder_lengths(workload, lens):
    sub = WORKLOADS[workload]['sub']
    out = list(lens)
    k = sub
    while k > 1:
        out = [int(n / 2) for n in out]
        k //= 2
    return out
