"""Experiment: cuDNN BiLSTM cost, packed (reference semantics) vs padded, fwd+bwd, encoder layer shapes of M2."""
import time
import torch
import torch.nn as nn
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

dev = torch.device("cuda:0")
torch.manual_seed(0)


def t(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


for (T, D, H) in [(1600, 120, 320), (800, 640, 320), (400, 640, 320)]:
    B = 32
    lstm = nn.LSTM(D, H, bidirectional=True, batch_first=True).to(dev)
    uni = nn.LSTM(D, H, bidirectional=False, batch_first=True).to(dev)
    x = torch.randn(B, T, D, device=dev, requires_grad=True)
    lens = sorted(torch.randint(int(T * 0.6), T + 1, (B,)).tolist(), reverse=True)
    lens[0] = T

    def packed():
        p = pack_padded_sequence(x, lens, batch_first=True)
        o, _ = lstm(p)
        o, _ = pad_packed_sequence(o, batch_first=True)
        o.sum().backward()

    def padded():
        o, _ = lstm(x)
        o.sum().backward()

    def padded_uni2():
        o1, _ = uni(x)
        o2, _ = uni(x.flip(1))
        (o1.sum() + o2.sum()).backward()

    print("T=%d D=%d: packed bi %.1f ms | padded bi %.1f ms | 2x padded uni %.1f ms" % (T, D, t(packed), t(padded), t(padded_uni2)))
    with torch.no_grad():
        def fwd_packed():
            p = pack_padded_sequence(x, lens, batch_first=True)
            lstm(p)
        def fwd_padded():
            lstm(x)
        print("   fwd only: packed %.1f ms | padded %.1f ms" % (t(fwd_packed), t(fwd_padded)))
