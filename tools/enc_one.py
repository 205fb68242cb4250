"""One forward+backward of an M2 Listener layer (T=400, B=32, 640->2x320) for ncu captures of the recurrence kernels."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
enc = importlib.import_module("end-to-end-asr-pytorch_b200._enc")
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, T, n_in, H = 32, 400, 640, 320
lens = sorted(torch.randint(250, T + 1, (B,)).tolist(), reverse=True)
lens[0] = T
layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
x = torch.randn(B, T, n_in, device=dev, requires_grad=True)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    out = enc.lstm_layer(layer, x, lens)
    out.square().sum().backward()
torch.cuda.synchronize()
enc.check_status(dev)
print("ok")
