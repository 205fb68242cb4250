"""One forward + backward of a Listener layer at the M2 shape of layers 2-4 (B=32, T=400, 640 -> 2x320) between
cudaProfilerStart/Stop, for one `ncu --profile-from-start off --set full -k regex:lstm_cl` capture of the recurrence
kernels (tools/gpu_ncu_enc.sh)."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "end-to-end-asr-pytorch_b200"
enc = importlib.import_module(PKG + "._enc")
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, T, n_in, H = 32, 400, 640, 320
lens = sorted(torch.randint(int(T * 0.625), T + 1, (B,)).tolist(), reverse=True)
lens[0] = T
layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
x = torch.randn(B, T, n_in, device=dev)
for b, n in enumerate(lens):
    x[b, n:] = 0
x.requires_grad_(True)
w = torch.randn(B, T, 2 * H, device=dev)


def step():
    out = enc.lstm_layer(layer, x, lens)
    (out * w).sum().backward()


step(); step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
step()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
enc.check_status(dev)
print("ok")
