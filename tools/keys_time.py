"""Device time of the key projection (csrc/keys2.cu) at M2 without the host in the way: 12 launches over six rotating
inputs (6 x 33 MB > L2) captured in one CUDA graph, replayed and timed with events.  usage: python tools/keys_time.py"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
keys = importlib.import_module("end-to-end-asr-pytorch_b200._keys")
dev = torch.device("cuda:0")
M, D, A = 12800, 640, 300
hs = [torch.randn(32, 400, D, device=dev) for _ in range(6)]
w = torch.randn(A, D, device=dev) * D ** -0.5
b = torch.randn(A, device=dev)
N = 12
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for i in range(3):
        keys.project(hs[i], w, b)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        outs = [keys.project(hs[i % 6], w, b) for i in range(N)]
for _ in range(2):
    graph.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
R = 10
for _ in range(R):
    graph.replay()
e1.record()
torch.cuda.synchronize()
t = e0.elapsed_time(e1) / (R * N) * 1e-3
fl, by = 2.0 * M * D * A, 4.0 * M * (D + A) + 4 * A * (D + 1)
print("keys M2: %.1f us per launch (graph replay, back to back)  %.1f TFLOP/s algorithmic (x3 issued), %.0f GB/s" % (t * 1e6, fl / t / 1e12, by / t / 1e9))
