"""Reads the per-role clock sums of a -DK2_TRACE build of csrc/keys2.cu (CTA 0, M2 shape).
usage: B2ASR_LIB=.../var/libb2asr_trace.so python tools/keys_trace.py"""
import ctypes
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
keys = importlib.import_module("end-to-end-asr-pytorch_b200._keys")
lib = importlib.import_module("end-to-end-asr-pytorch_b200._lib")
dev = torch.device("cuda:0")
h = torch.randn(32, 400, 640, device=dev)
w = torch.randn(300, 640, device=dev) * 640 ** -0.5
b = torch.randn(300, device=dev)
for _ in range(3):
    keys.project(h, w, b)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 64)()
L = ctypes.CDLL(lib.LIB_PATH)
assert L.b2asr_keys_trace(buf) == 0
names = {0: "mma: wait acc_empty", 1: "mma: wait conv", 2: "mma: issue 12 MMAs", 3: "mma: commit",
         8: "conv: wait tmem slot", 9: "conv: wait full", 10: "conv: convert+arrive",
         16: "epi: wait acc_full", 17: "epi: tmem loads", 18: "epi: tanh+stores"}
for k, n in names.items():
    print("%-24s %8d clk total, %6.0f clk per k-block (60 per CTA)" % (n, buf[k], buf[k] / 60.0))
