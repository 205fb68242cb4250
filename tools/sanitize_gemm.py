"""Small launches of the tcgen05 kernels against float64: key projection, TN and NT products incl. partial tiles, split-K,
a capped grid and the masked shifted views -- the cases to run first after touching csrc/keys2.cu / gemm_nt.cu (and
under compute-sanitizer where that is available).  usage: python tools/sanitize_gemm.py"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gm = importlib.import_module("end-to-end-asr-pytorch_b200._gemm")
keys = importlib.import_module("end-to-end-asr-pytorch_b200._keys")
gm.MIN_WORK = 0
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
worst = 0.0
for M, N, K in ((300, 168, 100), (129, 81, 36), (1000, 336, 640)):
    a = torch.randn(M, K, generator=g).to(dev)
    b = (torch.randn(N, K, generator=g) * K ** -0.5).to(dev)
    c = torch.randn(N, generator=g).to(dev)
    worst = max(worst, float((gm.mm_tn(a, b, c) - (a.double() @ b.double().t() + c.double()).float()).abs().max()))
    worst = max(worst, float((keys.project(a, b, c) - torch.tanh(a.double() @ b.double().t() + c.double()).float()).abs().max()))
for Kr, Ma, Nb in ((1000, 132, 72), (4100, 256, 64), (96, 128, 64)):
    a = (torch.randn(Kr, Ma, generator=g) * Kr ** -0.5).to(dev)
    b = torch.randn(Kr, Nb, generator=g).to(dev)
    worst = max(worst, float((gm.mm_nt(a, b) - (a.double().t() @ b.double()).float()).abs().max()))
    worst = max(worst, float((gm.mm_nt(a, b, max_ctas=3) - (a.double().t() @ b.double()).float()).abs().max()))
B, T, H = 5, 40, 32
dG = (torch.randn(B, T, 2, 4 * H, generator=g) * 0.1).to(dev)
out = torch.randn(B, T, 2 * H, generator=g).to(dev)
dG2, out2 = dG.view(B * T, 8 * H), out.view(B * T, 2 * H)
got = gm.mm_nt(dG2[1:, :4 * H], out2[:-1, :H], period=T)
ref = torch.bmm(dG[:, 1:, 0].double().transpose(1, 2), out[:, :T - 1, :H].double()).sum(0).float()
worst = max(worst, float((got - ref).abs().max()))
torch.cuda.synchronize()
print("max abs error %.2e" % worst)
assert worst < 1e-4
