"""Debug probe for csrc/gemm_nt.cu: one-hot a, ramp b -> which (k, n) element does each output column see?"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gm = importlib.import_module("end-to-end-asr-pytorch_b200._gemm")
gm.MIN_WORK = 0
dev = torch.device("cuda:0")
Kr, Ma, Nb = 32, 128, 64
k = torch.arange(Kr, device=dev, dtype=torch.float32)[:, None]
n = torch.arange(Nb, device=dev, dtype=torch.float32)[None, :]
b = (k * 100 + n).contiguous()
for k0, m0 in ((0, 0), (1, 0), (5, 3), (9, 40), (31, 127)):
    a = torch.zeros(Kr, Ma, device=dev)
    a[k0, m0] = 1.0
    out = gm.mm_nt(a, b)
    row = out[m0].cpu()
    nz = (out.abs().sum(1) > 0).nonzero().flatten().tolist()
    print("k0=%d m0=%d: nonzero rows %s; row[m0][:10]=%s  expect %s" % (k0, m0, nz[:6], [int(v) for v in row[:10].tolist()], [int(v) for v in b[k0, :10].tolist()]))
    print("     row[m0][30:36]=%s row[m0][60:64]=%s" % ([int(v) for v in row[30:36].tolist()], [int(v) for v in row[60:64].tolist()]))
out = gm.mm_nt(torch.ones(Kr, Ma, device=dev), torch.ones(Kr, Nb, device=dev))
print("ones x ones: min %.1f max %.1f nan %d (expect 32)" % (float(out.min()), float(out.max()), int(torch.isnan(out).sum())))
# all-ones a: column sums of b
a = torch.ones(Kr, Ma, device=dev)
out = gm.mm_nt(a, b)
print("ones: out[0][:8]", out[0][:8].tolist(), "expect", b.sum(0)[:8].tolist())
