"""CTC kernels alone at an M5 point (V=5000, B=64) for ncu.  usage: python tools/profile_ctc.py [T] [L]"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ctc = importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
T = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 150
B, V = 64, 5000
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
x = torch.randn(B, T, V, device=dev).requires_grad_(True)
il = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); il[0] = T
tl = torch.randint(L // 2, L + 1, (B,), generator=g); tl[0] = L
y = torch.randint(1, V, (B, L), generator=g)
for b in range(B):
    y[b, tl[b]:] = 0
yd, tld, ild = y.to(dev), tl.to(dev), il.to(dev)
for i in range(3):
    if i == 2:
        torch.cuda.synchronize(); torch.cuda.cudart().cudaProfilerStart()
    x.grad = None
    loss = ctc.ctc_loss(x, yd, ild, tld, from_logits=True, batch_first=True)
    loss.backward()
torch.cuda.synchronize(); torch.cuda.cudart().cudaProfilerStop()
print("loss", float(loss))
