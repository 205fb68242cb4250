"""Experiment: per-step cost of the CTC lattice walk (loss-only = no row stores) vs with gradient."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ctc = importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
lib = importlib.import_module("end-to-end-asr-pytorch_b200._lib")
dev = torch.device("cuda:0")
B, V = 64, 5000
for T, L in [(1000, 150), (1000, 20), (250, 150), (3000, 600)]:
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, T, V, device=dev)
    il = torch.full((B,), T); tl = torch.full((B,), L)
    y = torch.randint(1, V, (B, L), generator=g).to(dev)
    for grad in (False, True):
        xr = x.clone().requires_grad_(grad)
        def run():
            if grad:
                xr.grad = None
                ctc.ctc_loss(xr, y, il, tl.to(dev), from_logits=True, batch_first=True).backward()
            else:
                with torch.no_grad():
                    ctc.ctc_loss(xr, y, il, tl.to(dev), from_logits=True, batch_first=True)
        for _ in range(2): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): run()
        e1.record(); torch.cuda.synchronize()
        print("T=%d L=%d grad=%s: %.3f ms" % (T, L, grad, e0.elapsed_time(e1) / 5))
