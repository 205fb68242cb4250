import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ctc = importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
T = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 20
B, V = 64, 5000
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
x = torch.randn(B, T, V, device=dev)
il = torch.full((B,), T); tl = torch.full((B,), L).to(dev)
y = torch.randint(1, V, (B, L), generator=g).to(dev)
with torch.no_grad():
    for i in range(3):
        if i == 2:
            torch.cuda.synchronize(); torch.cuda.cudart().cudaProfilerStart()
        loss = ctc.ctc_loss(x, y, il, tl, from_logits=True, batch_first=True)
torch.cuda.synchronize(); torch.cuda.cudart().cudaProfilerStop()
print(float(loss))
