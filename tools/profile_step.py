"""One train step of the bench workload between cudaProfilerStart/Stop (for `ncu --profile-from-start off`).
usage: python tools/profile_step.py [warm_steps]"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
asr, solver, syn = (importlib.import_module(PKG + "." + m) for m in ("asr", "solver", "synthetic"))

warm = int(sys.argv[1]) if len(sys.argv) > 1 else 2
name = os.environ.get("B2ASR_WORKLOAD", "libri_m2")
w = syn.WORKLOADS[name]
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).train()
opt = solver.make_optimizer(model, w["cfg"]["optimizer"])
ts = solver.TrainStep(model, opt, w["cfg"]["optimizer"]["joint_ctc"])
x, y, _ = syn.make_batch(name, 1234)
x, y = x.to(dev), y.to(dev)
for _ in range(warm):
    ts.step(x, y, 1.0)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
out = ts.step(x, y, 1.0)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("loss", float(out["loss"]))
