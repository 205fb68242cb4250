"""Wall clock of the train step's phases with a device synchronise at every boundary (so each phase's wall time is
max(CPU launch time, GPU time) of that phase alone), next to the CPU-only time to enqueue the phase.  Shows which
phases are launch-bound.  usage: python tools/phase_wall.py [workload]"""
import importlib, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
asr, solver, syn = (importlib.import_module(PKG + "." + m) for m in ("asr", "solver", "synthetic"))
name = sys.argv[1] if len(sys.argv) > 1 else "libri_m2"
w = syn.WORKLOADS[name]
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).train()
opt = solver.make_optimizer(model, w["cfg"]["optimizer"])
ts = solver.TrainStep(model, opt, w["cfg"]["optimizer"]["joint_ctc"])
x, y, _ = syn.make_batch(name, 1234)
x, y = x.to(dev), y.to(dev)
for _ in range(3):
    ts.step(x, y, 1.0)
torch.cuda.synchronize()

marks = {}


def timed(label, fn):
    def wrap(*a, **k):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn(*a, **k)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        m = marks.setdefault(label, [0.0, 0.0])
        m[0] += t1 - t0
        m[1] += t2 - t0
        return r
    return wrap


model.encoder.forward = timed("encoder forward", model.encoder.forward)
model._decode_loop_fused = timed("decoder loop forward", model._decode_loop_fused)
solver.hybrid_loss = timed("losses (CE + CTC kernels)", solver.hybrid_loss)
orig_backward = torch.Tensor.backward
torch.Tensor.backward = timed("backward (whole)", orig_backward)
ts.grads.clip_ = timed("clip", ts.grads.clip_)
opt.step = timed("optimizer", opt.step)
N = 3
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(N):
    ts.step(x, y, 1.0)
torch.cuda.synchronize()
tot = (time.perf_counter() - t0) / N
print("workload %s: synchronised step %.1f ms" % (name, tot * 1e3))
for k, (cpu, wall) in marks.items():
    print("  %-28s enqueue (CPU) %7.1f ms   phase wall %7.1f ms" % (k, cpu / N * 1e3, wall / N * 1e3))
