"""Is the train step GPU-bound or launch-bound?  Sum of kernel durations (torch.profiler, warm, concurrent-capable)
against the step's wall clock, split by kernel family."""
import importlib, os, sys, time
import torch
from torch.profiler import profile, ProfilerActivity
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
asr, solver, syn = (importlib.import_module(PKG + "." + m) for m in ("asr", "solver", "synthetic"))
name = os.environ.get("B2ASR_WORKLOAD", "libri_m2")
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
t0 = time.perf_counter()
for _ in range(3):
    ts.step(x, y, 1.0)
torch.cuda.synchronize()
print("plain step %.1f ms" % ((time.perf_counter() - t0) / 3 * 1e3))
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    t0 = time.perf_counter()
    ts.step(x, y, 1.0)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = sum(e.device_time for e in ev)
print("profiled step wall %.1f ms; %d GPU kernels/memcpys, summed GPU time %.1f ms" % (wall * 1e3, len(ev), tot / 1e3))
fam = {}
for e in ev:
    n = e.name
    k = ("b2asr" if "b2asr" in n else "cudnn-rnn" if ("RNN" in n or "LSTM_elem" in n or "elemWise" in n) else
         "gemm" if ("gemm" in n or "cutlass" in n or "splitK" in n) else "other")
    a = fam.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += e.device_time
for k, (c, t) in sorted(fam.items(), key=lambda kv: -kv[1][1]):
    print("  %-10s %6d launches %8.1f ms" % (k, c, t / 1e3))
# time span of GPU activity vs summed: gaps = idle
ev.sort(key=lambda e: e.time_range.start)
span = ev[-1].time_range.end - ev[0].time_range.start
print("GPU activity span %.1f ms -> idle inside the span %.1f ms" % (span / 1e3, (span - tot) / 1e3))
# union of the kernel intervals (streams overlap since the weight-gradient side stream): true idle, and where
busy, cur_s, cur_e, gaps = 0.0, None, None, []
for e in ev:
    s0, e0 = e.time_range.start, e.time_range.end
    if cur_e is None or s0 > cur_e:
        if cur_e is not None:
            busy += cur_e - cur_s
            gaps.append((s0 - cur_e, prev_name, e.name))
        cur_s, cur_e = s0, e0
    else:
        cur_e = max(cur_e, e0)
    prev_name = e.name
busy += cur_e - cur_s
print("union of kernel intervals %.1f ms busy, %.1f ms with no kernel running (%d gaps)" % (busy / 1e3, (span - busy) / 1e3, len(gaps)))
after = {}
for g, a, b in gaps:
    k = ("attn_bwd" if "attn_step_bwd" in a else "attn_fwd" if "attn_step_fwd" in a else "dec_step" if "dec_step" in a else
         "lstm_cl" if "lstm_cl" in a else "b2asr other" if "b2asr" in a else "gemm" if ("gemm" in a or "splitK" in a) else "other")
    v = after.setdefault(k, [0, 0.0]); v[0] += 1; v[1] += g
for k, (c, t) in sorted(after.items(), key=lambda kv: -kv[1][1]):
    print("  gaps after %-12s %5d x %5.2f us = %6.2f ms" % (k, c, t / c, t / 1e3))

# the tail of the backward pass: what still runs after the last recurrence kernel until the optimizer launch
rec = [e for e in ev if "lstm_cl_bwd" in e.name]
opt_k = [e for e in ev if "adadelta" in e.name]
if rec and opt_k:
    t_end = max(e.time_range.end for e in rec)
    t_opt = opt_k[-1].time_range.start
    tail = [e for e in ev if e.time_range.end > t_end and e.time_range.start < t_opt]
    print("after the last recurrence kernel: %.2f ms until the optimizer launch, %d kernels, %.2f ms summed" % (
        (t_opt - t_end) / 1e3, len(tail), sum(min(e.time_range.end, t_opt) - max(e.time_range.start, t_end) for e in tail) / 1e3))
    agg = {}
    for e in tail:
        a = agg.setdefault(e.name[:70], [0, 0.0]); a[0] += 1; a[1] += e.device_time
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:8]:
        print("    %3d x %8.1f us  %s" % (c, t / c, k))
    # and per backward recurrence launch: when it ran and how much side-stream GEMM time overlapped it
    for e in sorted(rec, key=lambda e: e.time_range.start):
        ov = sum(max(0, min(g.time_range.end, e.time_range.end) - max(g.time_range.start, e.time_range.start))
                 for g in ev if g is not e and ("gemm" in g.name or "cutlass" in g.name or "reduce_kernel" in g.name))
        print("    recurrence bwd %7.2f ms long, GEMM/reduce time overlapping it %7.2f ms" % (
            (e.time_range.end - e.time_range.start) / 1e3, ov / 1e3))

byname = {}
for e in ev:
    a = byname.setdefault(e.name[:90], [0, 0.0]); a[0] += 1; a[1] += e.device_time
print("top kernels:")
for k, (c, t) in sorted(byname.items(), key=lambda kv: -kv[1][1])[:28]:
    print("  %6d x %7.1f us = %7.2f ms  %s" % (c, t / c, t / 1e3, k))
