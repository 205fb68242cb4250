"""One M4 beam decode under torch.profiler: summed kernel time vs wall clock, top kernels."""
import importlib, math, os, sys, time
import torch
from torch.profiler import profile, ProfilerActivity
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
PKG = "end-to-end-asr-pytorch_b200"
asr, rnnlm, syn, beam = (importlib.import_module(PKG + "." + m) for m in ("asr", "rnnlm", "synthetic", "beam"))
DEC = bench.DEC
w = syn.WORKLOADS[DEC["workload"]]
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).eval()
model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, DEC["ctc_weight"], DEC["lm_weight"]
model.rnn_lm = rnnlm.RNN_LM(DEC["lm"][0], DEC["lm"][1], w["vocab"], layers=DEC["lm"][2]).to(dev).eval()
x = torch.randn(1, DEC["frames"], w["d_in"], generator=torch.Generator().manual_seed(1)).to(dev)
steps = int(math.ceil(DEC["frames"] * DEC["ratio"]))
with torch.no_grad():
    for _ in range(2):
        model.beam_decode(x, steps, [DEC["frames"]], DEC["beam"])
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3):
        model.beam_decode(x, steps, [DEC["frames"]], DEC["beam"])
    torch.cuda.synchronize()
    print("plain decode %.1f ms per utterance (%d steps)" % ((time.perf_counter() - t0) / 3 * 1e3, steps))
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        model.beam_decode(x, steps, [DEC["frames"]], DEC["beam"])
        torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
tot = sum(e.device_time for e in ev)
ev.sort(key=lambda e: e.time_range.start)
span = ev[-1].time_range.end - ev[0].time_range.start
print("%d GPU kernels/memcpys, summed %.1f ms, span %.1f ms" % (len(ev), tot / 1e3, span / 1e3))
byname = {}
for e in ev:
    a = byname.setdefault(e.name[:100], [0, 0.0]); a[0] += 1; a[1] += e.device_time
for k, (c, t) in sorted(byname.items(), key=lambda kv: -kv[1][1])[:25]:
    print("  %6d x %7.1f us = %7.2f ms  %s" % (c, t / c, t / 1e3, k))
# timeline of one decode step in the middle of the utterance: every kernel between two beam steps
bsteps = [i for i, e in enumerate(ev) if "beam_step_kernel" in e.name]
if len(bsteps) > 82:
    i0, i1 = bsteps[80], bsteps[81]
    t0 = ev[i0].time_range.end
    print("one step (after beam step #80): start/end in us relative to the previous beam step's end")
    for e in ev[i0 + 1:i1 + 1]:
        print("  %7.1f .. %7.1f  (%6.1f us)  %s" % (e.time_range.start - t0, e.time_range.end - t0, e.device_time, e.name[:60]))
