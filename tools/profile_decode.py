"""Phase timing of one M4 beam decode (encoder / eager steps / graph capture / replay / host rebuild) and,
between cudaProfilerStart/Stop, a few eager decode steps for ncu.  usage: python tools/profile_decode.py"""
import importlib
import math
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

PKG = "end-to-end-asr-pytorch_b200"
asr, rnnlm, syn, beam, lib = (importlib.import_module(PKG + "." + m) for m in ("asr", "rnnlm", "synthetic", "beam", "_lib"))
DEC = bench.DEC
w = syn.WORKLOADS[DEC["workload"]]
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).eval()
model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, DEC["ctc_weight"], DEC["lm_weight"]
model.rnn_lm = rnnlm.RNN_LM(DEC["lm"][0], DEC["lm"][1], w["vocab"], layers=DEC["lm"][2]).to(dev).eval()
x = torch.randn(1, DEC["frames"], w["d_in"], generator=torch.Generator().manual_seed(1)).to(dev)
steps = int(math.ceil(DEC["frames"] * DEC["ratio"]))


def sync():
    torch.cuda.synchronize()
    return time.perf_counter()


with torch.no_grad():
    for graph in (True, False, True):
        t0 = sync()
        hyps = model.beam_decode(x, steps, [DEC["frames"]], DEC["beam"]) if graph else \
            beam.beam_decode(model, x, steps, [DEC["frames"]], DEC["beam"], use_graph=False)
        t1 = sync()
        print("graph=%s total %.1f ms, best len %d" % (graph, (t1 - t0) * 1e3, len(hyps[0].outIndex)))
    t0 = sync()
    h, el = model.encoder(x, [DEC["frames"]])
    t1 = sync()
    print("encoder %.1f ms" % ((t1 - t0) * 1e3))
    # the search kernel alone on random logits of the M4 shape
    V, T, n = w["vocab"], h.shape[1], DEC["beam"]
    x_ctc = torch.log_softmax(torch.randn(T, V, device=dev) * 2, -1)
    att = torch.log_softmax(torch.randn(n, V, device=dev) * 2, -1)
    lm = torch.log_softmax(torch.randn(n, V, device=dev) * 2, -1)
    bs = beam.BeamSearch(V, n, 64, dev, x_ctc=x_ctc, ctc_weight=0.5, lm_weight=0.5)
    for _ in range(4):
        bs.step(att, lm)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    e0.record()
    for _ in range(20):
        bs.step(att, lm)
    e1.record()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("beam_step kernel: %.1f us per step (T'=%d, V=%d, beam %d, K=%d)" % (e0.elapsed_time(e1) / 20 * 1e3, T, V, n, bs.K))
    import ctypes
    buf = (ctypes.c_longlong * 16)()
    if lib.lib().b2asr_debug_beam_trace(buf, 16) == 0:
        v = list(buf)
        names = ["load att", "top-K", "tables", "recursion", "fusion", "top-beam", "store", "", "rank in", "rank", "state", "r gather", "eos+final"]
        print("beam_step phases (CTA 0: 0-6; pruning CTA: 8-13), us at 1965 MHz:")
        for i in range(1, 7):
            print("  %-10s %7.2f" % (names[i - 1], (v[i] - v[i - 1]) / 1965.0))
        for i in range(9, 14):
            print("  %-10s %7.2f" % (names[i - 1], (v[i] - v[i - 1]) / 1965.0))
    t0 = sync()
    res = bs.results()
    t1 = sync()
    print("host rebuild of %d hypotheses: %.1f ms" % (len(res), (t1 - t0) * 1e3))
