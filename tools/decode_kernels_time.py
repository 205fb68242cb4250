"""Stand-alone device time of the decode step's dense kernels (csrc/decode.cu) at the M4 shapes: the Speller cell
(20 rows, [640 | 320] -> 4x320), an RNN-LM cell (20 rows, 650 -> 4x650) and the two output layers (V=5000), each
captured 20x in a CUDA graph and replayed.  usage: python tools/decode_kernels_time.py"""
import ctypes
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
m = importlib.import_module("end-to-end-asr-pytorch_b200._lib")
L = m.lib()
dev = torch.device("cuda:0")
n = 20


def cell(D, H, V=5000):
    D4, H4 = (D + 3) // 4 * 4, (H + 3) // 4 * 4
    table = torch.randn(V, 4 * H, device=dev)
    W = torch.randn(4 * H, D4 + H4, device=dev) * 0.02
    x1 = torch.randn(n, D4, device=dev) if D else None
    hd, cd = torch.randn(n, H4, device=dev), torch.randn(n, H4, device=dev)
    h_out, c_out = torch.zeros(n, H4, device=dev), torch.zeros(n, H4, device=dev)
    tok, par = torch.randint(0, V, (n,), device=dev), torch.randint(0, n, (n,), device=dev)
    cs = m.CellStep()
    cs.B, cs.H, cs.K1, cs.K2 = n, H, D4, H4
    cs.x1, cs.x1_ld = m.ptr(x1), D4
    cs.x2, cs.x2_ld, cs.map = hd.data_ptr(), H4, par.data_ptr()
    cs.W, cs.w_ld = W.data_ptr(), W.stride(0)
    cs.base, cs.base_ld, cs.base_row = table.data_ptr(), table.stride(0), tok.data_ptr()
    cs.c_prev, cs.c, cs.c_ld, cs.h, cs.h_ld = cd.data_ptr(), c_out.data_ptr(), H4, h_out.data_ptr(), H4
    keep = (table, W, x1, hd, cd, h_out, c_out, tok, par)
    return lambda st: m.check(L.b2asr_cell_step_fwd(ctypes.byref(cs), st), "cell"), keep, 4 * W.numel()


def linear(K, V=5000):
    K4 = (K + 3) // 4 * 4
    x = torch.randn(n, K4, device=dev)
    W = torch.randn(V, K4, device=dev) * 0.02
    b = torch.randn(V, device=dev)
    logits, lse = torch.empty(n, V, device=dev), torch.empty(n, device=dev)
    part = torch.empty(L.b2asr_linear_lse_ctas(V) * n * 2, device=dev)
    ticket = torch.zeros(4, dtype=torch.int32, device=dev)
    a = m.LinearLse()
    a.B, a.K, a.V = n, K4, V
    a.x, a.x_ld, a.W, a.w_ld, a.bias = x.data_ptr(), K4, W.data_ptr(), K4, b.data_ptr()
    a.logits, a.out_ld, a.lse, a.part, a.ticket = logits.data_ptr(), V, lse.data_ptr(), part.data_ptr(), ticket.data_ptr()
    keep = (x, W, b, logits, lse, part, ticket)
    return lambda st: m.check(L.b2asr_linear_lse_fwd(ctypes.byref(a), st), "linear"), keep, 4 * W.numel()


for name, (fn, keep, wbytes) in (("Speller cell  [640|320] -> 4x320", cell(640, 320)), ("RNN-LM cell   650 -> 4x650", cell(0, 650)),
                                 ("char_trans    320 -> 5000", linear(320)), ("LM output     650 -> 5000", linear(650))):
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            fn(side.cuda_stream)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(20):
                fn(torch.cuda.current_stream().cuda_stream)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 100 * 1e3
    print("%-36s %6.1f us per launch   weights %.1f MB -> %.0f GB/s (L2-resident)" % (name, t, wbytes / 1e6, wbytes / t / 1e3))
