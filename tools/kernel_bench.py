"""Stand-alone timing of each kernel group at the SURVEY.md section 8d shapes, with algorithmic bytes / flops
and the fraction of the measured peaks.  Inputs are larger than L2 or rotated so that every timed launch
streams from HBM unless noted.  usage: python tools/kernel_bench.py [ctc|keys|attn|ce|all]"""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
ctc, keys, attn, lib, ce = (importlib.import_module(PKG + "." + m) for m in ("ctc", "_keys", "_attn", "_lib", "ce"))
dev = torch.device("cuda:0")
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
which = sys.argv[1] if len(sys.argv) > 1 else "all"


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def bench_ctc():
    print("== kernel (1) CTC loss+grad from logits, V=5000, B=64 (workload M5); bytes = 4V(sum T_b + B*Tmax) + 8 sum L + 16B")
    g = torch.Generator().manual_seed(0)
    for T, L in [(500, 50), (500, 150), (1000, 150), (1000, 300), (2000, 300), (2000, 600), (3000, 600)]:
        B, V = 64, 5000
        x = torch.randn(B, T, V, device=dev)
        il = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); il[0] = T
        tl = torch.randint(L // 2, L + 1, (B,), generator=g); tl[0] = L
        y = torch.randint(1, V, (B, L), generator=g)
        for b in range(B):
            y[b, tl[b]:] = 0
        yd, tld, ild = y.to(dev), tl.to(dev), il.to(dev)
        xr = x.requires_grad_(True)

        def run():
            xr.grad = None
            loss = ctc.ctc_loss(xr, yd, ild, tld, from_logits=True, batch_first=True)
            loss.backward()
        lib.profile_start()
        t = timeit(run, iters=5, warm=2)
        prof = lib.profile_stop()
        kt = prof["b2asr_ctc_loss_fwd_bwd"][1] / prof["b2asr_ctc_loss_fwd_bwd"][0] * 1e-3
        nbytes = 4 * V * (int(il.sum()) + B * T) + 8 * int(tl.sum()) + 16 * B
        print("T'=%4d L=%3d: kernels %.3f ms (autograd round trip %.3f ms)  %.0f GB/s = %.2f of measured HBM peak, %.0f utt/s"
              % (T, L, kt * 1e3, t * 1e3, nbytes / kt / 1e9, nbytes / kt / 1e9 / PEAK["hbm_gbs"], B / kt))
        del x, xr


def bench_keys():
    print("== kernel (2a) key projection, M2: M=12800, D=640, A=300; flops = 2MDA (algorithmic, 1x), bytes = 4M(D+A)+4A(D+1)")
    M, D, A = 12800, 640, 300
    hs = [torch.randn(32, 400, D, device=dev) for _ in range(6)]      # rotate inputs (6 x 33 MB > L2)
    w = torch.randn(A, D, device=dev) * D ** -0.5
    b = torch.randn(A, device=dev)
    i = [0]

    def ours():
        i[0] += 1
        return keys.project(hs[i[0] % 6], w, b)

    def cublas():
        i[0] += 1
        return torch.tanh(torch.nn.functional.linear(hs[i[0] % 6], w, b))
    for name, fn, tf in (("tcgen05 3xTF32", ours, None), ("cuBLAS fp32 + tanh", cublas, False), ("cuBLAS TF32 + tanh", cublas, True)):
        if tf is not None:
            torch.backends.cuda.matmul.allow_tf32 = tf
        if tf is None:
            lib.profile_start()
        t = timeit(fn, iters=20)
        if tf is None:
            prof = lib.profile_stop()["b2asr_attn_keys"]
            print("%-20s %.1f us per launch on the device (events around the launch); the next line includes the host side" % ("", prof[1] / prof[0] * 1e3))
            t = min(t, prof[1] / prof[0] * 1e-3)
        fl, by = 2.0 * M * D * A, 4.0 * M * (D + A) + 4 * A * (D + 1)
        print("%-20s %.1f us  %.1f TFLOP/s algorithmic, %.0f GB/s = %.2f of measured HBM peak" % (name, t * 1e6, fl / t / 1e12, by / t / 1e9, by / t / 1e9 / PEAK["hbm_gbs"]))


def bench_attn():
    print("== kernel (2b) attention step, M2: B=32, T'=400, D=640, A=300, dec=320 (keys+features 48 MB: L2-resident when one batch is reused; 'rotating' uses 4 batches = 193 MB > L2)")
    B, T, D, A, dec = 32, 400, 640, 300, 320
    nb = 4
    hs = [torch.randn(B, T, D, device=dev) for _ in range(nb)]
    ks = [torch.tanh(torch.randn(B, T, A, device=dev)) for _ in range(nb)]
    lens = attn.lens_i32([T] * B, dev)
    s = torch.randn(B, dec, device=dev)
    g = torch.Generator().manual_seed(1)
    W = dict(phi=torch.randn(A, dec, generator=g) * dec ** -0.5, conv=torch.randn(10, 1, 201, generator=g) * 201 ** -0.5,
             lproj=torch.randn(A, 10, generator=g) * 10 ** -0.5, gw=torch.randn(1, A, generator=g) * A ** -0.5, gb=torch.zeros(1))
    W = {k: v.to(dev) for k, v in W.items()}
    prev = torch.softmax(torch.randn(B, T, device=dev), -1)
    by_f = 4 * B * T * (A + D + 2) + 4 * B * (dec + D)
    by_b = 4 * B * T * (3 * A + D + 4) + 4 * B * (2 * A + D + dec)
    for label, rot in (("same batch (L2-resident)", 1), ("rotating batches (HBM)", nb)):
        i = [0]

        def fwd():
            i[0] += 1
            j = i[0] % rot
            return attn.step_forward("loc", s, W["phi"], ks[j], hs[j], lens, prev, W["conv"], W["lproj"], W["gw"], W["gb"], 2.0, want_q=True)
        t = timeit(fwd, iters=20)
        print("fwd %-26s %.1f us  %.0f GB/s algorithmic = %.2f of measured HBM peak" % (label, t * 1e6, by_f / t / 1e9, by_f / t / 1e9 / PEAK["hbm_gbs"]))
        att, ctx, q = fwd()
        dk = [torch.zeros(B, T, A, device=dev) for _ in range(nb)]
        slots = attn.num_slots(B, T, D, A, dec, "loc", 10, 100)
        pg = torch.zeros(slots, attn.pgrad_layout(A, 10, 100)[-1], device=dev)
        dctx, datt = torch.randn(B, D, device=dev), torch.randn(B, T, device=dev) * 0.01

        def bwd():
            i[0] += 1
            j = i[0] % rot
            return attn.step_backward("loc", W["phi"], ks[j], hs[j], lens, prev, W["conv"], W["lproj"], W["gw"], 2.0, q, att, dctx, datt, dk[j], pg, dec)
        t = timeit(bwd, iters=20)
        print("bwd %-26s %.1f us  %.0f GB/s algorithmic = %.2f of measured HBM peak" % (label, t * 1e6, by_b / t / 1e9, by_b / t / 1e9 / PEAK["hbm_gbs"]))


def bench_ce():
    print("== sequence loss (section 8f-3): CE + gradient + argmax in one launch; bytes = 8 V N (logits read, gradient written)")
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(0)
    for B, L, V in [(32, 100, 5000), (64, 400, 5000), (32, 200, 31)]:
        nrot = max(2, int(3e8 // (8 * B * L * V)) + 1)                 # rotate inputs so launches stream from HBM
        xs = [torch.randn(B, L, V, device=dev).requires_grad_(True) for _ in range(min(nrot, 8))]
        y = torch.randint(2, V, (B, L), generator=g)
        tl = torch.randint(L // 2, L + 1, (B,), generator=g); tl[0] = L
        for b in range(B):
            y[b, tl[b]:] = 0
        yd, tld = y.to(dev), tl.to(dev)
        k = [0]

        def ours():
            x = xs[k[0] % len(xs)]; k[0] += 1
            x.grad = None
            ce.seq_loss_mean(x, yd, tld, 0, 0.5).backward()

        def stock():
            x = xs[k[0] % len(xs)]; k[0] += 1
            x.grad = None
            r = F.cross_entropy(x.reshape(B * L, V), yd.reshape(-1), ignore_index=0, reduction='none')
            (0.5 * torch.mean(r.view(B, L).sum(-1) / tld.float())).backward()
            x.detach().argmax(-1)
        lib.profile_start()
        t = timeit(ours, iters=10, warm=3)
        prof = lib.profile_stop()
        kt = prof["b2asr_ce_loss_fwd_bwd"][1] / prof["b2asr_ce_loss_fwd_bwd"][0] * 1e-3
        ts = timeit(stock, iters=10, warm=3)
        nbytes = 8 * V * B * L
        print("B=%2d L=%3d V=%4d: kernel %.1f us (with autograd %.1f us)  %.0f GB/s = %.2f of measured HBM peak; torch cross_entropy+argmax fwd+bwd %.1f us"
              % (B, L, V, kt * 1e6, t * 1e6, nbytes / kt / 1e9, nbytes / kt / 1e9 / PEAK["hbm_gbs"], ts * 1e6))
        del xs


if which in ("ce", "all"):
    bench_ce()
if which in ("ctc", "all"):
    bench_ctc()
if which in ("keys", "all"):
    bench_keys()
if which in ("attn", "all"):
    bench_attn()
