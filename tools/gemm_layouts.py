"""cuBLAS fp32 GEMM rates for the Listener's dense products at M2 shapes, per operand layout (which form to feed)."""
import torch
dev = torch.device("cuda:0")
def bench(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
N = 2560
for BT, K in ((51200, 120), (25600, 640), (12800, 640)):
    x = torch.randn(BT, K, device=dev); w = torch.randn(N, K, device=dev); wt = w.t().contiguous()
    dG = torch.randn(BT, N, device=dev); dGt = dG.t().contiguous(); xt = x.t().contiguous()
    fl = 2.0 * BT * K * N
    out = torch.empty(BT, N, device=dev); dx = torch.empty(BT, K, device=dev); dw = torch.empty(N, K, device=dev)
    res = {
        "fwd  x @ w.T            ": bench(lambda: torch.mm(x, w.t(), out=out)),
        "fwd  x @ wt (nn)        ": bench(lambda: torch.mm(x, wt, out=out)),
        "dx   dG @ w (nn)        ": bench(lambda: torch.mm(dG, w, out=dx)),
        "dx   dG @ wt.T (nt)     ": bench(lambda: torch.mm(dG, wt.t(), out=dx)),
        "dw   dG.T @ x (tn)      ": bench(lambda: torch.mm(dG.t(), x, out=dw)),
        "dw   dGt @ x (nn)       ": bench(lambda: torch.mm(dGt, x, out=dw)),
        "dwT  x.T @ dG -> [K,N]  ": bench(lambda: torch.mm(x.t(), dG)),
    }
    for k, ms in res.items():
        print("BT=%6d K=%4d  %s %7.3f ms  %5.1f TFLOP/s" % (BT, K, k, ms, fl / ms / 1e9), flush=True)
