"""One-off: timeline of the overlapped CTC call from %globaltimer stamps (build with B2ASR_NVCC_EXTRA=-DB2ASR_CTC_TRACE)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
os.environ["B2ASR_CTC_DEBUG"] = "1"
lib = importlib.import_module("end-to-end-asr-pytorch_b200._lib")
L = lib.lib()
dev = torch.device("cuda:0")
B, V = 64, 5000
for T, Lmax in [(1000, 150), (2000, 300), (3000, 600)]:
    g = torch.Generator().manual_seed(T)
    x = torch.randn(B, T, V, device=dev)
    il = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); il[0] = T
    tl = torch.randint(Lmax // 2, Lmax + 1, (B,), generator=g); tl[0] = Lmax
    y = torch.randint(1, V, (B, Lmax), generator=g)
    for b in range(B):
        y[b, tl[b]:] = 0
    yd, ild, tld = y.to(dev), il.to(dev), tl.to(dev)
    os.environ["B2ASR_CTC_OVERLAP"] = "3"
    n = L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, 1)
    ws = torch.empty(n, dtype=torch.uint8, device=dev)
    nll = torch.empty(B, device=dev); loss = torch.empty(1, device=dev); grad = torch.empty_like(x)
    for mode in ("2", "2", "0"):
        os.environ["B2ASR_CTC_OVERLAP"] = mode
        rc = L.b2asr_ctc_loss_fwd_bwd(x.data_ptr(), x.stride(1), x.stride(0), yd.data_ptr(), yd.stride(0), ild.data_ptr(), tld.data_ptr(),
                                      T, B, V, Lmax, 0, 1, 1, 1.0, nll.data_ptr(), loss.data_ptr(), grad.data_ptr(), grad.stride(1), grad.stride(0),
                                      ws.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        w = ws[64:64 + 11 * 8].view(torch.int64).cpu().tolist()
        t0 = w[0]
        names = ["rows first start", "rows last end", "walk first start", "walk first half", "walk last half", "walk first end",
                 "walk last end", "occ first start", "occ last end"]
        print("T'=%d L=%d overlap=%s rc=%d err=%d" % (T, Lmax, mode, rc, int(ws[4:8].view(torch.int32)[0])))
        for nm, v in zip(names, w):
            print("   %-18s %9.1f us" % (nm, (v - t0) / 1e3))
