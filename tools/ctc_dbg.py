import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
lib = importlib.import_module("end-to-end-asr-pytorch_b200._lib")
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_ctc_gpu import make_case
L = lib.lib()
dev = torch.device("cuda:0")
B, T, V, Lmax = 4, 200, 5000, 40
x, y, il, tl = make_case(B, T, V, Lmax, seed=B * 1000 + T)
xd, yd, ild, tld = x.to(dev), y.to(dev), il.to(dev), tl.to(dev)
res = {}
for mode in ("0", "1", "1"):
    os.environ["B2ASR_CTC_OVERLAP"] = mode
    n = L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, 1)
    ws = torch.full((n,), 0x55, dtype=torch.uint8, device=dev)
    nll = torch.empty(B, device=dev); loss = torch.empty(1, device=dev); grad = torch.empty_like(xd)
    rc = L.b2asr_ctc_loss_fwd_bwd(xd.data_ptr(), xd.stride(1), xd.stride(0), yd.data_ptr(), yd.stride(0), ild.data_ptr(), tld.data_ptr(),
                                  T, B, V, Lmax, 0, 1, 1, 1.0, nll.data_ptr(), loss.data_ptr(), grad.data_ptr(), grad.stride(1), grad.stride(0),
                                  ws.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    hdr = ws[:256 + 256 + 4 * B * 13].view(torch.int32).cpu()
    print("mode", mode, "rc", rc, "nll", nll.cpu().tolist(), "ticket/err", hdr[:2].tolist(), "prog", hdr[64:64 + 2 * B].tolist())
    print("   rdy", hdr[128:128 + 13 * B].tolist(), "in_len", il.tolist())
    res[mode] = (nll.clone(), grad.clone())
