"""Phase timeline of the attention step kernels (CTA 0), from clock64 stamps.  Run with B2ASR_ATT_TRACE=1."""
import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B2ASR_ATT_TRACE", "1")
PKG = "end-to-end-asr-pytorch_b200"
attn, lib = (importlib.import_module(PKG + "." + m) for m in ("_attn", "_lib"))
dev = torch.device("cuda:0")
B, T, D, A, dec = 32, 400, 640, 300, 320
h = torch.randn(B, T, D, device=dev); k = torch.tanh(torch.randn(B, T, A, device=dev))
lens = attn.lens_i32([T] * B, dev)
s = torch.randn(B, dec, device=dev)
g = torch.Generator().manual_seed(1)
W = dict(phi=torch.randn(A, dec, generator=g) * dec ** -0.5, conv=torch.randn(10, 1, 201, generator=g) * 201 ** -0.5,
         lproj=torch.randn(A, 10, generator=g) * 10 ** -0.5, gw=torch.randn(1, A, generator=g) * A ** -0.5, gb=torch.zeros(1))
W = {n: v.to(dev) for n, v in W.items()}
prev = torch.softmax(torch.randn(B, T, device=dev), -1)
for _ in range(3):
    att, ctx, q = attn.step_forward("loc", s, W["phi"], k, h, lens, prev, W["conv"], W["lproj"], W["gw"], W["gb"], 2.0, want_q=True)
dk = torch.zeros(B, T, A, device=dev)
slots = attn.num_slots(B, T, D, A, dec, "loc", 10, 100)
pg = torch.zeros(slots, attn.pgrad_layout(A, 10, 100)[-1], device=dev)
dctx, datt = torch.randn(B, D, device=dev), torch.randn(B, T, device=dev) * 0.01
for _ in range(3):
    attn.step_backward("loc", W["phi"], k, h, lens, prev, W["conv"], W["lproj"], W["gw"], 2.0, q, att, dctx, datt, dk, pg, dec)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 32)()
rc = lib.lib().b2asr_debug_attn_trace(buf, 32)
if rc != 0:
    print("trace disabled:", lib.lib().b2asr_last_error().decode())
    sys.exit(0)
mhz = torch.cuda.clock_rate() if hasattr(torch.cuda, "clock_rate") else 1900
names_f = ["", "init+issue", "stage+sync0", "query+conv", "sync1", "energy", "softmax", "context", "sync3", "dsmem reduce+sync4"]
names_b = ["init+issue", "stage+sync0+conv", "d_alpha", "softmax bwd", "key pass", "pgrad+sync2", "dq", "loc branch", "sync3", "d_s", "sync4"]
v = list(buf)
print("clock MHz", mhz)
print("forward:")
for i in range(1, 10):
    print("  %-22s %7d cycles  %6.2f us" % (names_f[i], v[i] - v[i - 1], (v[i] - v[i - 1]) / mhz))
print("  total %.2f us" % ((v[9] - v[0]) / mhz))
print("backward:")
for i in range(17, 28):
    print("  %-22s %7d cycles  %6.2f us" % (names_b[i - 17], v[i] - v[i - 1], (v[i] - v[i - 1]) / mhz))
print("  total %.2f us" % ((v[27] - v[16]) / mhz))
