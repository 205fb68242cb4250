"""GPU parity: kernel (2a) key projection K = tanh(psi(h)) (reference src/asr.py:419) on the tcgen05 tensor
cores with error-compensated TF32, vs the oracle (oracle/attention_oracle.project_keys) in float64 and vs
the keys the unmodified reference cached (tests/golden/attention_*.pt).

Tolerance (north_star: attention outputs within 1e-4 relative in fp32): |a-b| <= 1e-4*|b| + 1e-5*max|keys|
(keys lie in (-1,1); the floor covers entries near 0, same convention as the context vector in
test_attention_gpu.py).  The test also prints how far a true-fp32 library GEMM and a single-pass TF32
GEMM are from float64 on the same data, to show the compensation is what meets the bar."""
import importlib
import os

import pytest
import torch

from oracle import attention_oracle as ao

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _keys():
    return importlib.import_module("end-to-end-asr-pytorch_b200._keys")


def worst(a, b, rtol=1e-4, atol=1e-5):
    a, b = a.double().cpu(), b.double().cpu()
    return float(((a - b).abs() / (rtol * b.abs() + atol)).max())


@pytest.mark.parametrize("B,T,D,A", [(32, 400, 640, 300), (16, 97, 512, 256), (3, 37, 24, 12), (1, 750, 640, 300),
                                     (5, 129, 100, 40), (2, 50, 64, 176)])
def test_keys_match_float64_oracle(cuda, B, T, D, A):
    g = torch.Generator().manual_seed(B * T + A)
    h = torch.randn(B, T, D, generator=g)
    w = torch.randn(A, D, generator=g) * D ** -0.5
    b = torch.randn(A, generator=g) * 0.1
    ref = ao.project_keys(h.double(), w.double(), b.double())
    got = _keys().project(h.to(cuda), w.to(cuda), b.to(cuda))
    torch.cuda.synchronize()
    assert got.shape == ref.shape
    hd, wd, bd = h.to(cuda), w.to(cuda), b.to(cuda)
    flag = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    fp32 = torch.tanh(torch.nn.functional.linear(hd, wd, bd))
    torch.backends.cuda.matmul.allow_tf32 = True
    tf32 = torch.tanh(torch.nn.functional.linear(hd, wd, bd))
    torch.backends.cuda.matmul.allow_tf32 = flag
    err = lambda t: float((t.double().cpu() - ref).abs().max())
    print("max abs err vs float64: 3xTF32 tcgen05 %.2e | cuBLAS fp32 %.2e | cuBLAS TF32 %.2e" % (err(got), err(fp32), err(tf32)))
    assert worst(got, ref) <= 1.0, worst(got, ref)


@pytest.mark.parametrize("name", ["attention_loc.pt", "attention_dot.pt"])
def test_keys_match_reference_golden(cuda, name):
    g = torch.load(os.path.join(G, name), weights_only=False)
    sd = g["state_dict"]
    got = _keys().project(g["h"].to(cuda), sd["psi.weight"].to(cuda), sd["psi.bias"].to(cuda))
    assert worst(got, g["keys"]) <= 1.0


def test_keys_autograd_matches_torch(cuda):
    g = torch.Generator().manual_seed(3)
    h = torch.randn(4, 60, 64, generator=g).to(cuda).requires_grad_(True)
    w = (torch.randn(48, 64, generator=g) * 0.125).to(cuda).requires_grad_(True)
    b = (torch.randn(48, generator=g) * 0.1).to(cuda).requires_grad_(True)
    up = torch.randn(4, 60, 48, generator=g).to(cuda)
    k = _keys().project(h, w, b)
    (k * up).sum().backward()
    got = [t.grad.clone() for t in (h, w, b)]
    h2, w2, b2 = (t.detach().double().requires_grad_(True) for t in (h, w, b))
    (torch.tanh(torch.nn.functional.linear(h2, w2, b2)) * up.double()).sum().backward()
    for a, r in zip(got, (h2.grad, w2.grad, b2.grad)):
        assert torch.allclose(a.double(), r, rtol=2e-3, atol=2e-3 * float(r.abs().max()))   # library TF32 GEMMs in backward
