"""Persistent Listener recurrence (csrc/enc.cu, _enc.py) against the CPU oracle = the reference's own call sequence
pack_padded_sequence -> nn.LSTM -> pad_packed_sequence (src/asr.py:466-501) evaluated in float64.

Tolerance: 1e-4 relative (the bar BASELINE.json states for the fp32 kernels) with an absolute floor of 1e-4 times the
tensor's largest magnitude -- most padded/late-frame entries are exactly or nearly zero."""
import importlib

import pytest
import torch

from oracle import encoder_oracle as eo

PKG = "end-to-end-asr-pytorch_b200"
pytestmark = pytest.mark.gpu


def _close(a, b, what, rtol=1e-4):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    floor = rtol * float(b.abs().max()) + 1e-12
    err = (a - b).abs()
    bad = err > rtol * torch.maximum(a.abs(), b.abs()) + floor
    assert not bool(bad.any()), "%s: max abs err %.3e (floor %.3e), %d bad" % (what, float(err.max()), floor, int(bad.sum()))


def _lens(B, T, seed, sort=True):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(max(1, T // 3), T + 1, (B,), generator=g).tolist()
    lens[0] = T
    return sorted(lens, reverse=True) if sort else lens


CASES = [
    # B, T, n_in, H, bidir
    (3, 9, 8, 8, True),            # tiny, U=4 path with 2 CTAs per direction
    (5, 37, 24, 20, True),         # H not a multiple of the CTA's unit count or of 32 (padded k-slices)
    (16, 60, 39, 256, True),       # M1 layer 0 (timit_example.yaml: 39-dim MFCC, 256 units)
    (13, 45, 30, 320, False),      # ragged row group (13 = 8 + 5 rows), one direction
    (32, 120, 120, 320, True),     # M2 layer 0 shape, shortened in time
    (32, 50, 640, 320, True),      # M2 layers 1-4
    (7, 33, 16, 64, False),        # unidirectional ('RNN' enc_type)
    (40, 21, 12, 320, True),       # more rows than one cooperative launch holds -> two launches
    (4, 1, 8, 16, True),           # single frame
    (9, 1, 16, 320, True),         # single frame on the cluster kernels (no exchange at all)
    (12, 2, 16, 256, True),        # two frames: one exchange, buffer 1 only
    (20, 3, 24, 320, True),        # three frames: both staging buffers, 12 rows per cluster
    (6, 25, 40, 512, True),        # U=8 path
]


@pytest.fixture(params=["auto", "l2"])
def path(request):
    """auto = cluster kernels (enc_cluster.cu) for H in {256, 320}, L2-exchange kernels (enc.cu) otherwise;
    l2 = the L2-exchange kernels for every shape."""
    lib = importlib.import_module(PKG + "._lib").lib()
    prev = lib.b2asr_lstm_seq_path(1 if request.param == "l2" else 0)
    yield request.param
    lib.b2asr_lstm_seq_path(prev)


@pytest.mark.parametrize("B,T,n_in,H,bidir", CASES)
def test_lstm_layer_forward_backward_match_reference_call(B, T, n_in, H, bidir, path):
    if path == "l2" and H not in (256, 320):
        pytest.skip("already the L2-exchange path under auto")
    enc = importlib.import_module(PKG + "._enc")
    dev = torch.device("cuda:0")
    layer64 = eo.make_layer(n_in, H, bidir, seed=B * 1000 + T)
    lens = _lens(B, T, seed=7 + B)
    g = torch.Generator().manual_seed(3)
    x64 = torch.randn(B, T, n_in, generator=g, dtype=torch.float64)
    for b, n in enumerate(lens):
        x64[b, n:] = 0
    w64 = torch.randn(B, T, (2 if bidir else 1) * H, generator=g, dtype=torch.float64)
    x64.requires_grad_(True)
    ref = eo.reference_call(layer64, x64, lens)
    (ref * w64).sum().backward()

    layer = torch.nn.LSTM(n_in, H, bidirectional=bidir, num_layers=1, batch_first=True).to(dev)
    layer.load_state_dict({k: v.float() for k, v in layer64.state_dict().items()})
    x = x64.detach().float().to(dev).requires_grad_(True)
    assert enc.supported(layer, x)
    out = enc.lstm_layer(layer, x, lens)
    (out * w64.float().to(dev)).sum().backward()
    torch.cuda.synchronize()
    enc.check_status(dev)
    _close(out, ref, "out")
    for b, n in enumerate(lens):
        assert float(out[b, n:].abs().max() if n < T else 0.0) == 0.0       # pad_packed_sequence zeros
    _close(x.grad, x64.grad, "dx")
    for (name, p), (_, p64) in zip(layer.named_parameters(), layer64.named_parameters()):
        _close(p.grad, p64.grad, "d" + name)


def test_unsorted_lengths_and_listener_routes_through_kernel():
    """Rows need not be sorted for the kernel (the packed reference needs them sorted: compare per permutation)."""
    enc = importlib.import_module(PKG + "._enc")
    asr = importlib.import_module(PKG + ".asr")
    lib = importlib.import_module(PKG + "._lib")
    dev = torch.device("cuda:0")
    B, T, n_in, H = 6, 30, 12, 32
    layer64 = eo.make_layer(n_in, H, True, seed=11)
    lens = [30, 12, 25, 30, 7, 19]
    g = torch.Generator().manual_seed(5)
    x64 = torch.randn(B, T, n_in, generator=g, dtype=torch.float64)
    for b, n in enumerate(lens):
        x64[b, n:] = 0
    order = sorted(range(B), key=lambda i: -lens[i])
    ref_sorted = eo.reference_call(layer64, x64[order], [lens[i] for i in order])
    ref = torch.empty_like(ref_sorted)
    ref[order] = ref_sorted
    layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
    layer.load_state_dict({k: v.float() for k, v in layer64.state_dict().items()})
    out = enc.lstm_layer(layer, x64.float().to(dev), lens)
    torch.cuda.synchronize()
    enc.check_status(dev)
    _close(out, ref, "out (unsorted)")

    # RNNLayer.forward (asr.py:466-501) takes the kernel path on CUDA and keeps the reference's return convention
    rl = asr.RNNLayer(n_in, H, 2, sample_style="drop").to(dev)
    rl.layer.load_state_dict(layer.state_dict())
    before = lib.LAUNCHES
    y, hidden, new_len = rl(x64[order].float().to(dev), state_len=[lens[i] for i in order], pack_input=True)
    assert lib.LAUNCHES == before + 1 and hidden is None
    assert new_len == [int(lens[i] / 2) for i in order]
    _close(y, ref_sorted[:, ::2], "RNNLayer drop output")


def test_weight_gradients_on_side_stream_match_stream_order():
    """_enc.OVERLAP: a layer's weight-gradient GEMMs run on a side stream next to the next recurrence kernel and are
    accumulated into .grad directly; same numbers as the stream-ordered autograd path, also when .grad already holds
    a value (accumulation over two backward passes) and when two layers are stacked (the second flushes the first)."""
    enc = importlib.import_module(PKG + "._enc")
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    B, T, n_in, H = 12, 40, 24, 320
    lens = _lens(B, T, seed=5)
    x = torch.randn(B, T, n_in, device=dev)
    l0 = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
    l1 = torch.nn.LSTM(2 * H, H, bidirectional=True, batch_first=True).to(dev)
    w = torch.randn(B, T, 2 * H, device=dev)

    import contextlib

    def run(overlap):
        for p in list(l0.parameters()) + list(l1.parameters()):
            p.grad = None
        xin = x.clone().requires_grad_(True)
        for _ in range(2):                                   # second pass accumulates into existing .grad
            y = enc.lstm_layer(l1, torch.tanh(enc.lstm_layer(l0, xin, lens)), lens)
            # the side-stream path is taken only inside deferred_weight_grads() (solver.TrainStep enters it)
            with (enc.deferred_weight_grads() if overlap else contextlib.nullcontext()):
                (y * w).sum().backward()
        torch.cuda.synchronize()
        enc.check_status(dev)
        return [p.grad.clone() for p in list(l0.parameters()) + list(l1.parameters())] + [xin.grad.clone()]

    a, b = run(True), run(False)
    for ga, gb in zip(a, b):
        _close(ga, gb, "overlap vs stream order", rtol=1e-6)


def test_weight_gradients_reach_autograd_outside_the_train_step():
    """torch.autograd.grad / backward(inputs=...) must receive the LSTM weight gradients (the side-stream
    accumulation into .grad is armed by solver.TrainStep only), and a failed backward inside the armed region must
    not leave closures behind for the next one."""
    enc = importlib.import_module(PKG + "._enc")
    dev = torch.device("cuda:0")
    torch.manual_seed(4)
    B, T, n_in, H = 6, 20, 16, 320
    lens = _lens(B, T, seed=2)
    layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
    x = torch.randn(B, T, n_in, device=dev)
    out = enc.lstm_layer(layer, x, lens)
    grads = torch.autograd.grad(out.square().sum(), list(layer.parameters()))
    assert all(g is not None and torch.isfinite(g).all() and float(g.abs().sum()) > 0 for g in grads)
    assert all(p.grad is None for p in layer.parameters())          # nothing was accumulated behind autograd's back
    with pytest.raises(RuntimeError):
        with enc.deferred_weight_grads():
            enc._deferred.append(lambda side: None)
            raise RuntimeError("backward failed")
    assert not enc._deferred and not enc._armed[0]


def test_recurrence_kernels_are_bit_reproducible():
    """The exchanges land in any order, but every sum is formed in a fixed order (the owner adds the 16 partial dh in
    rank order): two runs on the same inputs give identical bits, forward and backward."""
    enc = importlib.import_module(PKG + "._enc")
    dev = torch.device("cuda:0")
    torch.manual_seed(11)
    B, T, n_in, H = 32, 64, 40, 320
    lens = _lens(B, T, seed=3)
    layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
    x = torch.randn(B, T, n_in, device=dev)
    w = torch.randn(B, T, 2 * H, device=dev)
    prev = enc.OVERLAP
    enc.OVERLAP = False                                          # gradients through autograd: one stream, one order
    try:
        runs = []
        for _ in range(3):
            xin = x.clone().requires_grad_(True)
            for p in layer.parameters():
                p.grad = None
            out = enc.lstm_layer(layer, xin, lens)
            (out * w).sum().backward()
            torch.cuda.synchronize()
            enc.check_status(dev)
            runs.append([out.detach().clone(), xin.grad.clone()] + [p.grad.clone() for p in layer.parameters()])
    finally:
        enc.OVERLAP = prev
    for other in runs[1:]:
        for a, b in zip(runs[0], other):
            assert torch.equal(a, b)


def test_tensor_core_recurrence_long_sequence_and_row_scales():
    """The cluster kernels multiply on mma.sync with fp16 hi/lo operands (enc_cluster.cu): 400 frames of error
    accumulation, and upstream gradients whose magnitude differs by up to 1e8 between batch rows and by 1e3 between
    frames of a row -- every ROW of dx must meet 1e-4 of ITS OWN largest entry (the backward kernel scales each row by
    its own power of two before the fp16 split; one scale per frame would flush the small rows)."""
    enc = importlib.import_module(PKG + "._enc")
    dev = torch.device("cuda:0")
    B, T, n_in, H = 12, 400, 32, 320
    layer64 = eo.make_layer(n_in, H, True, seed=11)
    lens = _lens(B, T, seed=5)
    g = torch.Generator().manual_seed(9)
    x64 = torch.randn(B, T, n_in, generator=g, dtype=torch.float64)
    for b, n in enumerate(lens):
        x64[b, n:] = 0
    row_scale = 10.0 ** torch.linspace(-8, 0, B, dtype=torch.float64)
    frame_scale = 10.0 ** (-3 * torch.rand(B, T, generator=g, dtype=torch.float64))
    w64 = torch.randn(B, T, 2 * H, generator=g, dtype=torch.float64) * (row_scale[:, None] * frame_scale)[:, :, None]
    x64.requires_grad_(True)
    ref = eo.reference_call(layer64, x64, lens)
    (ref * w64).sum().backward()

    layer = torch.nn.LSTM(n_in, H, bidirectional=True, num_layers=1, batch_first=True).to(dev)
    layer.load_state_dict({k: v.float() for k, v in layer64.state_dict().items()})
    x = x64.detach().float().to(dev).requires_grad_(True)
    out = enc.lstm_layer(layer, x, lens)
    (out * w64.float().to(dev)).sum().backward()
    torch.cuda.synchronize()
    enc.check_status(dev)
    _close(out, ref, "out")
    for b in range(B):
        _close(x.grad[b], x64.grad[b], "dx row %d (upstream scale %.0e)" % (b, float(row_scale[b])))
    for (name, p), (_, p64) in zip(layer.named_parameters(), layer64.named_parameters()):
        _close(p.grad, p64.grad, "d" + name)


def test_ffma_recurrence_kernels_still_match():
    """B2ASR_ENC_MMA=0 selects the packed-FFMA cluster kernels (read once per process, hence the child process)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B2ASR_ENC_MMA="0")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", os.path.abspath(__file__), "-k",
                        "match_reference_call and auto and (32-50-640-320 or 16-60-39-256 or 20-3-24-320)"],
                       cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "3 passed" in r.stdout, r.stdout[-500:]
