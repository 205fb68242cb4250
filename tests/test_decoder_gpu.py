"""Decoder-loop fusion (end-to-end-asr-pytorch_b200/_dec.py + csrc/lstm.cu) against the
reference-shaped loop (torch.nn.LSTMCell on cat([last_char, context]), src/asr.py:84-104) on the same weights."""
import importlib
import random

import pytest
import torch

PKG = "end-to-end-asr-pytorch_b200"
pytestmark = pytest.mark.gpu


def _cfg(att_mode):
    return {
        'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2', 'sample_style': 'drop', 'dim': '24_24', 'dropout': '0_0', 'rnn_cell': 'LSTM'},
        'attention': {'att_mode': att_mode, 'dim': 32, 'proj': True, 'num_head': 1},
        'decoder': {'dim': 40, 'layer': 1, 'dropout': 0, 'rnn_cell': 'LSTMCell'},
        'optimizer': {'joint_ctc': 0.5},
    }


def _run(model, x, y, lens, tf_rate, fused, seed):
    dec = importlib.import_module(PKG + "._dec")
    keep = dec.supported
    if not fused:
        dec.supported = lambda d: False
    try:
        random.seed(seed); torch.manual_seed(seed)
        model.zero_grad(set_to_none=True)
        ctc_out, enc_len, att_out, maps = model(x, y.shape[1] - 1, tf_rate=tf_rate, teacher=y, state_len=lens)
        loss = (att_out ** 2).mean() + 0.1 * (ctc_out ** 2).mean() + att_out[:, :, 3].sum() * 1e-2
        loss.backward()
        grads = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
        return att_out.detach(), maps[0].detach(), grads, random.random()
    finally:
        dec.supported = keep


@pytest.mark.parametrize("att_mode", ["loc", "dot"])
def test_fused_loop_matches_cell_loop_teacher_forced(cuda, att_mode):
    asr = importlib.import_module(PKG + ".asr")
    torch.manual_seed(0)
    B, T, D_in, V, L = 5, 64, 12, 17, 9
    model = asr.Seq2Seq(torch.zeros(1, 8, D_in), V, _cfg(att_mode)).to(cuda).train()
    x = torch.randn(B, T, D_in, device=cuda)
    lens = [64, 60, 52, 40, 33]
    for b, n in enumerate(lens):
        x[b, n:] = 0
    y = torch.randint(2, V, (B, L + 1), device=cuda); y[:, 0] = 0
    a0, m0, g0, r0 = _run(model, x, y, lens, 1.0, False, 7)
    a1, m1, g1, r1 = _run(model, x, y, lens, 1.0, True, 7)
    assert r0 == r1                                             # same number of coins drawn
    assert torch.allclose(a0, a1, rtol=1e-4, atol=1e-5)
    assert torch.allclose(m0, m1, rtol=1e-4, atol=1e-6)
    assert set(g0) == set(g1)
    for n in g0:
        tol = 1e-4 * float(g0[n].abs().max()) + 1e-7
        assert float((g0[n] - g1[n]).abs().max()) <= tol, n


def test_fused_loop_sampling_and_greedy_paths(cuda):
    asr = importlib.import_module(PKG + ".asr")
    torch.manual_seed(1)
    B, T, D_in, V, L = 4, 48, 12, 11, 7
    model = asr.Seq2Seq(torch.zeros(1, 8, D_in), V, _cfg("loc")).to(cuda).train()
    x = torch.randn(B, T, D_in, device=cuda)
    lens = [48, 44, 40, 30]
    y = torch.randint(2, V, (B, L + 1), device=cuda); y[:, 0] = 0
    # scheduled sampling: coins and torch RNG are consumed exactly as in the reference-shaped loop, so with the
    # same seeds both loops sample from (numerically almost) the same distributions; compare the first step, which
    # cannot have diverged yet, and require finite gradients everywhere
    a0, _, g0, r0 = _run(model, x, y, lens, 0.5, False, 3)
    a1, _, g1, r1 = _run(model, x, y, lens, 0.5, True, 3)
    assert r0 == r1
    assert torch.allclose(a0[:, 0], a1[:, 0], rtol=1e-4, atol=1e-5)
    assert all(torch.isfinite(v).all() for v in g1.values()) and set(g0) == set(g1)
    # greedy (no teacher), no grad: identical argmax chains
    model.eval()
    with torch.no_grad():
        dec = importlib.import_module(PKG + "._dec")
        _, _, o1, _ = model(x, L, state_len=lens)
        keep = dec.supported
        dec.supported = lambda d: False
        try:
            _, _, o0, _ = model(x, L, state_len=lens)
        finally:
            dec.supported = keep
    assert torch.allclose(o0, o1, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("att_mode,B,T,L", [("loc", 5, 64, 9), ("dot", 3, 40, 1), ("loc", 8, 200, 33)])
def test_native_loop_matches_per_step_path(cuda, monkeypatch, att_mode, B, T, L):
    """csrc/decloop.cu (one native call per direction) against the per-step autograd path (_dec.py) that it
    replaces for teacher-forced batches: same kernels, same cuBLAS calls -> agreement far inside the 1e-4 bar."""
    asr = importlib.import_module(PKG + ".asr")
    lib = importlib.import_module(PKG + "._lib")
    torch.manual_seed(2)
    D_in, V = 12, 17
    model = asr.Seq2Seq(torch.zeros(1, 8, D_in), V, _cfg(att_mode)).to(cuda).train()
    x = torch.randn(B, T, D_in, device=cuda)
    lens = sorted([int(v) for v in torch.randint(T // 2, T + 1, (B,))], reverse=True)
    lens[0] = T
    for b, n in enumerate(lens):
        x[b, n:] = 0
    y = torch.randint(2, V, (B, L + 1), device=cuda); y[:, 0] = 0
    monkeypatch.setenv("B2ASR_DECLOOP", "0")
    a0, m0, g0, r0 = _run(model, x, y, lens, 1.0, True, 7)
    monkeypatch.setenv("B2ASR_DECLOOP", "1")
    lib.profile_start()
    a1, m1, g1, r1 = _run(model, x, y, lens, 1.0, True, 7)
    prof = lib.profile_stop()
    assert "b2asr_decoder_loop_fwd" in prof and "b2asr_decoder_loop_bwd" in prof       # the native loop really ran
    assert r0 == r1
    assert torch.allclose(a0, a1, rtol=1e-5, atol=1e-6)
    assert torch.allclose(m0, m1, rtol=1e-5, atol=1e-7)
    assert set(g0) == set(g1)
    for n in g0:
        tol = 1e-4 * float(g0[n].abs().max()) + 1e-7        # summation orders differ (d_h paths, dK accumulation)
        assert float((g0[n] - g1[n]).abs().max()) <= tol, n
    # module state after the loop is what the per-step loop leaves behind
    assert model.decoder.state_list[0].shape == (B, 40)
    if att_mode == "loc":
        assert model.attention.prev_att.shape == (B, 1, T // 4)


@pytest.mark.parametrize("att_mode", ["loc", "dot"])
def test_native_loop_with_scheduled_sampling_matches_per_step_path(cuda, monkeypatch, att_mode):
    """Scheduled sampling (reference asr.py:96-101) through the native loop run range by range (csrc/decloop.cu
    b2asr_decoder_loop_fwd_range) against the per-step path.  The output layer is made so peaked that the Categorical
    draw is the argmax in both paths (the draws themselves consume the same RNG stream), so tokens agree and outputs
    and EVERY gradient -- including the embedding table's and the embedding half of weight_ih through the sampled
    rows -- must agree."""
    asr = importlib.import_module(PKG + ".asr")
    lib = importlib.import_module(PKG + "._lib")
    torch.manual_seed(5)
    B, T, D_in, V, L = 6, 64, 12, 13, 12
    model = asr.Seq2Seq(torch.zeros(1, 8, D_in), V, _cfg(att_mode)).to(cuda).train()
    with torch.no_grad():
        model.char_trans.weight.mul_(60.0)
    x = torch.randn(B, T, D_in, device=cuda)
    lens = [64, 61, 55, 50, 41, 33]
    for b, n in enumerate(lens):
        x[b, n:] = 0
    y = torch.randint(2, V, (B, L + 1), device=cuda); y[:, 0] = 0
    monkeypatch.setenv("B2ASR_DECLOOP_SAMPLING", "0")
    a0, m0, g0, r0 = _run(model, x, y, lens, 0.5, True, 11)
    monkeypatch.setenv("B2ASR_DECLOOP_SAMPLING", "1")
    lib.profile_start()
    a1, m1, g1, r1 = _run(model, x, y, lens, 0.5, True, 11)
    prof = lib.profile_stop()
    assert "b2asr_decoder_loop_fwd_range" in prof and prof["b2asr_decoder_loop_fwd_range"][0] >= 2   # really segmented
    assert r0 == r1
    assert torch.allclose(a0, a1, rtol=1e-4, atol=1e-4 * float(a0.abs().max()))
    assert torch.allclose(m0, m1, rtol=1e-4, atol=1e-6)
    assert set(g0) == set(g1)
    for n in g0:
        # (gen_energy.bias shifts every energy alike, its true gradient is 0: both paths return rounding noise there)
        tol = 1e-4 * float(g0[n].abs().max()) + 1e-6
        assert float((g0[n] - g1[n]).abs().max()) <= tol, n
    assert float(g1["embed.weight"].abs().max()) > 0
