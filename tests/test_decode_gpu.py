"""GPU parity of the decode step's neural kernels (csrc/decode.cu) through the C ABI against float64 evaluations of
the reference's operations: torch.nn.LSTMCell on cat([embed(last), context]) (reference src/asr.py:352-357, 208),
nn.LSTM one-token step (src/rnnlm.py:19-25), Linear + log_softmax (asr.py:210, 235).
Tolerance: 1e-4 relative (north_star) with an absolute floor of 1e-5 on O(1) activations / logits."""
import ctypes
import importlib

import pytest
import torch

PKG = "end-to-end-asr-pytorch_b200"
pytestmark = pytest.mark.gpu


def _lib():
    return importlib.import_module(PKG + "._lib")


def _pad(w, cols):
    out = torch.zeros(w.shape[0], cols, dtype=w.dtype, device=w.device)
    out[:, :w.shape[1]] = w
    return out


@pytest.mark.parametrize("n,V,E,D,H", [(20, 5000, 320, 640, 320), (20, 5000, 650, 0, 650), (7, 37, 10, 24, 18), (33, 50, 8, 12, 6)])
def test_cell_step_matches_lstm_cell_float64(cuda, n, V, E, D, H):
    """Table row (embedding half + biases) + [x1 | x2[map]] W^T + cell, rows through a parent map, H not a
    multiple of 4 (650 = the shipped RNN-LM), more rows than one warp pass (33)."""
    m = _lib()
    L = m.lib()
    g = torch.Generator().manual_seed(n * 100 + H)
    emb = torch.randn(V, E, generator=g, dtype=torch.float64) * 0.5
    cell = torch.nn.LSTMCell(E + D, H).double()
    x1 = torch.randn(n, D, generator=g, dtype=torch.float64)
    h_prev = torch.randn(n, H, generator=g, dtype=torch.float64) * 0.5
    c_prev = torch.randn(n, H, generator=g, dtype=torch.float64) * 0.5
    tok = torch.randint(0, V, (n,), generator=g)
    par = torch.randint(0, n, (n,), generator=g)
    with torch.no_grad():
        h_ref, c_ref = cell(torch.cat([emb[tok], x1], -1), (h_prev[par], c_prev[par]))
    D4, H4 = (D + 3) // 4 * 4, (H + 3) // 4 * 4
    f = lambda t: t.float().to(cuda).contiguous()
    table = f(emb @ cell.weight_ih.detach()[:, :E].t() + cell.bias_ih.detach() + cell.bias_hh.detach())
    W = f(torch.cat([_pad(cell.weight_ih.detach()[:, E:], D4), _pad(cell.weight_hh.detach(), H4)], 1))
    x1d, hd, cd = f(_pad(x1, D4)) if D else None, f(_pad(h_prev, H4)), f(_pad(c_prev, H4))
    h_out = torch.zeros(n, H4, device=cuda)
    c_out = torch.zeros(n, H4, device=cuda)
    tokd, pard = tok.to(cuda), par.to(cuda)
    cs = m.CellStep()
    cs.B, cs.H, cs.K1, cs.K2 = n, H, D4, H4
    cs.x1, cs.x1_ld = m.ptr(x1d), D4
    cs.x2, cs.x2_ld, cs.map = hd.data_ptr(), H4, pard.data_ptr()
    cs.W, cs.w_ld = W.data_ptr(), W.stride(0)
    cs.base, cs.base_ld, cs.base_row = table.data_ptr(), table.stride(0), tokd.data_ptr()
    cs.c_prev, cs.c, cs.c_ld, cs.h, cs.h_ld = cd.data_ptr(), c_out.data_ptr(), H4, h_out.data_ptr(), H4
    m.check(L.b2asr_cell_step_fwd(ctypes.byref(cs), torch.cuda.current_stream().cuda_stream), "cell_step")
    torch.cuda.synchronize()
    assert torch.allclose(h_out[:, :H].cpu().double(), h_ref, rtol=1e-4, atol=1e-5)
    assert torch.allclose(c_out[:, :H].cpu().double(), c_ref, rtol=1e-4, atol=1e-5)
    assert float(h_out[:, H:].abs().sum()) == 0.0                       # padding columns stay zero


@pytest.mark.parametrize("n,K,V", [(20, 320, 5000), (20, 650, 5000), (5, 18, 37), (33, 12, 1)])
def test_linear_lse_matches_float64(cuda, n, K, V):
    m = _lib()
    L = m.lib()
    g = torch.Generator().manual_seed(K + V)
    lin = torch.nn.Linear(K, V).double()
    x = torch.randn(n, K, generator=g, dtype=torch.float64)
    with torch.no_grad():
        ref = lin(x)
    K4 = (K + 3) // 4 * 4
    xd = _pad(x, K4).float().to(cuda).contiguous()
    W = _pad(lin.weight.detach(), K4).float().to(cuda).contiguous()
    b = lin.bias.detach().float().to(cuda)
    nct = L.b2asr_linear_lse_ctas(V)
    logits = torch.empty(n, V, device=cuda)
    lse = torch.empty(n, device=cuda)
    part = torch.empty(n, nct, 2, device=cuda)
    ticket = torch.zeros(1, dtype=torch.int32, device=cuda)
    a = m.LinearLse()
    a.B, a.K, a.V = n, K4, V
    a.x, a.x_ld, a.W, a.w_ld, a.bias = xd.data_ptr(), K4, W.data_ptr(), K4, b.data_ptr()
    a.logits, a.out_ld, a.lse, a.part, a.ticket = logits.data_ptr(), V, lse.data_ptr(), part.data_ptr(), ticket.data_ptr()
    for _ in range(2):                                                  # the ticket re-arms itself
        logits.fill_(float("nan"))
        m.check(L.b2asr_linear_lse_fwd(ctypes.byref(a), torch.cuda.current_stream().cuda_stream), "linear_lse")
        torch.cuda.synchronize()
        assert torch.allclose(logits.cpu().double(), ref, rtol=1e-4, atol=1e-5)
        assert torch.allclose(lse.cpu().double(), torch.logsumexp(ref, -1), rtol=1e-5, atol=1e-5)
        # log_softmax as the beam step forms it: logit - lse
        assert torch.allclose((logits - lse[:, None]).cpu().double(), torch.log_softmax(ref, -1), rtol=1e-4, atol=2e-5)
    assert int(ticket.item()) == 0


def test_native_decode_equals_per_step_path_at_m4_shape(cuda, monkeypatch):
    """The native loop (tables, row maps, second stream) against the per-step library path of beam.py on the shipped
    libri shape (V=5000, beam 20, RNN-LM 2x650, CTC + LM fusion): identical token sequences for every hypothesis.
    (Parity with the REFERENCE is pinned by the three beam goldens in test_beam_gpu.py, which run the native loop.)"""
    asr, rnnlm, syn = (importlib.import_module(PKG + "." + x) for x in ("asr", "rnnlm", "synthetic"))
    w = syn.WORKLOADS["libri_m2s"]
    torch.manual_seed(0)
    model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(cuda).eval()
    with torch.no_grad():
        model.char_trans.weight.mul_(3.0)                               # sharpen, so hypotheses differ in score
    model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, 0.5, 0.5
    model.rnn_lm = rnnlm.RNN_LM(650, 650, w["vocab"], layers=2).to(cuda).eval()
    x = torch.randn(1, 400, w["d_in"], generator=torch.Generator().manual_seed(3)).to(cuda)
    outs = {}
    for native in ("1", "0"):
        monkeypatch.setenv("B2ASR_DECODE_NATIVE", native)
        with torch.no_grad():
            hyps = model.beam_decode(x, 12, [400], 20)
        outs[native] = [(h.outIndex, float(h.avgScore())) for h in hyps]
    assert [a for a, _ in outs["1"]] == [a for a, _ in outs["0"]]
    assert all(abs(a - b) <= 1e-4 * abs(b) + 1e-5 for (_, a), (_, b) in zip(outs["1"], outs["0"]))


def test_concurrent_decode_equals_sequential(cuda):
    """beam.beam_decode_many (several utterances in flight on streams of their own) returns exactly what beam_decode
    returns for each utterance on its own: same tokens, same scores."""
    asr = importlib.import_module(PKG + ".asr")
    rnnlm = importlib.import_module(PKG + ".rnnlm")
    beam = importlib.import_module(PKG + ".beam")
    torch.manual_seed(3)
    cfg = {'encoder': {'enc_type': 'BiRNN', 'sample_rate': '2_2', 'sample_style': 'drop', 'dim': '32_32', 'dropout': '0_0', 'rnn_cell': 'LSTM'},
           'attention': {'att_mode': 'loc', 'dim': 32, 'proj': True, 'num_head': 1},
           'decoder': {'dim': 40, 'layer': 1, 'dropout': 0, 'rnn_cell': 'LSTMCell'}, 'optimizer': {'joint_ctc': 0.5}}
    V = 50
    model = asr.Seq2Seq(torch.zeros(1, 8, 20), V, cfg).to(cuda).eval()
    model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, 0.5, 0.3
    model.rnn_lm = rnnlm.RNN_LM(24, 24, V, layers=2).to(cuda).eval()
    utts = [(torch.randn(1, T, 20, device=cuda), 12, [T]) for T in (96, 120, 64, 200, 88, 144, 100)]
    with torch.no_grad():
        seq = [model.beam_decode(x, steps, lens, 5) for x, steps, lens in utts]
        par = beam.beam_decode_many(model, utts, 5, streams=3)
    assert len(par) == len(seq)
    for a, b in zip(seq, par):
        assert [h.outIndex for h in a] == [h.outIndex for h in b]
        assert [float(h.avgScore()) for h in a] == [float(h.avgScore()) for h in b]
