"""The fused decoder loop (csrc/decloop.cu + deccell.cu + attn_fwd.cu / attn_bwd.cu chained over 30 steps), the
Listener recurrence, CE and CTC kernels at the SHIPPED layer widths (attention 300, decoder 320, encoder 2x320
bidirectional -> D = 640, T' = 100) against what the UNMODIFIED reference computed on the same weights and batch
(tests/golden/midsize_{loc,dot}.pt from tools/make_golden.py::midsize_fixture): forward (reference asr.py:58-112),
both losses (solver.py:149-163) and every parameter gradient (solver.py:177).

Tolerances (north_star: 1e-4 relative in fp32): outputs rtol 1e-4 with an absolute floor of 2e-5 (logits of
magnitude ~1), losses 1e-4 relative, gradients 1e-4 of the tensor's largest entry (oracle/seeded.compare_gradient;
tensors above 65536 entries are compared on a fixed 16384-entry sample plus their norm and sum)."""
import importlib
import os

import pytest
import torch

from oracle import seeded

PKG = "end-to-end-asr-pytorch_b200"
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
pytestmark = pytest.mark.gpu


def _run(cuda, g, per_step):
    asr, solver = (importlib.import_module(PKG + "." + m) for m in ("asr", "solver"))
    dec = importlib.import_module(PKG + "._dec")
    model = asr.Seq2Seq(torch.zeros(1, 8, g["x"].shape[-1]), g["vocab"], g["cfg"])
    assert {k: tuple(p.shape) for k, p in model.named_parameters()} == g["param_shapes"]
    seeded.reseed_parameters(model, g["seed"])
    model = model.to(cuda).train()
    x, y = g["x"].to(cuda), g["y"].to(cuda)
    keep = dec.supported
    if per_step:
        dec.supported = lambda d: False
    try:
        ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
        ctc_pred, enc_len, att_pred, maps = model(x, ans_len, tf_rate=1.0, teacher=y, state_len=g["lens"])
        loss, att_loss, ctc_loss = solver.hybrid_loss(ctc_pred, att_pred, y, enc_len, 0.5)
        loss.backward()
    finally:
        dec.supported = keep
    return model, ctc_pred, enc_len, att_pred, maps, loss, att_loss, ctc_loss


@pytest.mark.parametrize("per_step", [False, True], ids=["fused_loop", "per_step"])
@pytest.mark.parametrize("mode", ["loc", "dot"])
def test_train_step_at_shipped_widths_matches_reference_golden(cuda, mode, per_step):
    g = torch.load(os.path.join(G, "midsize_%s.pt" % mode), weights_only=False)
    model, ctc_pred, enc_len, att_pred, maps, loss, att_loss, ctc_loss = _run(cuda, g, per_step)
    assert [int(v) for v in enc_len] == [int(v) for v in g["enc_len"]]
    assert torch.allclose(ctc_pred.detach().cpu(), g["ctc_pred"], atol=2e-5, rtol=1e-4)
    assert torch.allclose(att_pred.detach().cpu(), g["att_pred"], atol=2e-5, rtol=1e-4)
    assert torch.allclose(maps[0].detach().cpu(), g["att_maps"], atol=2e-6, rtol=1e-4)
    for ours, ref in ((att_loss, g["att_loss"]), (ctc_loss, g["ctc_loss"]), (loss, g["loss"])):
        assert abs(float(ours.detach()) - float(ref)) <= 1e-4 * abs(float(ref))
    bad = []
    for k, p in model.named_parameters():
        assert p.grad is not None, k
        msg = seeded.compare_gradient(k, p.grad, g["grads"][k], g["seed"], rel=1e-4)
        if msg:
            bad.append(msg)
    assert not bad, bad
