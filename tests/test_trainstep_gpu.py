"""Whole hybrid train step on the GPU against what the UNMODIFIED reference computed on the same weights and batch
(tests/golden/trainstep.pt, produced by tools/make_golden.py): Seq2Seq.forward (reference asr.py:58-112) through the
fused decoder loop + attention kernels, CE + CTC loss (solver.py:149-163) through the CTC kernels, and every
parameter gradient of loss.backward() (solver.py:177)."""
import importlib
import os

import pytest
import torch

PKG = "end-to-end-asr-pytorch_b200"
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
pytestmark = pytest.mark.gpu


def test_train_step_matches_reference_golden(cuda):
    asr, solver = (importlib.import_module(PKG + "." + m) for m in ("asr", "solver"))
    g = torch.load(os.path.join(G, "trainstep.pt"), weights_only=False)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False              # the reference ran fp32 on the CPU; cuDNN's RNN defaults to TF32
    try:
        model = asr.Seq2Seq(torch.zeros(1, 8, g["x"].shape[-1]), g["vocab"], g["cfg"])
        model.load_state_dict(g["state_dict"], strict=True)
        model = model.to(cuda).train()
        x, y = g["x"].to(cuda), g["y"].to(cuda)
        assert solver.infer_lengths(x) == g["lens"]
        ans_len = int(torch.max(torch.sum(y != 0, dim=-1)))
        ctc_pred, enc_len, att_pred, maps = model(x, ans_len, tf_rate=1.0, teacher=y, state_len=g["lens"])
        assert [int(v) for v in enc_len] == [int(v) for v in g["enc_len"]]
        assert torch.allclose(ctc_pred.cpu(), g["ctc_pred"], atol=2e-5, rtol=1e-4)
        assert torch.allclose(att_pred.cpu(), g["att_pred"], atol=2e-5, rtol=1e-4)
        assert torch.allclose(maps[0].cpu(), g["att_maps"], atol=2e-6, rtol=1e-4)
        loss, att_loss, ctc_loss = solver.hybrid_loss(ctc_pred, att_pred, y, enc_len, 0.5)
        # tolerance: 1e-4 relative (north_star) on the losses
        assert abs(float(att_loss) - float(g["att_loss"])) <= 1e-4 * abs(float(g["att_loss"]))
        assert abs(float(ctc_loss) - float(g["ctc_loss"])) <= 1e-4 * abs(float(g["ctc_loss"]))
        assert abs(float(loss) - float(g["loss"])) <= 1e-4 * abs(float(g["loss"]))
        loss.backward()
        for k, p in model.named_parameters():
            ref = g["grads"][k]
            tol = 1e-4 * float(ref.abs().max()) + 1e-7   # 1e-4 of the tensor's largest entry (most entries are ~0)
            assert p.grad is not None, k
            assert float((p.grad.cpu() - ref).abs().max()) <= tol, (k, float((p.grad.cpu() - ref).abs().max()), tol)
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
