"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/b2asr.h declares;
argument validation works without a GPU (no compute calls here); the product path refuses CPU tensors."""
import ctypes
import importlib
import os

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    b = importlib.import_module("end-to-end-asr-pytorch_b200.build")
    b.build()
    return importlib.import_module("end-to-end-asr-pytorch_b200._lib")


def test_every_declared_symbol_is_exported_and_bound(lib):
    L = lib.lib()
    declared = lib.declared_symbols()
    assert len(declared) >= 14
    for s in declared:
        assert hasattr(L, s), "libb2asr.so lacks %s" % s
        assert s in lib._SIGNATURES, "no ctypes signature for %s" % s
    assert L.b2asr_arch() == 100
    assert L.b2asr_beam_io_size() == ctypes.sizeof(lib.BeamIO)


def test_bad_arguments_are_rejected_with_codes_not_crashes(lib):
    L = lib.lib()
    assert L.b2asr_ctc_workspace_bytes(0, 1, 5, 2, 1) == 0
    assert L.b2asr_ctc_workspace_bytes(100, 4, 31, 20, 1) > 0
    rc = L.b2asr_ctc_loss_fwd_bwd(0, 0, 0, 0, 0, 0, 0, 10, 2, 5, 3, 0, 1, 1, 1.0, 0, 0, 0, 0, 0, 0, 0, 0)
    assert rc == -1 and b"null" in L.b2asr_last_error()
    rc = L.b2asr_attn_step_fwd(1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 10, 100, 0, 0, 0, 2.0, 0, 0, 0, 2, 8, 8, 8, 8, 0, 0)
    assert rc == -1
    assert L.b2asr_beam_step(None, 0) == -1
    assert L.b2asr_ctc_prefix_score(0, 10, 5, 0, 0, 0, 0, 4, 0, 0, 0) == -1
    assert L.b2asr_attn_pgrad_stride(300, 10, 100) == 300 * 10 + 10 * 201 + 300 + 1


def test_product_path_has_no_cpu_fallback(lib):
    ctc = importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
    asr = importlib.import_module("end-to-end-asr-pytorch_b200.asr")
    x = torch.randn(5, 2, 4)
    with pytest.raises(lib.B2asrError):
        ctc.ctc_loss(x, torch.ones(2, 2, dtype=torch.long), [5, 5], torch.tensor([2, 2]))
    att = asr.Attention(8, 8, 'loc', 8, True, 1)
    with pytest.raises(RuntimeError):
        att(torch.randn(2, 8), torch.randn(2, 5, 8), [5, 4])
    with pytest.raises(lib.B2asrError):
        ctc.CTCPrefixScore(torch.randn(1, 5, 4))
