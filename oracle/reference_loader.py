"""Import the UNMODIFIED upstream reference (Chung-I/End-to-end-ASR-Pytorch) in-process.

TEST INFRASTRUCTURE ONLY.  Used in the build container (where /root/reference is
mounted) to (a) pin the CPU restatements in this directory against the real
reference and (b) generate the golden vectors committed under tests/golden/.
/root/reference does not exist on the GPU box, so nothing that runs there may
call `load()`; `available()` tells a test whether it can.

The reference needs four in-process shims on torch 2.11 / this image
 -- none of them edits /root/reference:
  1. stub modules `editdistance`, `tensorboardX`, `librosa` (absent here);
  2. `torch.ByteTensor = torch.BoolTensor` (asr.py:418 builds a uint8 mask that
     `masked_fill_` rejects on torch >= 2.x);
  3. yaml / torch.load compat live in callers, not here.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("B2ASR_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "src", "asr.py"))


def _levenshtein(a, b):
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


def _install_stubs():
    if "editdistance" not in sys.modules:
        m = types.ModuleType("editdistance")
        m.eval = _levenshtein
        sys.modules["editdistance"] = m
    if "tensorboardX" not in sys.modules:
        m = types.ModuleType("tensorboardX")

        class SummaryWriter:  # no-op logger
            def __init__(self, *a, **k):
                pass

            def __getattr__(self, name):
                return lambda *a, **k: None

        m.SummaryWriter = SummaryWriter
        sys.modules["tensorboardX"] = m
    if "librosa" not in sys.modules:
        sys.modules["librosa"] = types.ModuleType("librosa")


_cache = {}


def load():
    """Return a namespace with the reference's modules: .asr .ctc .postprocess .rnnlm .solver"""
    if "ns" in _cache:
        return _cache["ns"]
    if not available():
        raise RuntimeError("reference not mounted at %s" % REFERENCE_ROOT)
    import importlib
    import torch

    _install_stubs()
    torch.ByteTensor = torch.BoolTensor  # shim 2
    # The reference imports itself as the top-level package `src`; our repo also
    # ships a `src` shim package, so import the reference's under a private root.
    saved_path = list(sys.path)
    saved_src = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
    for k in saved_src:
        del sys.modules[k]
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        ns = types.SimpleNamespace()
        ns.ctc = importlib.import_module("src.ctc")
        ns.postprocess = importlib.import_module("src.postprocess")
        ns.asr = importlib.import_module("src.asr")
        ns.rnnlm = importlib.import_module("src.rnnlm")
        ns.solver = importlib.import_module("src.solver")
        ns.modules = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
    finally:
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            del sys.modules[k]
        sys.modules.update(saved_src)
        sys.path[:] = saved_path
    _cache["ns"] = ns
    return ns
