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

_HERE = os.path.dirname(os.path.abspath(__file__))
# where the reference can be: the build container's read-only mount, or the copy oracle/make_ref.py places under
# oracle/_ref/ (git-ignored, travels with the gpurun snapshot) so bench.py's CPU arm can time the real reference
UPSTREAM_ROOT = "/root/reference"
SHIPPED_ROOT = os.path.join(_HERE, "_ref")


def reference_root():
    env = os.environ.get("B2ASR_REFERENCE_ROOT")
    for cand in ([env] if env else []) + [UPSTREAM_ROOT, SHIPPED_ROOT]:
        if os.path.isfile(os.path.join(cand, "src", "asr.py")):
            return os.path.realpath(cand)
    return None


REFERENCE_ROOT = reference_root() or UPSTREAM_ROOT


def available() -> bool:
    return reference_root() is not None


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
_MODULES = ("preprocess", "ctc", "postprocess", "dataset", "asr", "rnnlm", "clm", "solver")


def load(root=None):
    """Return a namespace with the reference's modules: .asr .ctc .postprocess .rnnlm .solver ...

    The reference imports itself as the top-level package `src` (a NAMESPACE package: upstream has no
    src/__init__.py).  This repo ships a regular `src` shim package, and Python prefers a regular package anywhere on
    sys.path over a namespace portion, so `import src.asr` with both on the path would return the B200 mirror.
    Here the reference's files are therefore imported BY PATH: a private package object whose __path__ is the
    reference's src/ directory stands in as sys.modules['src'] for the duration of the import, every module's
    __file__ is asserted to lie under the reference root, and the caller's own `src` modules are put back after."""
    root = os.path.realpath(root) if root else reference_root()
    if root is None:
        raise RuntimeError("reference not found at %s or %s" % (UPSTREAM_ROOT, SHIPPED_ROOT))
    if root in _cache:
        return _cache[root]
    import importlib
    import importlib.machinery
    import importlib.util
    import torch

    _install_stubs()
    torch.ByteTensor = torch.BoolTensor  # shim 2
    src_dir = os.path.join(root, "src")
    saved = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
    for k in saved:
        del sys.modules[k]
    private = types.ModuleType("src")
    private.__path__ = [src_dir]
    private.__package__ = "src"
    private.__spec__ = importlib.machinery.ModuleSpec("src", None, is_package=True)
    private.__spec__.submodule_search_locations = [src_dir]
    sys.modules["src"] = private
    try:
        ns = types.SimpleNamespace(root=root)
        for name in _MODULES:
            path = os.path.join(src_dir, name + ".py")
            if not os.path.isfile(path):
                continue
            full = "src." + name
            mod = sys.modules.get(full)
            if mod is None:
                spec = importlib.util.spec_from_file_location(full, path)
                mod = importlib.util.module_from_spec(spec)
                sys.modules[full] = mod
                spec.loader.exec_module(mod)
                setattr(private, name, mod)
            setattr(ns, name, mod)
        ns.modules = {k: v for k, v in sys.modules.items() if k == "src" or k.startswith("src.")}
        for k, v in ns.modules.items():
            f = getattr(v, "__file__", None)
            if k != "src" and not (f and os.path.realpath(f).startswith(root + os.sep)):
                raise RuntimeError("reference_loader: %s resolved to %s, outside the reference root %s" % (k, f, root))
    finally:
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    _cache[root] = ns
    return ns
