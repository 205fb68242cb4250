"""Drop-in shim: `from src.solver import ...` (the import path of the reference's src/solver.py) resolves to the
B200-native mirror in end-to-end-asr-pytorch_b200/solver.py."""
import importlib as _importlib

_impl = _importlib.import_module("end-to-end-asr-pytorch_b200.solver")
globals().update({_k: _v for _k, _v in vars(_impl).items() if not _k.startswith("__")})
