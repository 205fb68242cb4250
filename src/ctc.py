"""Drop-in shim: `from src.ctc import ...` (the import path of the reference's src/ctc.py) resolves to the
B200-native mirror in end-to-end-asr-pytorch_b200/ctc.py."""
import importlib as _importlib

_impl = _importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
globals().update({_k: _v for _k, _v in vars(_impl).items() if not _k.startswith("__")})
