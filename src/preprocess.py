"""Drop-in shim: `from src.preprocess import ...` (the import path of the reference's src/preprocess.py) resolves to the
B200-native mirror in end-to-end-asr-pytorch_b200/preprocess.py."""
import importlib as _importlib

_impl = _importlib.import_module("end-to-end-asr-pytorch_b200.preprocess")
globals().update({_k: _v for _k, _v in vars(_impl).items() if not _k.startswith("__")})
