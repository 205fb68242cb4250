"""B200-native hybrid CTC/attention hot path behind the module API of
Chung-I/End-to-end-ASR-Pytorch (src/asr.py, src/ctc.py, src/solver.py).

The directory name carries a hyphen, so import it with
    importlib.import_module("end-to-end-asr-pytorch_b200")
or through the `src` shim package at the repo root (`from src.asr import Seq2Seq`), which is
what main.py / eval.py use.
"""
__version__ = "0.1.0"
