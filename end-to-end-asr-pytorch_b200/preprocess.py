"""Run-time helpers of the reference's src/preprocess.py (only the two padding functions are used while
training / decoding, dataset.py:8).  The offline feature extraction (librosa fbank/mfcc, preprocess.py:21-52)
and target encoding are corpus preparation and out of scope; `extract_feature` is kept as
a lazily-importing stub so the module imports without librosa, which this image lacks."""
import numpy as np


def zero_padding(x, pad_len):
    """list of [T_i, D] arrays -> [N, pad_len, D] float array, zero padded / cropped (preprocess.py:104-110);
    pad_len 0 means the longest item."""
    features = x[0].shape[-1]
    if pad_len == 0:
        pad_len = max(len(v) for v in x)
    out = np.zeros((len(x), pad_len, features))
    for i, item in enumerate(x):
        n = min(len(item), pad_len)
        out[i, :n, :] = item[:n, :]
    return out


def target_padding(y, max_len):
    """list of label lists -> [N, max_len] int array right padded with 0 (preprocess.py:118-123)."""
    if max_len == 0:
        max_len = max(len(v) for v in y)
    out = np.zeros((len(y), max_len), dtype=int)
    for i, seq in enumerate(y):
        out[i, :len(seq)] = np.array(seq)
    return out


def extract_feature(*args, **kwargs):
    try:
        import librosa  # noqa: F401
    except ImportError as e:
        raise ImportError("offline feature extraction needs librosa (reference src/preprocess.py:21-52); "
                          "it is corpus preparation and not part of the accelerated path") from e
    raise NotImplementedError("feature extraction is out of scope; use the reference's data/preprocess_*.py")
