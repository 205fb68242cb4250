"""ctypes binding of csrc/libb2asr.so (the C ABI declared in include/b2asr.h).

There is deliberately NO fallback: if the shared library is missing or was not built for
sm_100a, importing a kernel raises.  Tensors cross the boundary as raw device pointers plus
sizes/strides; the CUDA stream is torch's current stream so launches are stream-ordered and
CUDA-graph capturable.
"""
import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2ASR_LIB: another build of the same library (A/B experiments with tools/keys_variants.sh); default the in-tree one
LIB_PATH = os.environ.get("B2ASR_LIB") or os.path.join(_HERE, "csrc", "libb2asr.so")
HEADER_PATH = os.path.join(_HERE, "..", "include", "b2asr.h")

_lib = None

c_p = ctypes.c_void_p
c_i = ctypes.c_int
c_i64 = ctypes.c_int64
c_f = ctypes.c_float
c_sz = ctypes.c_size_t



class BeamIO(ctypes.Structure):
    """Mirror of `b2asr_beam_io` (include/b2asr.h); field order and types must match exactly."""
    _fields_ = [
        ("x_ctc", c_p), ("T", ctypes.c_int32), ("V", ctypes.c_int32), ("n", ctypes.c_int32), ("K", ctypes.c_int32),
        ("max_steps", ctypes.c_int32), ("w_att", c_f), ("w_ctc", c_f), ("w_lm", c_f),
        ("att_logp", c_p), ("lm_logp", c_p),
        ("valid_in", c_p), ("last_in", c_p), ("r_in", c_p), ("ctc_prob_in", c_p), ("sum_in", c_p),
        ("valid_out", c_p), ("last_out", c_p), ("r_out", c_p), ("ctc_prob_out", c_p), ("sum_out", c_p),
        ("parent_out", c_p),
        ("cand", c_p), ("psi", c_p), ("r_all", c_p), ("topv", c_p), ("topi", c_p), ("step", c_p), ("ticket", c_p),
        ("hist_parent", c_p), ("hist_tok", c_p), ("hist_score", c_p),
        ("fin_count", c_p), ("fin_step", c_p), ("fin_row", c_p), ("fin_score", c_p), ("fin_sum", c_p),
        ("att_lse", c_p), ("lm_lse", c_p),
        ("x_ctc_t", c_p), ("x_ctc_t_ld", c_i64),
    ]


class CellStep(ctypes.Structure):
    """Mirror of `b2asr_cell_step` (include/b2asr.h)."""
    _fields_ = ([(n, ctypes.c_int32) for n in ("B", "H", "K1", "K2")] +
                [("x1", c_p), ("x1_ld", c_i64), ("x2", c_p), ("x2_ld", c_i64), ("map", c_p), ("W", c_p), ("w_ld", c_i64),
                 ("base", c_p), ("base_ld", c_i64), ("base_row", c_p), ("b1", c_p), ("b2", c_p),
                 ("c_prev", c_p), ("c", c_p), ("c_ld", c_i64), ("h", c_p), ("h_ld", c_i64),
                 ("after_linear", ctypes.c_int32), ("pad_", ctypes.c_int32)])


class LinearLse(ctypes.Structure):
    """Mirror of `b2asr_linear_lse` (include/b2asr.h)."""
    _fields_ = ([(n, ctypes.c_int32) for n in ("B", "K", "V", "pad_")] +
                [("x", c_p), ("x_ld", c_i64), ("W", c_p), ("w_ld", c_i64), ("bias", c_p), ("logits", c_p),
                 ("out_ld", c_i64), ("lse", c_p), ("part", c_p), ("ticket", c_p)])


class AttnCall(ctypes.Structure):
    """Mirror of `b2asr_attn_call` (include/b2asr.h)."""
    _fields_ = ([(n, ctypes.c_int32) for n in ("mode", "C", "Kw", "B", "T", "D", "A", "dec", "cluster")] +
                [("scale", c_f), ("s", c_p), ("s_stride", c_i64), ("Wphi", c_p), ("keys", c_p), ("k_bstride", c_i64),
                 ("h", c_p), ("h_bstride", c_i64), ("lens", c_p), ("prev_att", c_p), ("row_map", c_p),
                 ("Wc", c_p), ("Wl", c_p), ("wg", c_p), ("bg", c_p), ("att", c_p), ("ctx", c_p), ("ctx_stride", c_i64)])


class DecodeProg(ctypes.Structure):
    """Mirror of `b2asr_decode_prog` (include/b2asr.h)."""
    _fields_ = [("attn", AttnCall), ("n_cells", ctypes.c_int32), ("n_linear", ctypes.c_int32),
                ("cells", CellStep * 8), ("linear", LinearLse * 2), ("beam", ctypes.POINTER(BeamIO))]


class DecLoop(ctypes.Structure):
    """Mirror of `b2asr_decloop` (include/b2asr.h); field order and types must match exactly."""
    _fields_ = ([(n, ctypes.c_int32) for n in ("B", "T", "D", "A", "H", "L", "mode", "C", "Kw", "cluster")] +
                [("scale", c_f), ("pad_", ctypes.c_int32)] +
                [(n, c_p) for n in ("keys", "h", "lens", "w_phi", "w_conv", "w_lproj", "w_gen", "b_gen", "w_ctx")] +
                [("w_ctx_ld", c_i64)] +
                [(n, c_p) for n in ("w_hh", "b_ih", "b_hh", "h0", "c0", "att", "ctx", "q", "gates", "hs", "cs",
                                    "d_hs", "d_h_last", "d_c_last", "d_gates", "d_ctx", "d_zq", "d_keys", "pgrad")] +
                [("pgrad_stride", c_i64)] +
                [(n, c_p) for n in ("d_prev", "d_s", "d_hrec", "d_c", "xcat", "w_cat", "d_xcat")])


class LstmSeq(ctypes.Structure):
    """Mirror of `b2asr_lstm_seq` (include/b2asr.h); field order and types must match exactly."""
    _fields_ = ([(n, ctypes.c_int32) for n in ("B", "T", "H", "ND")] +
                [("gates", c_p), ("w_hh", c_p * 2), ("b_ih", c_p * 2), ("b_hh", c_p * 2),
                 ("lens", c_p), ("out", c_p), ("cells", c_p), ("d_out", c_p), ("d_gates", c_p), ("status", c_p)])


class ParamChunk(ctypes.Structure):
    """Mirror of `b2asr_param_chunk` (include/b2asr.h)."""
    _fields_ = [("param", c_p), ("offset", ctypes.c_int64), ("count", ctypes.c_int32), ("reserved", ctypes.c_int32)]


# name -> (restype, argtypes); must list every symbol include/b2asr.h declares.
_SIGNATURES = {
    "b2asr_attn_keys": (c_i, [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p]),
    "b2asr_gemm_tn_3xtf32": (c_i, [c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p]),
    "b2asr_gemm_nt_3xtf32": (c_i, [c_p, c_i64, c_p, c_i64, c_p, c_i, c_i, c_i64, c_i, c_i, c_i, c_p]),
    "b2asr_gemm_nt_3xtf32_slices": (c_i, [c_i, c_i, c_i64, c_i]),
    "b2asr_ctc_prefix_init": (c_i, [c_p, c_i, c_i, c_i, c_p, c_p]),
    "b2asr_ctc_prefix_score": (c_i, [c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_i, c_p, c_p, c_p]),
    "b2asr_beam_io_size": (c_sz, []),
    "b2asr_beam_step": (c_i, [ctypes.POINTER(BeamIO), c_p]),
    "b2asr_attn_step_fwd_rows": (c_i, [c_i, c_p, c_i64, c_p, c_p, c_i64, c_p, c_i64, c_p, c_p, c_p,
                                       c_p, c_i, c_i, c_p, c_p, c_p, c_f, c_p, c_p, c_i64, c_p,
                                       c_i, c_i, c_i, c_i, c_i, c_i, c_p]),
    "b2asr_cell_step_size": (c_sz, []),
    "b2asr_linear_lse_size": (c_sz, []),
    "b2asr_decode_prog_size": (c_sz, []),
    "b2asr_linear_lse_ctas": (c_i, [c_i]),
    "b2asr_cell_step_fwd": (c_i, [ctypes.POINTER(CellStep), c_p]),
    "b2asr_linear_lse_fwd": (c_i, [ctypes.POINTER(LinearLse), c_p]),
    "b2asr_decode_run": (c_i, [ctypes.POINTER(DecodeProg), c_i, c_i, c_p]),
    "b2asr_debug_attn_trace": (c_i, [ctypes.POINTER(ctypes.c_longlong), c_i]),
    "b2asr_debug_beam_trace": (c_i, [ctypes.POINTER(ctypes.c_longlong), c_i]),
    "b2asr_lstm_cell_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_p]),
    "b2asr_lstm_cell_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_i64, c_p, c_i, c_i, c_p]),
    "b2asr_decloop_size": (c_sz, []),
    "b2asr_decoder_loop_fwd": (c_i, [ctypes.POINTER(DecLoop), c_p]),
    "b2asr_decoder_loop_fwd_range": (c_i, [ctypes.POINTER(DecLoop), c_i, c_i, c_p]),
    "b2asr_decoder_loop_bwd": (c_i, [ctypes.POINTER(DecLoop), c_p]),
    "b2asr_lstm_seq_size": (c_sz, []),
    "b2asr_lstm_seq_path": (c_i, [c_i]),
    "b2asr_lstm_seq_fwd": (c_i, [ctypes.POINTER(LstmSeq), c_p]),
    "b2asr_lstm_seq_bwd": (c_i, [ctypes.POINTER(LstmSeq), c_p]),
    "b2asr_param_chunk_size": (c_sz, []),
    "b2asr_adadelta_step": (c_i, [c_p, c_i, c_p, c_p, c_p, c_f, c_f, c_f, c_f, c_p]),
    "b2asr_adadelta_step_clipped": (c_i, [c_p, c_i, c_p, c_p, c_p, c_f, c_f, c_f, c_f, c_p, c_f, c_f, c_p, c_p]),
    "b2asr_ce_loss_fwd_bwd": (c_i, [c_p, c_i64, c_p, c_i64, c_i, c_i, c_p, c_i64, c_f, c_p, c_p, c_p, c_i64, c_p, c_p]),
    "b2asr_edit_distance": (c_i64, [c_p, c_i64, c_p, c_i64]),
    "b2asr_last_error": (ctypes.c_char_p, []),
    "b2asr_version": (c_i, []),
    "b2asr_arch": (c_i, []),
    "b2asr_ctc_workspace_bytes": (c_sz, [c_i, c_i, c_i, c_i, c_i]),
    "b2asr_ctc_loss_fwd_bwd": (c_i, [c_p, c_i64, c_i64, c_p, c_i64, c_p, c_p,
                                     c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_f,
                                     c_p, c_p, c_p, c_i64, c_i64, c_p, c_sz, c_p]),
    "b2asr_ctc_loss_fwd_bwd_static": (c_i, [c_p, c_i64, c_i64, c_p, c_i64, c_p, c_p,
                                            c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_f,
                                            c_p, c_p, c_p, c_i64, c_i64, c_p]),
    "b2asr_rescale_if_needed": (c_i, [c_p, c_i64, c_p, c_f, c_p]),
    "b2asr_attn_step_fwd": (c_i, [c_i, c_p, c_i64, c_p, c_p, c_i64, c_p, c_i64, c_p, c_p,
                                  c_p, c_i, c_i, c_p, c_p, c_p, c_f, c_p, c_p, c_p,
                                  c_i, c_i, c_i, c_i, c_i, c_i, c_p]),
    "b2asr_attn_pgrad_stride": (c_i64, [c_i, c_i, c_i]),
    "b2asr_attn_slots": (c_i, [c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i, c_i]),
    "b2asr_attn_step_bwd": (c_i, [c_i, c_p, c_p, c_i64, c_p, c_i64, c_p, c_p,
                                  c_p, c_i, c_i, c_p, c_p, c_f,
                                  c_p, c_p, c_p, c_p,
                                  c_p, c_i64, c_p, c_p, c_p, c_p, c_i64,
                                  c_i, c_i, c_i, c_i, c_i, c_i, c_p]),
}


class B2asrError(RuntimeError):
    pass


def declared_symbols():
    """Every function name include/b2asr.h declares (used by the CPU-side ABI test)."""
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2asr_[a-z0-9_]+)\s*\(", text)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2asrError(
                "libb2asr.so is missing (%s). Build it with `python __graft_entry__.py build`; "
                "there is no CPU or PyTorch fallback for the CUDA hot path." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.b2asr_beam_io_size() != ctypes.sizeof(BeamIO):
            raise B2asrError("b2asr_beam_io layout mismatch: library %d bytes, binding %d bytes"
                             % (L.b2asr_beam_io_size(), ctypes.sizeof(BeamIO)))
        if L.b2asr_decloop_size() != ctypes.sizeof(DecLoop):
            raise B2asrError("b2asr_decloop layout mismatch: library %d bytes, binding %d bytes"
                             % (L.b2asr_decloop_size(), ctypes.sizeof(DecLoop)))
        if L.b2asr_lstm_seq_size() != ctypes.sizeof(LstmSeq):
            raise B2asrError("b2asr_lstm_seq layout mismatch: library %d bytes, binding %d bytes"
                             % (L.b2asr_lstm_seq_size(), ctypes.sizeof(LstmSeq)))
        for cname, cls in (("b2asr_cell_step_size", CellStep), ("b2asr_linear_lse_size", LinearLse),
                           ("b2asr_decode_prog_size", DecodeProg)):
            if getattr(L, cname)() != ctypes.sizeof(cls):
                raise B2asrError("%s layout mismatch: library %d bytes, binding %d bytes"
                                 % (cname[:-5], getattr(L, cname)(), ctypes.sizeof(cls)))
        if L.b2asr_param_chunk_size() != ctypes.sizeof(ParamChunk):
            raise B2asrError("b2asr_param_chunk layout mismatch: library %d bytes, binding %d bytes"
                             % (L.b2asr_param_chunk_size(), ctypes.sizeof(ParamChunk)))
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().b2asr_last_error().decode("utf-8", "replace")
        if rc in (-1, -2, -3):
            raise ValueError("%s failed (%d): %s" % (what, rc, msg))
        raise B2asrError("%s failed (%d): %s" % (what, rc, msg))


# ---- launch accounting (bench.py's `gpu_launches`) and optional per-call CUDA-event timing ----------
LAUNCHES = 0          # kernels of libb2asr.so launched by this process
_prof = None          # None, or {name: [(start_event, end_event), ...]} while profiling is on


def profile_start():
    global _prof
    _prof = {}


def profile_stop():
    """-> {name: (calls, total_ms)}; synchronises the device."""
    global _prof
    rec, _prof = _prof or {}, None
    torch.cuda.synchronize()
    return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in rec.items()}


def call(name, n_kernels, fn, *args):
    """Launch through the C ABI: count the kernels, time them with events on the launching (current)
    stream when profiling is on, and turn a non-zero return code into an exception."""
    global LAUNCHES
    LAUNCHES += n_kernels
    if _prof is None:
        rc = fn(*args)
    else:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args)
        e1.record()
        _prof.setdefault(name, []).append((e0, e1))
    check(rc, name)


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return 0 if t is None else t.data_ptr()


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B2asrError("b2asr kernels need CUDA tensors; got a %s tensor (no CPU fallback)" % t.device)
