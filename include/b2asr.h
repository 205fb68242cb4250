/*
 * b2asr.h -- C ABI of the B200-native hybrid CTC/attention hot path.
 *
 * Drop-in boundary for the path BASELINE.json:north_star names in
 * Chung-I/End-to-end-ASR-Pytorch.  The reference is pure Python and has no FFI
 * of its own; every entry point below cites the reference call site whose
 * arithmetic it replaces (file:line into the upstream tree), and
 * INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 * Conventions
 *   - plain C types only: device pointers, sizes, strides (in ELEMENTS) and a
 *     CUDA stream handle passed as void* (cudaStream_t).  No torch types.
 *   - the caller owns every buffer (inputs, outputs, workspace); the library
 *     never allocates or frees device memory and keeps no global device state.
 *   - every launch is stream-ordered on `stream`, performs no host
 *     synchronisation and is CUDA-graph capturable.
 *   - return value: 0 on success, negative B2ASR_E* on failure;
 *     b2asr_last_error() returns a thread-local message.  Nothing throws.
 *   - all floating point buffers are IEEE fp32 unless noted.
 */
#ifndef B2ASR_H_
#define B2ASR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2ASR_OK            0
#define B2ASR_EINVAL       -1   /* bad argument (null pointer, negative size, misaligned buffer) */
#define B2ASR_EUNSUPPORTED -2   /* shape outside what the kernels support */
#define B2ASR_EWORKSPACE   -3   /* workspace too small */
#define B2ASR_ECUDA        -4   /* CUDA runtime error (message in b2asr_last_error) */

#define B2ASR_REDUCE_NONE 0
#define B2ASR_REDUCE_MEAN 1     /* mean_b( nll_b / max(target_len_b,1) )  == torch 'mean' */
#define B2ASR_REDUCE_SUM  2

const char* b2asr_last_error(void);
int         b2asr_version(void);
/* compiled-for architecture, e.g. 100 for sm_100a */
int         b2asr_arch(void);

/* ------------------------------------------------------------------------
 * Kernel group (1): CTC loss + gradient.
 * Replaces torch.nn.CTCLoss(blank=0, reduction='mean') as called at reference
 * src/solver.py:93,160 (train) and :253-254 (validation), optionally fused with
 * the F.log_softmax(ctc_pred.transpose(0,1), dim=-1) of src/solver.py:160.
 *
 *   x          element (t,b,v) at x[t*sx_t + b*sx_b + v]; v is contiguous.
 *              from_logits=1: raw logits (log-softmax fused; grad is w.r.t. logits)
 *              from_logits=0: log-probabilities (torch.nn.CTCLoss-compatible;
 *                             grad follows ATen's convention exp(lp) - occupancy)
 *   targets    [B, >=Lmax] int64, row stride tgt_stride, right padded (padding never read)
 *   in_len     [B] int64 DEVICE  (post-subsampling encoder lengths, <= T)
 *   tgt_len    [B] int64 DEVICE  (<= Lmax)
 *   nll        [B] out: -log p(l|x) per utterance (+inf when no alignment exists)
 *   loss       [1] out or NULL: reduced loss per `reduction`
 *   grad       out or NULL (loss only).  Same indexing as x with strides sg_t, sg_b.
 *              Holds prescale * d(reduced loss)/dx; rows t >= in_len[b] are 0;
 *              an infeasible utterance gets NaN on its valid rows (as ATen does).
 *   prescale   upstream gradient folded into grad so backward needs no extra pass
 *   ws         workspace of b2asr_ctc_workspace_bytes(), 256-byte aligned
 * ---------------------------------------------------------------------- */
size_t b2asr_ctc_workspace_bytes(int T, int B, int V, int Lmax, int want_grad);

int b2asr_ctc_loss_fwd_bwd(const float* x, int64_t sx_t, int64_t sx_b,
                           const int64_t* targets, int64_t tgt_stride,
                           const int64_t* in_len, const int64_t* tgt_len,
                           int T, int B, int V, int Lmax, int blank,
                           int from_logits, int reduction, float prescale,
                           float* nll, float* loss,
                           float* grad, int64_t sg_t, int64_t sg_b,
                           void* ws, size_t ws_bytes, void* stream);

/* buf[i] *= (*upstream / assumed) for i < n, but only when *upstream != assumed
 * (device-side test; no host sync).  Used by autograd backward when the
 * upstream gradient differs from the prescale folded in at forward time. */
int b2asr_rescale_if_needed(float* buf, int64_t n, const float* upstream, float assumed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2ASR_H_ */
