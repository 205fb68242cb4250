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

/* ------------------------------------------------------------------------
 * Kernel group (2b): one decoder step of Attention.forward, reference
 * src/asr.py:410-462 (query projection :422, location conv + projection :452,
 * tanh energy :453, mask + softmax(scale*e) :454-455, context bmm :457; dot
 * mode :426-432).  One launch per step; a thread-block cluster of `cluster`
 * CTAs (0 = choose; else 1,2,4,8) works on each batch row.
 *
 *   mode       0 = dot, 1 = loc
 *   s          [B,dec] decoder state, row stride s_stride
 *   Wphi       [A,dec] phi.weight, or NULL when proj=False (then A == dec, q = s)
 *   keys       [*,T,A] cached tanh(psi(h)) (asr.py:419); batch stride k_bstride
 *              (0 = every row attends over the same utterance: beam decode)
 *   h          [*,T,D] encoder features, batch stride h_bstride
 *   lens       [B] int32 valid frames per row (mask of asr.py:415-418)
 *   prev_att   [B,T] previous alignment, or NULL = uniform 1/len (asr.py:444-449)
 *   Wc [C,2Kw+1] loc_conv.weight, Wl [A,C] loc_proj.weight, wg [A] / bg [1] gen_energy
 *   att [B,T] out (0 on padding), ctx [B,D] out, q_out [B,A] out or NULL
 * ---------------------------------------------------------------------- */
int b2asr_attn_step_fwd(int mode, const float* s, int64_t s_stride, const float* Wphi,
                        const float* keys, int64_t k_bstride, const float* h, int64_t h_bstride,
                        const int32_t* lens, const float* prev_att,
                        const float* Wc, int C, int Kw, const float* Wl, const float* wg, const float* bg,
                        float scale, float* att, float* ctx, float* q_out,
                        int B, int T, int D, int A, int dec, int cluster, void* stream);

/* ------------------------------------------------------------------------
 * Backward of one attention step (the reference obtains it from autograd over
 * the ops of src/asr.py:422-457).  Activations are recomputed from q_saved
 * (q_out of the forward), prev_att and att_saved (the forward's att).
 *
 *   d_ctx [B,D], d_att [B,T] or NULL   upstream gradients of ctx / att
 *   d_keys [B,T,A]   accumulated IN PLACE (+=), batch stride dk_bstride
 *   d_prev_att [B,T] out or NULL (loc: gradient w.r.t. prev_att)
 *   d_zq [B,A] out or NULL: gradient w.r.t. the pre-tanh query phi(s)
 *   d_s  [B,dec] out or NULL: gradient w.r.t. the decoder state
 *   pgrad [b2asr_attn_slots(), pgrad_stride] accumulated IN PLACE, one private
 *        slot per CTA laid out [dWl A*C | dWc C*(2Kw+1) | dwg A | dbg 1];
 *        the caller zeroes it once per batch and sums the slots at the end.
 * ---------------------------------------------------------------------- */
int64_t b2asr_attn_pgrad_stride(int A, int C, int Kw);
int b2asr_attn_slots(int B, int T, int D, int A, int dec, int mode, int C, int Kw, int cluster);
int b2asr_attn_step_bwd(int mode, const float* Wphi, const float* keys, int64_t k_bstride,
                        const float* h, int64_t h_bstride, const int32_t* lens, const float* prev_att,
                        const float* Wc, int C, int Kw, const float* Wl, const float* wg, float scale,
                        const float* q_saved, const float* att_saved, const float* d_ctx, const float* d_att,
                        float* d_keys, int64_t dk_bstride, float* d_prev_att, float* d_zq, float* d_s,
                        float* pgrad, int64_t pgrad_stride,
                        int B, int T, int D, int A, int dec, int cluster, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2ASR_H_ */
