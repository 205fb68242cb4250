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
 *     never calls cudaMalloc / cudaFree.  What it does keep, per device and
 *     created on first use: one cuBLAS handle (csrc/decloop.cu; cuBLAS owns a
 *     workspace; used only when the fused decoder-step kernels do not take a
 *     shape or B2ASR_DEC_STEP=0), one extra CUDA stream + two events for the
 *     CTC call's overlapped schedule (csrc/ctc.cu) and one for the decode
 *     loop's LM chain (csrc/decode.cu), the 32 MiB static workspace of
 *     b2asr_ctc_loss_fwd_bwd_static, and two small debug-trace arrays.
 *   - every launch is stream-ordered on `stream` (work forked to the
 *     library's own stream is joined back before the call returns control of
 *     `stream`) and performs no host synchronisation.  Calls that use neither
 *     cuBLAS nor the first-use creation above are CUDA-graph capturable.
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

/* Allocation-free variant of b2asr_ctc_loss_fwd_bwd.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same arguments and results as b2asr_ctc_loss_fwd_bwd minus
 * ws / ws_bytes: the scratch lives in a file-scope device array of B2ASR_CTC_STATIC_WS_BYTES inside the library
 * (every GPU that loads the library holds it until exit).  The call itself performs no cudaMalloc / cudaFree and
 * calls no allocating library routine.  Shapes whose b2asr_ctc_workspace_bytes() exceeds the array are rejected with
 * B2ASR_EWORKSPACE and a message naming both sizes.  ALL calls on a device share the one array: nothing in it has
 * to be cleared between calls (each call rewrites what it reads), but two calls that overlap on the GPU (different
 * streams, or graph branches without a dependency) race on it -- keep them ordered. */
#define B2ASR_CTC_STATIC_WS_BYTES (32u << 20)
int b2asr_ctc_loss_fwd_bwd_static(const float* x, int64_t sx_t, int64_t sx_b,
                                  const int64_t* targets, int64_t tgt_stride,
                                  const int64_t* in_len, const int64_t* tgt_len,
                                  int T, int B, int V, int Lmax, int blank,
                                  int from_logits, int reduction, float prescale,
                                  float* nll, float* loss,
                                  float* grad, int64_t sg_t, int64_t sg_b, void* stream);

/* buf[i] *= (*upstream / assumed) for i < n, but only when *upstream != assumed
 * (device-side test; no host sync).  Used by autograd backward when the
 * upstream gradient differs from the prescale folded in at forward time. */
int b2asr_rescale_if_needed(float* buf, int64_t n, const float* upstream, float assumed, void* stream);

/* ------------------------------------------------------------------------
 * Kernel group (2a): key projection, reference src/asr.py:419
 *     self.comp_listener_feature = torch.tanh(self.psi(listener_feature))
 * out[m,a] = tanh( sum_d h[m,d] * W[a,d] + bias[a] ),  h [M,D] (M = B*T' rows, contiguous), W [A,D]
 * (psi.weight), bias [A] (psi.bias), out [M,A].  tcgen05 tensor cores, error-compensated TF32
 * (hi/lo operand split, three MMAs per k-step, fp32 accumulation in tensor memory) so the result
 * meets the 1e-4 fp32 parity bar.  D % 4 == 0; h and W 16-byte aligned.
 * ---------------------------------------------------------------------- */
int b2asr_attn_keys(const float* h, const float* W, const float* bias, float* out,
                    int M, int D, int A, void* stream);

/* The same tensor-core pipeline without the tanh: c[M,N] = a[M,K] . b[N,K]^T (+ bias[N], may be NULL), all row-major
 * and contiguous, fp32 in and out, error-compensated TF32 (max abs error ~2x cuBLAS fp32's on O(1) results).
 * Replaces the dense products around the Listener recurrence, reference src/asr.py:466-501 (nn.LSTM's input
 * projection x W_ih^T and, in backward, dG W_ih).  Returns B2ASR_EUNSUPPORTED (nothing launched) when K % 4 != 0 or
 * a / b are not 16-byte aligned: the caller then uses a library GEMM. */
int b2asr_gemm_tn_3xtf32(const float* a, const float* b, const float* bias, float* c, int M, int N, int K, void* stream);

/* Weight-gradient form on the same pipeline: c[Ma,Nb] = sum_{k < Kr} a[k,0:Ma]^T b[k,0:Nb]; a / b are row-major with
 * pitch_a / pitch_b floats between rows (views into wider matrices are fine), the reduction runs over the rows.
 * Replaces what autograd computes for nn.LSTM's weight_ih / weight_hh, reference src/asr.py:466-501 (dW_ih = dG^T x,
 * dW_hh = sum_t dG_{t+1}^T h_t).  period != 0: rows with k % period == period - 1 do not count (the dW_hh pairing of
 * row r with row r + 1 is a pointer offset; pairs straddling two utterances of `period` frames are excluded); period
 * must be 0 or >= 32.  slices >= 1 = split-K: c holds `slices` partial products [slices][Ma][Nb] which the caller adds
 * (b2asr_gemm_nt_3xtf32_slices proposes a count).  max_ctas > 0 caps the persistent grid (a launch that shares the
 * device with another persistent kernel).  Returns B2ASR_EUNSUPPORTED (nothing launched) for pitches that are not
 * multiples of 4 floats, bases that are not 16-byte aligned, or 0 < period < 32. */
int b2asr_gemm_nt_3xtf32(const float* a, long long pitch_a, const float* b, long long pitch_b, float* c, int Ma, int Nb,
                         long long Kr, int period, int slices, int max_ctas, void* stream);
int b2asr_gemm_nt_3xtf32_slices(int Ma, int Nb, long long Kr, int max_ctas);

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

/* The same step with two extras the beam decode uses (reference src/asr.py:201-207: every hypothesis restores ITS
 * parent's decoder state and previous alignment before attending):
 *   row_map    [B] int64 DEVICE or NULL: row b reads s and prev_att of row row_map[b] (the surviving child's parent
 *              row, written by b2asr_beam_step) -- no gather of the state between steps
 *   ctx_stride row stride of ctx in elements (0 = D): lets ctx land inside a wider row, e.g. [ctx | h] of the cell */
int b2asr_attn_step_fwd_rows(int mode, const float* s, int64_t s_stride, const float* Wphi,
                             const float* keys, int64_t k_bstride, const float* h, int64_t h_bstride,
                             const int32_t* lens, const float* prev_att, const int64_t* row_map,
                             const float* Wc, int C, int Kw, const float* Wl, const float* wg, const float* bg,
                             float scale, float* att, float* ctx, int64_t ctx_stride, float* q_out,
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

/* ------------------------------------------------------------------------
 * Kernel group (3): CTC prefix scorer and the joint CTC/attention/RNN-LM beam
 * step.
 *
 * b2asr_ctc_prefix_init  == CTCPrefixScore.init_state, reference src/ctc.py:19-27:
 *     r[t,1] = running fp32 sum of x[0..t, blank] (sequential order), r[t,0] = -1e8.
 * b2asr_ctc_prefix_score == CTCPrefixScore.cheap_compute, src/ctc.py:65-101
 *     (full_compute, src/ctc.py:29-63, is the same call with cands = 0..V-1 in
 *     chunks of <= 32): float32 logaddexp recursion over frames in increasing t,
 *     -1e8 "log zero" sentinel, repeated-label rule of ctc.py:87-91.
 *   x [T,V] log-posteriors (row stride V); prefix_len = len(g); last = g[-1] (0 if empty)
 *   r_prev [T,2]; cands [K] int32, K <= 32; psi [K] out; r_out [K,T,2] out.
 * ---------------------------------------------------------------------- */
int b2asr_ctc_prefix_init(const float* x, int T, int V, int blank, float* r_out, void* stream);
int b2asr_ctc_prefix_score(const float* x, int T, int V, int prefix_len, int last,
                           const float* r_prev, const int32_t* cands, int K,
                           float* psi, float* r_out, void* stream);

/* One decode step of Seq2Seq.beam_decode for ALL live hypotheses of one utterance
 * (reference src/asr.py:210-256 + Hypothesis.addTopk/avgScore, src/postprocess.py:71-106),
 * one launch, device-resident state, no host synchronisation:
 *   per row (hypothesis): top-K CTC candidates of the attention log-probs (asr.py:215),
 *   prefix scores of row.prefix+[c] (ctc.py:65-101), fused = (1-w)*att + w*scatter(ctc,-1e6),
 *   fused[0] = -1e7 (asr.py:222-229), fused += w_lm*lm (asr.py:235), top-`beam` (asr.py:238);
 *   then, by the last CTA to finish: <eos>(=1) finalises its parent, all other picks become
 *   children ranked by mean score (fp32 left-to-right sum / count), stable in (row, rank)
 *   order, best `beam` survive (asr.py:251-252); survivors' CTC state is gathered.
 * Rows are fixed at n = beam; `valid` marks live ones (step 0: only row 0).
 * All arrays are device memory owned by the caller; `in`/`out` state must not alias. */
typedef struct b2asr_beam_io {
  /* utterance constants */
  const float* x_ctc;      /* [T,V] CTC log-posteriors (asr.py:180) or NULL: joint CTC off */
  int32_t T, V;
  int32_t n;               /* rows == decode_beam_size, <= 32 */
  int32_t K;               /* CTC candidates per row = int(1.5*beam) (asr.py:160), <= 32 */
  int32_t max_steps;       /* capacity of the history arrays */
  float w_att, w_ctc, w_lm;/* float32(1-ctc_weight), float32(ctc_weight), float32(lm_weight; 0 = off) */
  /* per step inputs */
  const float* att_logp;   /* [n,V] log_softmax(char_trans(dec_out)) (asr.py:210) */
  const float* lm_logp;    /* [n,V] log_softmax(lm_out) (asr.py:235) or NULL */
  /* state in */
  const int32_t* valid_in;     /* [n] */
  const int64_t* last_in;      /* [n] last token of the prefix (0 = <sos>) */
  const float*   r_in;         /* [n,T,2] CTC prefix state */
  const float*   ctc_prob_in;  /* [n] cumulative CTC prefix log-prob (root 0) */
  const float*   sum_in;       /* [n] fp32 left-to-right sum of the prefix's token scores */
  /* state out */
  int32_t* valid_out; int64_t* last_out; float* r_out; float* ctc_prob_out; float* sum_out;
  int64_t* parent_out;     /* [n] row each survivor descends from (gather index for the neural state) */
  /* scratch */
  int32_t* cand;           /* [n,K] */
  float*   psi;            /* [n,K] */
  float*   r_all;          /* [n,K,T,2] scratch: every candidate's rows, candidate-major */
  float*   topv;           /* [n,n] */
  int32_t* topi;           /* [n,n] */
  int32_t* step;           /* [1] prefix length of the live rows; the kernel increments it */
  uint32_t* ticket;        /* [1] zero before the first step; the kernel re-arms it */
  /* history (for rebuilding token sequences on the host after the last step) */
  int32_t* hist_parent;    /* [max_steps,n]  -1 = empty slot */
  int32_t* hist_tok;       /* [max_steps,n] */
  float*   hist_score;     /* [max_steps,n] fused score of the token */
  int32_t* fin_count;      /* [1] */
  int32_t* fin_step;       /* [max_steps*n] step at which the parent was finalised */
  int32_t* fin_row;        /* [max_steps*n] parent row at that step */
  float*   fin_score;      /* [max_steps*n] fused score of <eos> */
  float*   fin_sum;        /* [max_steps*n] fp32 score sum including <eos> */
  /* optional: att_logp / lm_logp hold raw LOGITS and these the rows' log-sum-exp, so that the log_softmax of
   * asr.py:210,235 is `logit - lse` (torch's own epilogue form) inside the kernel; NULL = the inputs are log-probs */
  const float* att_lse;    /* [n] or NULL */
  const float* lm_lse;     /* [n] or NULL */
  /* optional: x_ctc transposed, [V,T] with row pitch x_ctc_t_ld >= T (multiple of 4 floats), or NULL.  With it the
   * per-step gather of the K candidates' frame columns reads K contiguous rows instead of T*K scattered sectors, and
   * utterances whose frame tables exceed the kernel's shared memory (V + 35 T floats > 200 KB) are still decoded: the
   * candidates' rows are then read from this table directly (without it such a call returns B2ASR_EUNSUPPORTED). */
  const float* x_ctc_t; int64_t x_ctc_t_ld;
} b2asr_beam_io;

size_t b2asr_beam_io_size(void);   /* sizeof(b2asr_beam_io), for binding sanity checks */
int    b2asr_beam_step(const b2asr_beam_io* io, void* stream);

/* ------------------------------------------------------------------------
 * the neural half of a decode step, batched over the beam, and the native decode loop.
 * Reference src/asr.py:201-235 (per hypothesis: attention -> Speller -> char_trans -> log_softmax -> RNN-LM step,
 * src/rnnlm.py:19-25).  csrc/decode.cu.
 *
 * b2asr_cell_step_fwd: one LSTM(Cell) layer for B rows (torch gate order i,f,g,o):
 *     pre = base[base_row[b]] + [x1[b] | x2[map[b]]] W^T + b1 + b2 ;  c = f*c_prev[map[b]] + i*g ;  h = o*tanh(c)
 *   x1 [B,K1] (row stride x1_ld), x2 [*,K2] rows taken through `map` (the surviving child's parent row; NULL = identity),
 *   W [4H, K1+K2] (row stride w_ld); K1, K2 and all strides multiples of 4, operands 16-byte aligned (pad with zeros);
 *   base [*,4H] (row stride base_ld): e.g. the embedding half of the layer folded into a [V,4H] table, row = the
 *   hypothesis' last token (base_row int64; NULL = row b; base NULL = none); c_prev / c rows have stride c_ld.
 *   after_linear: position of this layer in b2asr_decode_run's step (0 = before the first output layer, 1 = after).
 * b2asr_linear_lse_fwd: logits[b] = x[b] W^T + bias for B rows and V outputs, plus lse[b] = logsumexp(logits[b]);
 *   part [B, b2asr_linear_lse_ctas(V), 2] scratch, ticket [1] zeroed once.
 * b2asr_decode_run: enqueue decode steps first_step .. first_step+steps-1; step t runs prog[t & 1] (double-buffered
 *   state): attention (b2asr_attn_step_fwd_rows), cells before the first output layer, output layer 0, remaining
 *   cells, output layer 1, b2asr_beam_step.  No host work or synchronisation between steps.
 * ---------------------------------------------------------------------- */
typedef struct b2asr_cell_step {
  int32_t B, H, K1, K2;
  const float* x1; int64_t x1_ld;
  const float* x2; int64_t x2_ld;
  const int64_t* map;
  const float* W; int64_t w_ld;
  const float* base; int64_t base_ld; const int64_t* base_row;
  const float* b1; const float* b2;
  const float* c_prev; float* c; int64_t c_ld;
  float* h; int64_t h_ld;
  int32_t after_linear; int32_t pad_;
} b2asr_cell_step;

typedef struct b2asr_linear_lse {
  int32_t B, K, V, pad_;
  const float* x; int64_t x_ld;
  const float* W; int64_t w_ld;
  const float* bias;
  float* logits; int64_t out_ld;
  float* lse; float* part; uint32_t* ticket;
} b2asr_linear_lse;

typedef struct b2asr_attn_call {          /* the arguments of b2asr_attn_step_fwd_rows */
  int32_t mode, C, Kw, B, T, D, A, dec, cluster; float scale;
  const float* s; int64_t s_stride; const float* Wphi; const float* keys; int64_t k_bstride;
  const float* h; int64_t h_bstride; const int32_t* lens; const float* prev_att; const int64_t* row_map;
  const float* Wc; const float* Wl; const float* wg; const float* bg;
  float* att; float* ctx; int64_t ctx_stride;
} b2asr_attn_call;

typedef struct b2asr_decode_prog {
  b2asr_attn_call attn;
  int32_t n_cells, n_linear;
  b2asr_cell_step cells[8];
  b2asr_linear_lse linear[2];
  const b2asr_beam_io* beam;
} b2asr_decode_prog;

size_t b2asr_cell_step_size(void);
size_t b2asr_linear_lse_size(void);
size_t b2asr_decode_prog_size(void);
int b2asr_linear_lse_ctas(int V);
int b2asr_cell_step_fwd(const b2asr_cell_step* a, void* stream);
int b2asr_linear_lse_fwd(const b2asr_linear_lse* a, void* stream);
int b2asr_decode_run(const b2asr_decode_prog* prog, int first_step, int steps, void* stream);

/* ------------------------------------------------------------------------
 * Decoder-loop fusion: pointwise half of the
 * Speller's torch.nn.LSTMCell, reference src/asr.py:352-357 stepped from the
 * decoder loop asr.py:84-104.  The host mirror (_dec.py) batches the
 * embedding half of the input projection and all weight gradients outside the
 * loop; per step it needs two small GEMMs into `gates` and one of these.
 *
 *   gates   [B,4H] in: pre-activation sum W_ih x + W_hh h (gate order i,f,g,o
 *           as torch.nn.LSTMCell); out: the four activations, kept for backward
 *   b_ih, b_hh  [4H] or NULL
 *   c_prev  [B,H];  h, c [B,H] out
 * backward:
 *   d_h, d_c  [B,H] or NULL (= 0);  act = the forward's `gates` output
 *   d_gates   [B,4H] out with row stride dg_stride (>= 4H) so a step can write
 *             its slice of a [B,L,4H] buffer;  d_c_prev [B,H] out
 * ---------------------------------------------------------------------- */
int b2asr_lstm_cell_fwd(float* gates, const float* b_ih, const float* b_hh, const float* c_prev,
                        float* h, float* c, int B, int H, void* stream);
int b2asr_lstm_cell_bwd(const float* d_h, const float* d_c, const float* act, const float* c_prev, const float* c,
                        float* d_gates, int64_t dg_stride, float* d_c_prev, int B, int H, void* stream);

/* ------------------------------------------------------------------------
 * the whole teacher-forced decoder loop of Seq2Seq.forward, reference
 * src/asr.py:84-108 (per step: Attention.forward :86, Speller.forward :87 = nn.LSTMCell on cat([last_char, context])
 * :352-357), and its backward, enqueued by ONE call each on the caller's stream (no host work between steps).
 * Per step: b2asr_attn_step_fwd -> gates[:,t] += ctx_t W_ctx^T + h_{t-1} W_hh^T (cuBLAS SGEMM) -> LSTM cell kernel;
 * backward walks t = L-1..0: cell backward -> d_ctx_t, d_h_{t-1} (SGEMM) -> b2asr_attn_step_bwd.
 * The embedding half of W_ih is applied by the caller for all steps at once (gates arrives holding it), char_trans
 * and every weight gradient are batched GEMMs on the caller's side over the stacks below.
 *
 *   sizes     B rows, T frames, D encoder width, A attention width, H decoder width (= dec), L steps,
 *             mode 0 dot / 1 loc, C conv channels, Kw conv half width, cluster (0 = choose), scale (2.0)
 *   operands  keys [B,T,A], h [B,T,D], lens [B] int32, attention weights as in b2asr_attn_step_fwd,
 *             w_ctx = &W_ih[0][emb] with row stride w_ctx_ld (= emb + D), w_hh [4H,H], b_ih/b_hh [4H] or NULL,
 *             h0, c0 [B,H]
 *   stacks    att [L,B,T], ctx [L,B,D], q [L,B,A], gates [B,L,4H] (in: embedding projection; out: activations
 *             i,f,g,o), hs [B,L,H] (h_t, the input of char_trans), cs [L,B,H]
 *   backward  d_hs [B,L,H] upstream of hs; d_h_last / d_c_last [B,H] upstream of the final state or NULL;
 *             out: d_gates [B,L,4H] (= gradient of the embedding projection), d_ctx [L,B,D], d_zq [L,B,A],
 *             d_keys [B,T,A] and pgrad (accumulated, caller zeroes), scratch d_prev [2,B,T], d_s [B,H],
 *             d_hrec [B,H], d_c [2,B,H].  On return d_s + d_hrec is the gradient of h0 and d_c[0] that of c0.
 * ---------------------------------------------------------------------- */
typedef struct b2asr_decloop {
  int32_t B, T, D, A, H, L, mode, C, Kw, cluster;
  float scale;
  int32_t pad_;
  const float* keys; const float* h; const int32_t* lens;
  const float* w_phi; const float* w_conv; const float* w_lproj; const float* w_gen; const float* b_gen;
  const float* w_ctx; int64_t w_ctx_ld; const float* w_hh; const float* b_ih; const float* b_hh;
  const float* h0; const float* c0;
  float* att; float* ctx; float* q; float* gates; float* hs; float* cs;
  const float* d_hs; const float* d_h_last; const float* d_c_last;
  float* d_gates; float* d_ctx; float* d_zq; float* d_keys; float* pgrad; int64_t pgrad_stride;
  float* d_prev; float* d_s; float* d_hrec; float* d_c;
  /* optional fused operand layout (all three NULL = the two-GEMM form above): xcat [L,B,D+H] holds [ctx_t | h_{t-1}]
   * side by side (ctx may then be NULL), w_cat [4H,D+H] = [W_ctx | W_hh] contiguous (w_ctx / w_hh may then be NULL),
   * d_xcat [L,B,D+H] receives [d_ctx_t | recurrent d_h_{t-1}] (d_ctx / d_hrec may then be NULL; the recurrent part
   * of the gradient of h0 is d_xcat[0][:, D:]).  One product per step and direction instead of two; the forward's is
   * fused with the cell in csrc/deccell.cu when H is a multiple of 4 (B2ASR_DEC_STEP=0: cuBLAS + cell kernel). */
  float* xcat; const float* w_cat; float* d_xcat;
} b2asr_decloop;
size_t b2asr_decloop_size(void);
int b2asr_decoder_loop_fwd(const b2asr_decloop* io, void* stream);
/* Steps [t0, t1) only; the state of step t0 - 1 is read from the stacks (hs, cs, att, xcat), so consecutive ranges add
 * up to b2asr_decoder_loop_fwd.  Scheduled sampling (reference asr.py:96-101: some steps feed back a sampled character
 * instead of the teacher's) runs the loop range by range and writes the sampled step's embedding projection into
 * gates[:, t1] before continuing. */
int b2asr_decoder_loop_fwd_range(const b2asr_decloop* io, int t0, int t1, void* stream);
int b2asr_decoder_loop_bwd(const b2asr_decloop* io, void* stream);

/* ------------------------------------------------------------------------
 * The recurrence of one Listener layer, reference src/asr.py:466-501 (RNNLayer.forward: pack_padded_sequence ->
 * nn.LSTM(bidirectional, batch_first) -> pad_packed_sequence), the producer of the features kernels (1) and (2)
 * consume.  One persistent cooperative launch walks every frame of both directions (csrc/enc.cu).
 *   sizes     B rows (lengths sorted or not), T frames (= max length), H units per direction, ND directions (1|2)
 *   gates     [B,T,ND,4H] in: x W_ih^T (the caller's single GEMM; biases NOT included); out: activations i,f,g,o
 *             (PyTorch gate order) at valid frames -- the buffer the backward call reads
 *   w_hh      per direction [4H,H]; b_ih/b_hh per direction [4H] or NULL
 *   lens      [B] int32 valid frames per row
 *   forward   out [B,T,ND*H] (h_t; zeros past a row's length, like pad_packed_sequence), cells [B,T,ND,H]
 *   backward  d_out [B,T,ND*H] upstream of out -> d_gates [B,T,ND,4H] (pre-activation gradients, zeros past a
 *             row's length): the caller forms dx, dW_ih, dW_hh, db with batched GEMMs over it
 *   status    [16] int32 device words the caller zeroes once: [0] set to 1 if a launch gave up waiting for a peer
 *             CTA; diagnostics: [1] / [2] running counts of failed exchange polls forward / backward, [4..6] /
 *             [8..10] CTA 0's clock64 cycles of the last forward / backward launch in {exchange, MACs, rest}
 * out (forward) and d_gates (backward) double as the exchange buffers between CTAs and are overwritten entirely.
 * ---------------------------------------------------------------------- */
typedef struct b2asr_lstm_seq {
  int32_t B, T, H, ND;
  float* gates;
  const float* w_hh[2]; const float* b_ih[2]; const float* b_hh[2];
  const int32_t* lens;
  float* out; float* cells;
  const float* d_out; float* d_gates;
  int32_t* status;
} b2asr_lstm_seq;
size_t b2asr_lstm_seq_size(void);
int b2asr_lstm_seq_fwd(const b2asr_lstm_seq* io, void* stream);
int b2asr_lstm_seq_bwd(const b2asr_lstm_seq* io, void* stream);
/* Two implementations sit behind the calls above: 16-CTA clusters exchanging through distributed shared memory
 * (H = 256 or 320, csrc/enc_cluster.cu) and a cooperative grid exchanging through L2 (any H, csrc/enc.cu).
 * path 0 = choose (default), 1 = L2 exchange only; any other value only queries.  Returns the previous setting. */
int b2asr_lstm_seq_path(int path);

/* ------------------------------------------------------------------------
 * the attention branch's sequence loss, reference src/solver.py:90,149-155
 *     seq_loss = CrossEntropyLoss(ignore_index=0, reduction='none')
 *     att_loss = seq_loss(att_pred.view(b*t,c), label.view(-1))
 *     att_loss = torch.sum(att_loss.view(b,t),dim=-1) / torch.sum(y!=0,dim=-1).float();  att_loss = torch.mean(att_loss)
 * and the argmax cal_acc / cal_cer take of the same logits (src/postprocess.py:121-146), in ONE pass.
 *   logits [N,V] fp32 (row stride in elements), labels [N] int64, rows_per_group = decoder steps per utterance
 *   (row n belongs to group n / rows_per_group), group_len [G] int64 divisors (target lengths; NULL -> 1)
 *   loss_rows [N] out: lse - x[label], 0 where label == ignore_index        (the reduction='none' vector)
 *   argmax    [N] out (NULL to skip): first index of the row maximum
 *   grad      [N,V] out (NULL to skip): grad_prescale / (group_len * G) * (softmax - onehot), zero rows where ignored
 *   loss      [1] out (NULL to skip): mean over groups of (sum of the group's rows / group_len), fixed summation
 *             order (bit-reproducible); a second one-CTA launch
 * ---------------------------------------------------------------------- */
int b2asr_ce_loss_fwd_bwd(const float* logits, int64_t stride_row, const int64_t* labels, int64_t N, int V,
                          int rows_per_group, const int64_t* group_len, int64_t ignore_index, float grad_prescale,
                          float* loss_rows, int64_t* argmax, float* grad, int64_t grad_stride_row,
                          float* loss, void* stream);

/* ------------------------------------------------------------------------
 * SURVEY.md section 8(f) rank 4 (host code, no device work): the Levenshtein distance that eval.py:17-18 and
 * src/postprocess.py:135-146 (cal_cer) take from the third-party `editdistance` package -- `ed.eval(a, b)`.
 * a [na], b [nb]: int32 ids of the sequence items (code points, or one id per distinct word); unit costs.
 * Returns the distance, or -1 on a bad argument.
 * ---------------------------------------------------------------------- */
int64_t b2asr_edit_distance(const int32_t* a, int64_t na, const int32_t* b, int64_t nb);

/* ------------------------------------------------------------------------
 * SURVEY.md section 8 "next" (the caller after the path): the optimizer step of the train step, reference
 * src/solver.py:101-106 (asr_opt = torch.optim.Adadelta(params, lr, eps=1e-8)) and solver.py:178-182 (clip, NaN
 * guard, asr_opt.step()).  One launch over the flat gradient buffer instead of torch's seven multi-tensor passes;
 * arithmetic and order of operations of torch/optim/adadelta.py (csrc/optim.cu).
 *   chunks      [n_chunks] DEVICE table: `count` consecutive elements of parameter tensor `param` sit at `offset`
 *               of the three flat vectors (a tensor longer than one chunk appears several times)
 *   grad        flat gradients (read), square_avg / acc_delta flat optimizer state (updated), parameters updated
 *               in place through the table
 * ---------------------------------------------------------------------- */
typedef struct b2asr_param_chunk {
  float*  param;     /* first element of this chunk inside its parameter tensor */
  int64_t offset;    /* the same element's index in grad / square_avg / acc_delta */
  int32_t count;
  int32_t reserved;
} b2asr_param_chunk;
size_t b2asr_param_chunk_size(void);
int b2asr_adadelta_step(const b2asr_param_chunk* chunks, int n_chunks, const float* grad, float* square_avg,
                        float* acc_delta, float lr, float rho, float eps, float weight_decay, void* stream);

/* The same step with the clip and the NaN guard of reference src/solver.py:178-182 on the device, so that the train
 * step needs no host read of the gradient norm:
 *   grad_norm  [1] DEVICE: 2-norm of `grad` as it stands (after a data-parallel all-reduce: of the SUM over ranks)
 *   pre_scale  1 / world size (1 on one GPU): the gradient that is clipped and applied is grad * pre_scale
 *   NaN norm   -> nothing is updated and *skipped (DEVICE int, may be NULL) is incremented  (solver.py:179-180)
 *   otherwise  -> grad <- grad * pre_scale * min(max_norm / (norm * pre_scale + 1e-6), 1)   (torch clip_grad_norm_,
 *                 solver.py:178; written back, as the reference leaves the clipped gradient in .grad), then the update. */
int b2asr_adadelta_step_clipped(const b2asr_param_chunk* chunks, int n_chunks, float* grad, float* square_avg,
                                float* acc_delta, float lr, float rho, float eps, float weight_decay,
                                const float* grad_norm, float pre_scale, float max_norm, int* skipped, void* stream);

/* Debug aid, not part of the drop-in path: with B2ASR_ATT_TRACE=1 in the environment CTA 0 of every attention
 * launch stamps clock64() at its phase boundaries (forward: slots 0-9, backward: 16-27); this copies the first n
 * slots of the last launch to the host (synchronous). */
int    b2asr_debug_attn_trace(long long* host_out, int n);
/* Same for the beam step: CTA 0's phases in slots 0-7, the pruning CTA's in 8-13 (always on: 14 stores per launch). */
int    b2asr_debug_beam_trace(long long* host_out, int n);

#ifdef __cplusplus
}
#endif
#endif /* B2ASR_H_ */
