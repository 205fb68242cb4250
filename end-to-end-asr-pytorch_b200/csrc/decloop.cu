// Native decoder loop, SURVEY.md section 8(f) rank 1: the L-step teacher-forced loop of Seq2Seq.forward
// (reference src/asr.py:84-108: attend -> LSTMCell -> next step) and its backward, enqueued from C++.
//
// Why: with every kernel of the step tuned, the Python loop became the bound -- 133 us of host time per decoder
// step against ~55 us of GPU work forward, and as much again in autograd's backward walk (tools/phase_wall.py:
// every phase's wall time equalled its enqueue time).  Here ONE C call enqueues all L steps on the caller's stream:
//   forward, per step t:   attention step kernel (attn_fwd.cu)  ->  gates[:,t] += ctx_t W_ctx^T + h_{t-1} W_hh^T
//                          (two plain cuBLAS SGEMMs accumulating into the slice of the batched embedding
//                          projection)  ->  LSTM cell kernel (lstm.cu) writing h_t straight into the [B,L,H]
//                          stack that char_trans consumes
//   backward, t = L-1..0:  LSTM cell backward (sums the three paths into h_t: output layer, recurrence, next
//                          step's attention query)  ->  d_ctx_t, d_h_{t-1} (two SGEMMs)  ->  attention step backward
// Every per-step tensor lives in a stack allocated once per batch, so no allocator or Python work happens inside
// the loop; the weight gradients stay deferred to a few batched GEMMs on the host side (_decloop.py).
// The arithmetic per step is the same kernels and the same cuBLAS calls as the per-step path (_dec.py), which
// stays for scheduled sampling (a step whose coin fails needs the output layer and a sample before the next step).
#include <cublas_v2.h>

#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace b2asr {
void lstm_cell_fwd_launch(float* gates, long long g_stride, const float* b_ih, const float* b_hh, const float* c_prev,
                          float* h, long long h_stride, float* c, int B, int H, cudaStream_t st, float* h2,
                          long long h2_stride);
void lstm_cell_bwd_launch(const float* d_h, long long dh_stride, const float* d_h2, const float* d_h3, const float* d_c,
                          const float* act, long long act_stride, const float* c_prev, const float* c, float* d_gates,
                          long long dg_stride, float* d_c_prev, int B, int H, cudaStream_t st, long long dh2_stride);
extern thread_local long long g_attn_row_stride;   // attn.cuh: row stride of ctx / d_ctx for the next attention launch
// deccell.cu: the forward step's GEMM + cell as one hand-written launch
bool dec_step_supported(int H, int K, long long x_ld, long long g_ld);
bool dec_step_bwd_supported(int N, int X, long long dg_ld, long long w_ld);
int dec_step_bwd(const float* dg, long long dg_ld, const float* W, long long w_ld, int B, int N, int X, float* out,
                 long long out_ld, cudaStream_t st);
int dec_step_fwd(const float* x, long long x_ld, const float* w_cat, int B, int K, int H, float* gates, long long g_ld,
                 const float* b_ih, const float* b_hh, const float* c_prev, float* c, float* h, long long h_ld, float* h2,
                 long long h2_ld, cudaStream_t st);
}  // namespace b2asr

namespace {

std::mutex g_mu;
cublasHandle_t g_handle[64] = {};

int blas_handle(cublasHandle_t* out, cudaStream_t st) {
  int dev = 0;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < 64, B2ASR_EINVAL, "decoder loop: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_handle[dev]) {
    const cublasStatus_t rc = cublasCreate(&g_handle[dev]);
    B2_REQUIRE(rc == CUBLAS_STATUS_SUCCESS, B2ASR_ECUDA, "cublasCreate failed (%d)", (int)rc);
    cublasSetMathMode(g_handle[dev], CUBLAS_DEFAULT_MATH);           // plain fp32 SGEMM (no TF32), as torch's default
    cublasSetAtomicsMode(g_handle[dev], CUBLAS_ATOMICS_NOT_ALLOWED);
  }
  const cublasStatus_t rc = cublasSetStream(g_handle[dev], st);
  B2_REQUIRE(rc == CUBLAS_STATUS_SUCCESS, B2ASR_ECUDA, "cublasSetStream failed (%d)", (int)rc);
  *out = g_handle[dev];
  return B2ASR_OK;
}

#define B2_BLAS_TRY(expr)                                                                     \
  do {                                                                                        \
    const cublasStatus_t rc_ = (expr);                                                        \
    B2_REQUIRE(rc_ == CUBLAS_STATUS_SUCCESS, B2ASR_ECUDA, "cuBLAS call failed (%d): %s", (int)rc_, #expr); \
  } while (0)

// row-major C[B,N] (ldc) = beta*C + X[B,K] (ldx) * W[N,K]^T (ldw)
inline cublasStatus_t gemm_xwt(cublasHandle_t hd, int Bn, int N, int K, const float* X, int64_t ldx, const float* W,
                               int64_t ldw, float beta, float* Cm, int64_t ldc) {
  const float one = 1.f;
  return cublasSgemm(hd, CUBLAS_OP_T, CUBLAS_OP_N, N, Bn, K, &one, W, (int)ldw, X, (int)ldx, &beta, Cm, (int)ldc);
}
// row-major C[B,K] (ldc) = G[B,N] (ldg) * W[N,K] (ldw)
inline cublasStatus_t gemm_gw(cublasHandle_t hd, int Bn, int N, int K, const float* G, int64_t ldg, const float* W,
                              int64_t ldw, float* Cm, int64_t ldc) {
  const float one = 1.f, zero = 0.f;
  return cublasSgemm(hd, CUBLAS_OP_N, CUBLAS_OP_N, K, Bn, N, &one, W, (int)ldw, G, (int)ldg, &zero, Cm, (int)ldc);
}

// B2ASR_DEC_STEP=0 keeps the per-step products on cuBLAS (A/B runs)
bool fused_step_enabled() {
  static const bool on = [] { const char* e = getenv("B2ASR_DEC_STEP"); return !(e && e[0] == '0'); }();
  return on;
}

int check_common(const b2asr_decloop* a) {
  B2_REQUIRE(a, B2ASR_EINVAL, "decoder loop: null argument block");
  B2_REQUIRE(a->B > 0 && a->T > 0 && a->D > 0 && a->A > 0 && a->H > 0 && a->L > 0, B2ASR_EINVAL,
             "decoder loop: bad sizes B=%d T=%d D=%d A=%d H=%d L=%d", a->B, a->T, a->D, a->A, a->H, a->L);
  B2_REQUIRE(a->keys && a->h && a->lens && ((a->w_ctx && a->w_hh) || a->w_cat) && a->h0 && a->c0, B2ASR_EINVAL,
             "decoder loop: null operand");
  B2_REQUIRE(a->att && (a->ctx || (a->xcat && a->w_cat)) && a->q && a->gates && a->hs && a->cs, B2ASR_EINVAL,
             "decoder loop: null step stack");
  B2_REQUIRE(a->w_cat || a->w_ctx_ld >= a->D, B2ASR_EINVAL, "decoder loop: w_ctx_ld smaller than D");
  return B2ASR_OK;
}

}  // namespace

using namespace b2asr;

extern "C" size_t b2asr_decloop_size(void) { return sizeof(b2asr_decloop); }

extern "C" int b2asr_decoder_loop_fwd(const b2asr_decloop* a, void* stream) {
  return b2asr_decoder_loop_fwd_range(a, 0, a ? a->L : 0, stream);
}

extern "C" int b2asr_decoder_loop_fwd_range(const b2asr_decloop* a, int t0, int t1, void* stream) {
  if (int rc = check_common(a)) return rc;
  B2_REQUIRE(0 <= t0 && t0 <= t1 && t1 <= a->L, B2ASR_EINVAL, "decoder loop forward: steps [%d,%d) of %d", t0, t1, a->L);
  cudaStream_t st = (cudaStream_t)stream;
  cublasHandle_t hd;
  if (int rc = blas_handle(&hd, st)) return rc;
  const int B = a->B, T = a->T, D = a->D, A = a->A, H = a->H, L = a->L;
  const int64_t gl = (int64_t)L * 4 * H, hl = (int64_t)L * H;
  // fused operand layout: xcat[t] = [ctx_t | h_{t-1}] (row stride D+H) against w_cat = [W_ctx | W_hh]: ONE GEMM per step
  const bool cat = a->xcat && a->w_cat;
  const int X = D + H;
  const bool fused = cat && fused_step_enabled() && dec_step_supported(H, X, X, gl);
  if (cat && t0 == 0 && t1 > 0)
    B2_CUDA_TRY(cudaMemcpy2DAsync(a->xcat + D, sizeof(float) * X, a->h0, sizeof(float) * H, sizeof(float) * H, B,
                                  cudaMemcpyDeviceToDevice, st));
  for (int t = t0; t < t1; ++t) {
    const float* s = t ? a->hs + (int64_t)(t - 1) * H : a->h0;
    const int64_t ss = t ? hl : H;
    const float* prev = (a->mode == 1 && t) ? a->att + (int64_t)(t - 1) * B * T : nullptr;
    float* att_t = a->att + (int64_t)t * B * T;
    float* ctx_t = cat ? a->xcat + (int64_t)t * B * X : a->ctx + (int64_t)t * B * D;
    if (cat) g_attn_row_stride = X;
    if (int rc = b2asr_attn_step_fwd(a->mode, s, ss, a->w_phi, a->keys, (int64_t)T * A, a->h, (int64_t)T * D, a->lens, prev,
                                     a->w_conv, a->C, a->Kw, a->w_lproj, a->w_gen, a->b_gen, a->scale, att_t, ctx_t,
                                     a->q + (int64_t)t * B * A, B, T, D, A, H, a->cluster, stream))
      return rc;
    float* g_t = a->gates + (int64_t)t * 4 * H;
    float* h2 = nullptr;
    if (cat) {
      if (t + 1 < L) h2 = a->xcat + (int64_t)(t + 1) * B * X + D;
      if (fused) {                              // GEMM + cell in one launch (deccell.cu)
        if (int rc = dec_step_fwd(ctx_t, X, a->w_cat, B, X, H, g_t, gl, a->b_ih, a->b_hh,
                                  t ? a->cs + (int64_t)(t - 1) * B * H : a->c0, a->cs + (int64_t)t * B * H,
                                  a->hs + (int64_t)t * H, hl, h2, X, st))
          return rc;
        continue;
      }
      B2_BLAS_TRY(gemm_xwt(hd, B, 4 * H, X, ctx_t, X, a->w_cat, X, 1.f, g_t, gl));
    } else {
      B2_BLAS_TRY(gemm_xwt(hd, B, 4 * H, D, ctx_t, D, a->w_ctx, a->w_ctx_ld, 1.f, g_t, gl));
      B2_BLAS_TRY(gemm_xwt(hd, B, 4 * H, H, s, ss, a->w_hh, H, 1.f, g_t, gl));
    }
    lstm_cell_fwd_launch(g_t, gl, a->b_ih, a->b_hh, t ? a->cs + (int64_t)(t - 1) * B * H : a->c0, a->hs + (int64_t)t * H, hl,
                         a->cs + (int64_t)t * B * H, B, H, st, h2, X);
  }
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

extern "C" int b2asr_decoder_loop_bwd(const b2asr_decloop* a, void* stream) {
  if (int rc = check_common(a)) return rc;
  B2_REQUIRE(a->d_hs && a->d_gates && ((a->d_ctx && a->d_hrec) || (a->d_xcat && a->w_cat)) && a->d_zq && a->d_keys &&
                 a->d_s && a->d_c, B2ASR_EINVAL, "decoder loop backward: null gradient buffer");
  B2_REQUIRE(a->mode != 1 || (a->pgrad && a->d_prev), B2ASR_EINVAL, "decoder loop backward: loc mode needs pgrad and d_prev");
  cudaStream_t st = (cudaStream_t)stream;
  cublasHandle_t hd;
  if (int rc = blas_handle(&hd, st)) return rc;
  const int B = a->B, T = a->T, D = a->D, A = a->A, H = a->H, L = a->L;
  const int64_t gl = (int64_t)L * 4 * H, hl = (int64_t)L * H;
  const bool loc = a->mode == 1;
  // fused layout: d_xcat[t] = d_gates_t [W_ctx | W_hh] = [d_ctx_t | d_h_{t-1} (recurrent path)], ONE GEMM per step
  const bool cat = a->d_xcat && a->w_cat;
  const int X = D + H;
  for (int t = L - 1; t >= 0; --t) {
    const bool last = t == L - 1;
    float* dc_in = a->d_c + (int64_t)((t + 1) & 1) * B * H;
    float* dc_out = a->d_c + (int64_t)(t & 1) * B * H;
    float* dg_t = a->d_gates + (int64_t)t * 4 * H;
    const float* d_rec = cat ? a->d_xcat + (int64_t)(t + 1) * B * X + D : a->d_hrec;   // only read when !last
    lstm_cell_bwd_launch(a->d_hs + (int64_t)t * H, hl, last ? a->d_h_last : d_rec, last ? nullptr : a->d_s,
                         last ? a->d_c_last : dc_in, a->gates + (int64_t)t * 4 * H, gl,
                         t ? a->cs + (int64_t)(t - 1) * B * H : a->c0, a->cs + (int64_t)t * B * H, dg_t, gl, dc_out, B, H, st,
                         (cat && !last) ? X : 0);
    float* dctx_t = cat ? a->d_xcat + (int64_t)t * B * X : a->d_ctx + (int64_t)t * B * D;
    if (cat) {
      if (fused_step_enabled() && dec_step_bwd_supported(4 * H, X, gl, X)) {
        if (int rc = dec_step_bwd(dg_t, gl, a->w_cat, X, B, 4 * H, X, dctx_t, X, st)) return rc;
      } else {
        B2_BLAS_TRY(gemm_gw(hd, B, 4 * H, X, dg_t, gl, a->w_cat, X, dctx_t, X));
      }
      g_attn_row_stride = X;
    } else {
      B2_BLAS_TRY(gemm_gw(hd, B, 4 * H, D, dg_t, gl, a->w_ctx, a->w_ctx_ld, dctx_t, D));
      B2_BLAS_TRY(gemm_gw(hd, B, 4 * H, H, dg_t, gl, a->w_hh, H, a->d_hrec, H));
    }
    const float* prev = (loc && t) ? a->att + (int64_t)(t - 1) * B * T : nullptr;
    const float* d_att = (loc && !last) ? a->d_prev + (int64_t)((t + 1) & 1) * B * T : nullptr;
    float* d_prev = (loc && t) ? a->d_prev + (int64_t)(t & 1) * B * T : nullptr;
    if (int rc = b2asr_attn_step_bwd(a->mode, a->w_phi, a->keys, (int64_t)T * A, a->h, (int64_t)T * D, a->lens, prev, a->w_conv,
                                     a->C, a->Kw, a->w_lproj, a->w_gen, a->scale, a->q + (int64_t)t * B * A,
                                     a->att + (int64_t)t * B * T, dctx_t, d_att, a->d_keys, (int64_t)T * A, d_prev,
                                     a->d_zq + (int64_t)t * B * A, a->d_s, a->pgrad, a->pgrad_stride, B, T, D, A, H,
                                     a->cluster, stream))
      return rc;
  }
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
