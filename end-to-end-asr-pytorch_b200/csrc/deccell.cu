// Decoder step, fused: the Speller's LSTMCell of reference src/asr.py:352-357 (nn.LSTMCell on cat([last_char, context]))
// as the native loop of decloop.cu steps it -- gates_t = E_t + [ctx_t | h_{t-1}] [W_ctx | W_hh]^T, cell, h_t -- in
// ONE launch per step instead of a cuBLAS SGEMM (8.5 us at 32 x 960 x 1280: latency-bound) plus the cell kernel.
//
//   grid = H/4 CTAs, a CTA owns 4 hidden units = 16 gate rows of the concatenated weight (staged once in shared
//   memory with cp.async, 61 KB at M2); warp w owns batch rows 4w..4w+3 and reads them straight from global memory
//   two k-blocks ahead (every x element is needed by exactly one warp of the CTA, so staging it would only add a
//   shared-memory round trip); the 32 lanes split K, a 5-round transposing butterfly leaves the four gates of one
//   (row, unit) on a lane pair, whose even lane applies the cell and writes activations, c_t, h_t (twice: the
//   [B,L,H] stack and next step's operand).
//
// Measured at M2 (B200): decoder-loop forward 10.4 -> 8.25 ms per batch of 189 steps.  The same body without the cell
// for the backward's product [d_ctx_t | d_h_{t-1}] = d_gates_t [W_ctx | W_hh] was tried in two decompositions (16
// and 8 output columns per CTA over the transposed weight, operands staged with chunked cp.async) and did not beat
// cuBLAS's split-K (15.8 / 16.0 vs 16.0 ms): every CTA re-reads the whole 164 KB d_gates operand from L2.  The
// backward therefore keeps the library call.
#include "common.cuh"

namespace b2asr {

bool pdl_enabled();                           // attn_fwd.cu

namespace {

constexpr int kRowsPerWarp = 4;
constexpr int kCols = 16;                     // 4 units x 4 gates per CTA

struct StepArgs {
  const float* x; long long x_ld;            // [B,K] rows (K contiguous): [ctx_t | h_{t-1}]
  const float* W; long long w_ld;            // [4H,K] rows (K contiguous): [W_ctx | W_hh]
  int B, K, H;
  float* gates; long long g_ld;              // in: embedding projection, out: activations
  const float* b_ih; const float* b_hh;
  const float* c_prev; float* c;             // [B,H]
  float* h; long long h_ld; float* h2; long long h2_ld;
};

__device__ __forceinline__ float sigmoid_d(float v) { return __fdividef(1.f, 1.f + ex2(-v * kLog2e)); }

__device__ __forceinline__ float4 ld_x(const StepArgs& a, int b, int k) {
  return (b < a.B && k < a.K) ? __ldg(reinterpret_cast<const float4*>(a.x + (long long)b * a.x_ld + k))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
}

// column c of the CTA is gate row  g*H + 4*blockIdx.x + uu  with c = uu*4 + g
__global__ void __launch_bounds__(256) dec_step_kernel(const StepArgs a) {
  extern __shared__ __align__(16) float W_s[];             // [16][Kp]
  pdl_trigger_small();
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int K = a.K, Kp = (K + 127) & ~127;
  for (int idx = tid; idx < kCols * (Kp >> 2); idx += 256) {
    const int c = idx / (Kp >> 2), k = (idx - c * (Kp >> 2)) << 2;
    const long long row = (long long)(c & 3) * a.H + 4 * blockIdx.x + (c >> 2);
    float* dst = W_s + c * Kp + k;
    if (k < K) cp_async16(dst, a.W + row * a.w_ld + k);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cp_async_commit();
  pdl_wait();                             // the weights above do not depend on the attention kernel before this one
  pdl_launch();

  // the first two k-blocks of this warp's first rows are in flight while the weights land
  float4 x0[kRowsPerWarp], x1[kRowsPerWarp];
#pragma unroll
  for (int i = 0; i < kRowsPerWarp; ++i) {
    x0[i] = ld_x(a, kRowsPerWarp * w + i, 4 * lane);
    x1[i] = ld_x(a, kRowsPerWarp * w + i, 128 + 4 * lane);
  }
  cp_async_wait_all();
  __syncthreads();

  for (int b0 = kRowsPerWarp * w; b0 < a.B; b0 += kRowsPerWarp * 8) {
    float acc[kRowsPerWarp][kCols];
#pragma unroll
    for (int i = 0; i < kRowsPerWarp; ++i)
#pragma unroll
      for (int c = 0; c < kCols; ++c) acc[i][c] = 0.f;
    if (b0 != kRowsPerWarp * w) {                           // later row tiles (B > 32)
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) {
        x0[i] = ld_x(a, b0 + i, 4 * lane);
        x1[i] = ld_x(a, b0 + i, 128 + 4 * lane);
      }
    }
    for (int kb = 0; kb < Kp; kb += 128) {
      float4 x2[kRowsPerWarp];
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) x2[i] = ld_x(a, b0 + i, kb + 256 + 4 * lane);
      const float* wp = W_s + kb + 4 * lane;
#pragma unroll
      for (int c = 0; c < kCols; ++c) {
        const float4 wv = *reinterpret_cast<const float4*>(wp + c * Kp);
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          acc[i][c] = fmaf(x0[i].x, wv.x, acc[i][c]);
          acc[i][c] = fmaf(x0[i].y, wv.y, acc[i][c]);
          acc[i][c] = fmaf(x0[i].z, wv.z, acc[i][c]);
          acc[i][c] = fmaf(x0[i].w, wv.w, acc[i][c]);
        }
      }
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) { x0[i] = x1[i]; x1[i] = x2[i]; }
    }
    // ---- transposing butterfly over the 32 k-lanes: 64 sums -> 2 per lane -----------------------------------
    float* v = &acc[0][0];                                  // v[i*16 + c]
#pragma unroll
    for (int m = 16, n = 32; m >= 1; m >>= 1, n >>= 1) {
      const bool hi = lane & m;
#pragma unroll
      for (int q = 0; q < n; ++q) {
        const float send = hi ? v[q] : v[q + n];
        const float keep = hi ? v[q + n] : v[q];
        v[q] = keep + __shfl_xor_sync(kFull, send, m);
      }
    }
    // lane bits: [4:3] = row i, [2:1] = unit uu, [0] = gate bit 1; v[0], v[1] = gate bit 0
    const int i = lane >> 3, b = b0 + i;
    const float o0 = __shfl_xor_sync(kFull, v[0], 1), o1 = __shfl_xor_sync(kFull, v[1], 1);
    if ((lane & 1) == 0 && b < a.B) {
      const int uu = (lane >> 1) & 3, u = 4 * blockIdx.x + uu, H = a.H;
      float* g = a.gates + (long long)b * a.g_ld;
      float bi = 0.f, bf = 0.f, bg = 0.f, bo = 0.f;
      if (a.b_ih) { bi += a.b_ih[u]; bf += a.b_ih[H + u]; bg += a.b_ih[2 * H + u]; bo += a.b_ih[3 * H + u]; }
      if (a.b_hh) { bi += a.b_hh[u]; bf += a.b_hh[H + u]; bg += a.b_hh[2 * H + u]; bo += a.b_hh[3 * H + u]; }
      const float gi = sigmoid_d(g[u] + v[0] + bi);
      const float gf = sigmoid_d(g[H + u] + v[1] + bf);
      const float gg = tanh_acc(g[2 * H + u] + o0 + bg);
      const float go = sigmoid_d(g[3 * H + u] + o1 + bo);
      const float cn = fmaf(gf, a.c_prev[(long long)b * H + u], gi * gg);
      const float hn = go * tanh_acc(cn);
      g[u] = gi; g[H + u] = gf; g[2 * H + u] = gg; g[3 * H + u] = go;
      a.c[(long long)b * H + u] = cn;
      a.h[(long long)b * a.h_ld + u] = hn;
      if (a.h2) a.h2[(long long)b * a.h2_ld + u] = hn;
    }
  }
}

size_t step_smem(int K) { return (size_t)kCols * ((K + 127) & ~127) * sizeof(float); }

// ---------------------------------------------------------------------------------------------------------------
// Backward product of the step: [d_ctx_t | d_h_{t-1}] = d_gates_t [W_ctx | W_hh]  (out[b][j] = sum_k dg[b][k] W[k][j]).
//   grid = X/8 CTAs, a CTA owns 8 output columns: its 8 columns of the weight (N x 8 floats, 41 KB at M2) are staged in
//   shared memory column-major BEFORE the dependency wait -- they are constants of the loop, so under a programmatic
//   dependent launch the staging overlaps the previous kernel's tail -- then warp w takes batch rows 4w..4w+3, the 32
//   lanes split N, d_gates rows are read straight from L2 two k-blocks ahead, and a 5-round transposing butterfly
//   leaves ONE output per lane.  Replaces cuBLAS split-K SGEMM + splitKreduce (7.1 + 3.2 us and two launch gaps).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kBwdCols = 8;

struct StepBwdArgs {
  const float* dg; long long dg_ld;          // [B,N] rows: d_gates_t
  const float* W; long long w_ld;            // [N,X] rows (X contiguous)
  int B, N, X;
  float* out; long long out_ld;              // [B,X]
};

__device__ __forceinline__ float4 ld_dg(const StepBwdArgs& a, int b, int k) {
  return (b < a.B && k < a.N) ? __ldcg(reinterpret_cast<const float4*>(a.dg + (long long)b * a.dg_ld + k))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
}

__global__ void __launch_bounds__(256) dec_step_bwd_kernel(const StepBwdArgs a) {
  extern __shared__ __align__(16) float W_s[];             // [8][Np + 4]: the +4 spreads a k-row's 8 columns over banks
  // (no early trigger here or in the cell backward: with the attention backward's 256 CTAs scheduled next to these
  // two kernels the backward walk took 19.6 ms per M2 batch instead of 13.4)
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int N = a.N, Np = (N + 127) & ~127, ld = Np + 4;
  const int j0 = blockIdx.x * kBwdCols;
  for (int idx = tid; idx < Np * kBwdCols; idx += 256) {
    const int k = idx >> 3, c = idx & 7;
    float* dst = W_s + c * ld + k;
    if (k < N && j0 + c < a.X) cp_async4(dst, a.W + (long long)k * a.w_ld + j0 + c);
    else *dst = 0.f;
  }
  cp_async_commit();
  pdl_wait();                             // d_gates comes from the cell-backward kernel before this one
  pdl_launch();
  float4 x0[kRowsPerWarp], x1[kRowsPerWarp];
#pragma unroll
  for (int i = 0; i < kRowsPerWarp; ++i) {
    x0[i] = ld_dg(a, kRowsPerWarp * w + i, 4 * lane);
    x1[i] = ld_dg(a, kRowsPerWarp * w + i, 128 + 4 * lane);
  }
  cp_async_wait_all();
  __syncthreads();
  for (int b0 = kRowsPerWarp * w; b0 < a.B; b0 += kRowsPerWarp * 8) {
    float acc[kRowsPerWarp][kBwdCols];
#pragma unroll
    for (int i = 0; i < kRowsPerWarp; ++i)
#pragma unroll
      for (int c = 0; c < kBwdCols; ++c) acc[i][c] = 0.f;
    if (b0 != kRowsPerWarp * w) {
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) {
        x0[i] = ld_dg(a, b0 + i, 4 * lane);
        x1[i] = ld_dg(a, b0 + i, 128 + 4 * lane);
      }
    }
    for (int kb = 0; kb < Np; kb += 128) {
      float4 x2[kRowsPerWarp];
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) x2[i] = ld_dg(a, b0 + i, kb + 256 + 4 * lane);
      const float* wp = W_s + kb + 4 * lane;
#pragma unroll
      for (int c = 0; c < kBwdCols; ++c) {
        const float4 wv = *reinterpret_cast<const float4*>(wp + c * ld);
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          acc[i][c] = fmaf(x0[i].x, wv.x, acc[i][c]);
          acc[i][c] = fmaf(x0[i].y, wv.y, acc[i][c]);
          acc[i][c] = fmaf(x0[i].z, wv.z, acc[i][c]);
          acc[i][c] = fmaf(x0[i].w, wv.w, acc[i][c]);
        }
      }
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) { x0[i] = x1[i]; x1[i] = x2[i]; }
    }
    float* v = &acc[0][0];                                  // v[i*8 + c]; after the butterfly lane L holds element L
#pragma unroll
    for (int m = 16, n = 16; m >= 1; m >>= 1, n >>= 1) {
      const bool hi = lane & m;
#pragma unroll
      for (int q = 0; q < n; ++q) {
        const float send = hi ? v[q] : v[q + n];
        const float keep = hi ? v[q + n] : v[q];
        v[q] = keep + __shfl_xor_sync(kFull, send, m);
      }
    }
    const int b = b0 + (lane >> 3), j = j0 + (lane & 7);
    if (b < a.B && j < a.X) a.out[(long long)b * a.out_ld + j] = v[0];
  }
}

size_t step_bwd_smem(int N) { return (size_t)kBwdCols * (((N + 127) & ~127) + 4) * sizeof(float); }

}  // namespace

// true when the fused kernel takes these shapes (whole units per CTA, 16-byte aligned rows)
bool dec_step_supported(int H, int K, long long x_ld, long long g_ld) {
  return H % 4 == 0 && K % 4 == 0 && x_ld % 4 == 0 && g_ld % 4 == 0 && step_smem(K) <= 200 * 1024;
}

// gates[:, t] (holding E_t) -> activations, c_t, h_t (+ copy) from x = [ctx_t | h_{t-1}] and W = [W_ctx | W_hh]
int dec_step_fwd(const float* x, long long x_ld, const float* w_cat, int B, int K, int H, float* gates, long long g_ld,
                 const float* b_ih, const float* b_hh, const float* c_prev, float* c, float* h, long long h_ld, float* h2,
                 long long h2_ld, cudaStream_t st) {
  StepArgs a = {};
  a.x = x; a.x_ld = x_ld; a.W = w_cat; a.w_ld = K; a.B = B; a.K = K; a.H = H;
  a.gates = gates; a.g_ld = g_ld; a.b_ih = b_ih; a.b_hh = b_hh; a.c_prev = c_prev; a.c = c;
  a.h = h; a.h_ld = h_ld; a.h2 = h2; a.h2_ld = h2_ld;
  const size_t smem = step_smem(K);
  static size_t configured = 0;
  if (smem > configured) {
    B2_CUDA_TRY(cudaFuncSetAttribute((const void*)dec_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(H / 4)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, dec_step_kernel, a));
  return B2ASR_OK;
}

bool dec_step_bwd_supported(int N, int X, long long dg_ld, long long w_ld) {
  return N % 4 == 0 && dg_ld % 4 == 0 && X > 0 && w_ld >= X && step_bwd_smem(N) <= 200 * 1024;
}

// out[B,X] = dg[B,N] W[N,X]
int dec_step_bwd(const float* dg, long long dg_ld, const float* W, long long w_ld, int B, int N, int X, float* out,
                 long long out_ld, cudaStream_t st) {
  StepBwdArgs a = {};
  a.dg = dg; a.dg_ld = dg_ld; a.W = W; a.w_ld = w_ld; a.B = B; a.N = N; a.X = X; a.out = out; a.out_ld = out_ld;
  const size_t smem = step_bwd_smem(N);
  static size_t configured = 0;
  if (smem > configured) {
    B2_CUDA_TRY(cudaFuncSetAttribute((const void*)dec_step_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((X + kBwdCols - 1) / kBwdCols)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, dec_step_bwd_kernel, a));
  return B2ASR_OK;
}

}  // namespace b2asr
