// Neural half of one joint beam-decode step as kernels, and the native decode loop.
//
// Reference src/asr.py:201-253: for EVERY live hypothesis, one after the other (batch 1): restore its decoder /
// attention / LM state from CPU clones, Attention.forward -> Speller (nn.LSTMCell on cat([embed(last), context])) ->
// char_trans -> log_softmax, RNN_LM.forward (src/rnnlm.py:19-25: emb -> LSTM stack -> Linear) -> log_softmax, CTC prefix
// scores, fusion, top-k.  Here the whole beam is the batch (rows = hypotheses) and one decode step is
//
//   b2asr_attn_step_fwd_rows   attention, rows read their PARENT's state through the survivor map of the last step
//   cell_step_kernel           one launch per LSTM(Cell) layer of the Speller and of the LM: gate pre-activations
//                              base[token] + [x1 | x2] W^T (+ biases) and the cell, for all rows.  `base` is the
//                              embedding half of the layer folded into a [V, 4H] table once per model (the row gather
//                              replaces embed() and its product), x2 / c_prev rows come through the survivor map.
//   linear_lse_kernel          output layer (char_trans / LM out) for all rows + the rows' log-sum-exp, so that
//                              log_softmax is `logit - lse` inside the beam step kernel; partial (max, sum) per CTA,
//                              combined in a fixed order by the last CTA to finish (bit-reproducible).
//   b2asr_beam_step            kernel (3), beam.cu.  Its parent_out is the next step's survivor map.
//
// b2asr_decode_run enqueues `steps` decode steps from two such per-parity programs (the state is double buffered:
// step t reads buffers t&1 and writes the others) with no host work in between.  All kernels but the beam step are
// launched with the programmatic-dependent-launch attribute and stage their weights before the dependency wait.
#include "common.cuh"
#include <atomic>
#include <mutex>

namespace b2asr {

bool pdl_enabled();                           // attn_fwd.cu

namespace {

constexpr int kRows = 4;                      // batch rows per warp
constexpr int kUnits = 4;                     // hidden units per CTA (16 gate rows)
constexpr int kCols = 16;                     // output columns per CTA of the linear kernel

__device__ __forceinline__ float sigmoid_d(float v) { return __fdividef(1.f, 1.f + ex2(-v * kLog2e)); }

// x row b = [x1[b] (K1) | x2[map[b]] (K2)]; K1, K2 and both row strides are multiples of 4
__device__ __forceinline__ float4 ld_x(const b2asr_cell_step& a, int b, long long bsrc, int k) {
  if (b >= a.B) return make_float4(0.f, 0.f, 0.f, 0.f);
  if (k < a.K1) return __ldcg(reinterpret_cast<const float4*>(a.x1 + (long long)b * a.x1_ld + k));
  if (k < a.K1 + a.K2) return __ldcg(reinterpret_cast<const float4*>(a.x2 + bsrc * a.x2_ld + (k - a.K1)));
  return make_float4(0.f, 0.f, 0.f, 0.f);
}

// column c of the CTA is gate row  g*H + 4*blockIdx.x + uu  with c = uu*4 + g  (torch gate order i, f, g, o)
__global__ void __launch_bounds__(256) cell_step_kernel(const b2asr_cell_step a) {
  extern __shared__ __align__(16) float W_s[];             // [16][Kp]
  pdl_trigger_small();                    // the next kernel of the step may start staging ITS weights right away
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int K = a.K1 + a.K2, Kp = (K + 127) & ~127, H = a.H;
  for (int idx = tid; idx < kCols * (Kp >> 2); idx += 256) {
    const int c = idx / (Kp >> 2), k = (idx - c * (Kp >> 2)) << 2;
    const int u = kUnits * blockIdx.x + (c >> 2);
    float* dst = W_s + c * Kp + k;
    if (k < K && u < H) cp_async16(dst, a.W + ((long long)(c & 3) * H + u) * a.w_ld + k);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cp_async_commit();
  pdl_wait();                             // the weights above are constants; everything below depends on earlier kernels
  cp_async_wait_all();
  __syncthreads();

  for (int b0 = kRows * w; b0 < a.B; b0 += kRows * 8) {
    long long bsrc[kRows];
    float4 x0[kRows], x1[kRows];
#pragma unroll
    for (int i = 0; i < kRows; ++i) {
      bsrc[i] = (b0 + i < a.B) ? (a.map ? a.map[b0 + i] : (long long)(b0 + i)) : 0ll;
      x0[i] = ld_x(a, b0 + i, bsrc[i], 4 * lane);
      x1[i] = ld_x(a, b0 + i, bsrc[i], 128 + 4 * lane);
    }
    float acc[kRows][kCols];
#pragma unroll
    for (int i = 0; i < kRows; ++i)
#pragma unroll
      for (int c = 0; c < kCols; ++c) acc[i][c] = 0.f;
    for (int kb = 0; kb < Kp; kb += 128) {
      float4 x2[kRows];
#pragma unroll
      for (int i = 0; i < kRows; ++i) x2[i] = ld_x(a, b0 + i, bsrc[i], kb + 256 + 4 * lane);
      const float* wp = W_s + kb + 4 * lane;
#pragma unroll
      for (int c = 0; c < kCols; ++c) {
        const float4 wv = *reinterpret_cast<const float4*>(wp + c * Kp);
#pragma unroll
        for (int i = 0; i < kRows; ++i) {
          acc[i][c] = fmaf(x0[i].x, wv.x, acc[i][c]);
          acc[i][c] = fmaf(x0[i].y, wv.y, acc[i][c]);
          acc[i][c] = fmaf(x0[i].z, wv.z, acc[i][c]);
          acc[i][c] = fmaf(x0[i].w, wv.w, acc[i][c]);
        }
      }
#pragma unroll
      for (int i = 0; i < kRows; ++i) { x0[i] = x1[i]; x1[i] = x2[i]; }
    }
    // transposing butterfly over the 32 k-lanes: 64 sums -> 2 per lane
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
    const int u = kUnits * blockIdx.x + ((lane >> 1) & 3);
    if ((lane & 1) == 0 && b < a.B && u < H) {
      const long long bs = a.map ? a.map[b] : (long long)b;
      float gi = v[0], gf = v[1], gg = o0, go = o1;
      if (a.base) {
        const float* br = a.base + (a.base_row ? a.base_row[b] : (long long)b) * a.base_ld;
        gi += br[u]; gf += br[H + u]; gg += br[2 * H + u]; go += br[3 * H + u];
      }
      if (a.b1) { gi += a.b1[u]; gf += a.b1[H + u]; gg += a.b1[2 * H + u]; go += a.b1[3 * H + u]; }
      if (a.b2) { gi += a.b2[u]; gf += a.b2[H + u]; gg += a.b2[2 * H + u]; go += a.b2[3 * H + u]; }
      gi = sigmoid_d(gi); gf = sigmoid_d(gf); gg = tanh_acc(gg); go = sigmoid_d(go);
      const float cn = fmaf(gf, a.c_prev[bs * a.c_ld + u], gi * gg);
      a.c[(long long)b * a.c_ld + u] = cn;
      a.h[(long long)b * a.h_ld + u] = go * tanh_acc(cn);
    }
  }
}

// logits[b][v] = x[b] . W[v] + bias[v]; CTA = 16 consecutive v for all rows; partial (max, sum exp) per (row, CTA)
__global__ void __launch_bounds__(256) linear_lse_kernel(const b2asr_linear_lse a) {
  extern __shared__ __align__(16) float W_s[];             // [16][Kp]
  __shared__ unsigned is_last;
  pdl_trigger_small();
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int K = a.K, Kp = (K + 127) & ~127, V = a.V;
  const int v0 = blockIdx.x * kCols, ncta = gridDim.x;
  for (int idx = tid; idx < kCols * (Kp >> 2); idx += 256) {
    const int c = idx / (Kp >> 2), k = (idx - c * (Kp >> 2)) << 2;
    float* dst = W_s + c * Kp + k;
    if (k < K && v0 + c < V) cp_async16(dst, a.W + (long long)(v0 + c) * a.w_ld + k);
    else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  cp_async_commit();
  pdl_wait();
  cp_async_wait_all();
  __syncthreads();

  auto ldx = [&](int b, int k) {
    return (b < a.B && k < K) ? __ldcg(reinterpret_cast<const float4*>(a.x + (long long)b * a.x_ld + k))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  for (int b0 = kRows * w; b0 < a.B; b0 += kRows * 8) {
    float4 x0[kRows], x1[kRows];
#pragma unroll
    for (int i = 0; i < kRows; ++i) { x0[i] = ldx(b0 + i, 4 * lane); x1[i] = ldx(b0 + i, 128 + 4 * lane); }
    float acc[kRows][kCols];
#pragma unroll
    for (int i = 0; i < kRows; ++i)
#pragma unroll
      for (int c = 0; c < kCols; ++c) acc[i][c] = 0.f;
    for (int kb = 0; kb < Kp; kb += 128) {
      float4 x2[kRows];
#pragma unroll
      for (int i = 0; i < kRows; ++i) x2[i] = ldx(b0 + i, kb + 256 + 4 * lane);
      const float* wp = W_s + kb + 4 * lane;
#pragma unroll
      for (int c = 0; c < kCols; ++c) {
        const float4 wv = *reinterpret_cast<const float4*>(wp + c * Kp);
#pragma unroll
        for (int i = 0; i < kRows; ++i) {
          acc[i][c] = fmaf(x0[i].x, wv.x, acc[i][c]);
          acc[i][c] = fmaf(x0[i].y, wv.y, acc[i][c]);
          acc[i][c] = fmaf(x0[i].z, wv.z, acc[i][c]);
          acc[i][c] = fmaf(x0[i].w, wv.w, acc[i][c]);
        }
      }
#pragma unroll
      for (int i = 0; i < kRows; ++i) { x0[i] = x1[i]; x1[i] = x2[i]; }
    }
    float* v = &acc[0][0];
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
    // lane L holds row i = L>>3 and columns c = 2*(L&7), 2*(L&7)+1
    const int i = lane >> 3, b = b0 + i, c0 = 2 * (lane & 7);
    float l0 = -INFINITY, l1 = -INFINITY;
    if (b < a.B) {
      if (v0 + c0 < V) { l0 = v[0] + (a.bias ? a.bias[v0 + c0] : 0.f); a.logits[(long long)b * a.out_ld + v0 + c0] = l0; }
      if (v0 + c0 + 1 < V) { l1 = v[1] + (a.bias ? a.bias[v0 + c0 + 1] : 0.f); a.logits[(long long)b * a.out_ld + v0 + c0 + 1] = l1; }
    }
    float m = fmaxf(l0, l1);
    m = fmaxf(m, __shfl_xor_sync(kFull, m, 1));
    m = fmaxf(m, __shfl_xor_sync(kFull, m, 2));
    m = fmaxf(m, __shfl_xor_sync(kFull, m, 4));
    float s = (l0 > -INFINITY ? ex2((l0 - m) * kLog2e) : 0.f) + (l1 > -INFINITY ? ex2((l1 - m) * kLog2e) : 0.f);
    s += __shfl_xor_sync(kFull, s, 1);
    s += __shfl_xor_sync(kFull, s, 2);
    s += __shfl_xor_sync(kFull, s, 4);
    if ((lane & 7) == 0 && b < a.B) {
      a.part[((long long)b * ncta + blockIdx.x) * 2] = m;
      a.part[((long long)b * ncta + blockIdx.x) * 2 + 1] = s;
    }
  }
  // ---- the last CTA to finish combines the partials of every row, in CTA order ---------------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = (atomicAdd(a.ticket, 1u) == (unsigned)ncta - 1u);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  constexpr int kPairs = 12;                  // (max, sum) pairs a lane holds per row: ncta <= 384, i.e. V <= 6144
  if (ncta <= 32 * kPairs && a.B <= 32) {
    // every load of the warp's (up to four) rows is issued before the first is used: the combine is ONE round trip
    // to L2 instead of two per row (it sat at the end of both chains of a decode step).  Same operations in the same
    // order as the loop below, so the result is bit-identical.
    float2 pv[4][kPairs];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int b = w + 8 * r;
      const float2* pr = reinterpret_cast<const float2*>(a.part) + (long long)min(b, a.B - 1) * ncta;
#pragma unroll
      for (int q = 0; q < kPairs; ++q) {
        const int j = lane + 32 * q;
        pv[r][q] = (b < a.B && j < ncta) ? __ldcg(pr + j) : make_float2(-INFINITY, 0.f);
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int b = w + 8 * r;
      if (b >= a.B) break;
      float m = -INFINITY;
#pragma unroll
      for (int q = 0; q < kPairs; ++q) if (lane + 32 * q < ncta) m = fmaxf(m, pv[r][q].x);
      m = warp_max(m);
      float sum = 0.f;
#pragma unroll
      for (int q = 0; q < kPairs; ++q) if (lane + 32 * q < ncta) sum += pv[r][q].y * ex2((pv[r][q].x - m) * kLog2e);
      sum = warp_sum(sum);
      if (lane == 0) a.lse[b] = m + logf(sum);
    }
  } else {
    for (int b = w; b < a.B; b += 8) {
      const float* pr = a.part + (long long)b * ncta * 2;
      float m = -INFINITY;
      for (int j = lane; j < ncta; j += 32) m = fmaxf(m, __ldcg(pr + 2 * j));
      m = warp_max(m);
      float s = 0.f;
      for (int j = lane; j < ncta; j += 32) s += __ldcg(pr + 2 * j + 1) * ex2((__ldcg(pr + 2 * j) - m) * kLog2e);
      s = warp_sum(s);
      if (lane == 0) a.lse[b] = m + logf(s);
    }
  }
  if (tid == 0) *a.ticket = 0u;                                   // re-armed for the next step
}

size_t tile_smem(int K) { return (size_t)16 * ((K + 127) & ~127) * sizeof(float); }

// The LM's chain of a step (its layers and output layer) depends only on the hypotheses' last tokens and on its own
// state, not on the attention / Speller chain: it runs on a second stream next to it.  One stream + fork / join events
// per (device, caller stream), created on first use (host-side driver objects).
struct LmStream { cudaStream_t stream; cudaEvent_t fork, join; int dev; cudaStream_t owner; };
// one side stream per (device, caller stream): utterances decoded concurrently on different streams of one GPU
// (beam.beam_decode_many) must not queue their LM chains behind each other
constexpr int kMaxLmStreams = 64;
LmStream g_lm[kMaxLmStreams];
int g_lm_n = 0;
std::mutex g_lm_mu;

int lm_stream(LmStream** out, cudaStream_t owner) {
  int dev = -1;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_lm_mu);
  for (int i = 0; i < g_lm_n; ++i)
    if (g_lm[i].dev == dev && g_lm[i].owner == owner) { *out = &g_lm[i]; return B2ASR_OK; }
  B2_REQUIRE(g_lm_n < kMaxLmStreams, B2ASR_EUNSUPPORTED, "decode: more than %d (device, stream) pairs in use", kMaxLmStreams);
  LmStream& s = g_lm[g_lm_n];
  B2_CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
  B2_CUDA_TRY(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
  B2_CUDA_TRY(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  s.dev = dev; s.owner = owner;
  ++g_lm_n;
  *out = &s;
  return B2ASR_OK;
}

template <typename Args>
int launch_tile(void (*kern)(const Args), const Args& a, unsigned grid, int K, cudaStream_t st) {
  const size_t smem = tile_smem(K);
  B2_REQUIRE(smem <= 200 * 1024, B2ASR_EUNSUPPORTED, "decode: inner dimension %d too wide for the weight tile", K);
  B2_CUDA_TRY(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a));
  return B2ASR_OK;
}

}  // namespace
}  // namespace b2asr

using namespace b2asr;

extern "C" size_t b2asr_cell_step_size(void) { return sizeof(b2asr_cell_step); }
extern "C" size_t b2asr_linear_lse_size(void) { return sizeof(b2asr_linear_lse); }
extern "C" size_t b2asr_decode_prog_size(void) { return sizeof(b2asr_decode_prog); }

extern "C" int b2asr_cell_step_fwd(const b2asr_cell_step* a, void* stream) {
  B2_REQUIRE(a, B2ASR_EINVAL, "cell_step: null argument block");
  B2_REQUIRE(a->B > 0 && a->H > 0 && a->K1 >= 0 && a->K2 >= 0 && a->K1 + a->K2 > 0, B2ASR_EINVAL, "cell_step: bad sizes");
  B2_REQUIRE(a->K1 % 4 == 0 && a->K2 % 4 == 0 && a->x1_ld % 4 == 0 && a->x2_ld % 4 == 0 && a->w_ld % 4 == 0, B2ASR_EINVAL,
             "cell_step: K1, K2 and the row strides must be multiples of 4 (pad the operands)");
  B2_REQUIRE((a->K1 == 0 || a->x1) && (a->K2 == 0 || a->x2) && a->W && a->c_prev && a->c && a->h, B2ASR_EINVAL,
             "cell_step: null operand");
  B2_REQUIRE(((((uintptr_t)a->x1) | ((uintptr_t)a->x2) | ((uintptr_t)a->W)) & 15) == 0, B2ASR_EINVAL,
             "cell_step: operands must be 16-byte aligned");
  return launch_tile(cell_step_kernel, *a, (unsigned)((a->H + kUnits - 1) / kUnits), a->K1 + a->K2, (cudaStream_t)stream);
}

extern "C" int b2asr_linear_lse_fwd(const b2asr_linear_lse* a, void* stream) {
  B2_REQUIRE(a, B2ASR_EINVAL, "linear_lse: null argument block");
  B2_REQUIRE(a->B > 0 && a->K > 0 && a->V > 0, B2ASR_EINVAL, "linear_lse: bad sizes");
  B2_REQUIRE(a->K % 4 == 0 && a->x_ld % 4 == 0 && a->w_ld % 4 == 0, B2ASR_EINVAL,
             "linear_lse: K and the row strides must be multiples of 4 (pad the operands)");
  B2_REQUIRE(a->x && a->W && a->logits && a->lse && a->part && a->ticket, B2ASR_EINVAL, "linear_lse: null operand");
  B2_REQUIRE(((((uintptr_t)a->x) | ((uintptr_t)a->W)) & 15) == 0, B2ASR_EINVAL, "linear_lse: operands must be 16-byte aligned");
  return launch_tile(linear_lse_kernel, *a, (unsigned)((a->V + kCols - 1) / kCols), a->K, (cudaStream_t)stream);
}

extern "C" int b2asr_linear_lse_ctas(int V) { return (V + kCols - 1) / kCols; }

extern "C" int b2asr_decode_run(const b2asr_decode_prog* prog, int first_step, int steps, void* stream) {
  B2_REQUIRE(prog && steps >= 0 && first_step >= 0, B2ASR_EINVAL, "decode_run: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  static const bool split = [] { const char* e = getenv("B2ASR_DECODE_LM_STREAM"); return !(e && e[0] == '0'); }();
  LmStream* lm = nullptr;
  const bool has_lm = prog[0].n_linear > 1;
  if (has_lm && split) {
    if (int rc = lm_stream(&lm, st)) return rc;
  }
  for (int t = first_step; t < first_step + steps; ++t) {
    const b2asr_decode_prog& p = prog[t & 1];
    B2_REQUIRE(p.n_cells >= 0 && p.n_cells <= 8 && p.n_linear >= 0 && p.n_linear <= 2 && p.beam, B2ASR_EINVAL,
               "decode_run: malformed program");
    // chain B (the LM's layers and output layer: cells[i].after_linear) forks off before the attention kernel
    int first_lm = 0;
    while (first_lm < p.n_cells && !p.cells[first_lm].after_linear) ++first_lm;
    void* stb = stream;
    if (lm) {
      B2_CUDA_TRY(cudaEventRecord(lm->fork, st));
      B2_CUDA_TRY(cudaStreamWaitEvent(lm->stream, lm->fork, 0));
      stb = (void*)lm->stream;
      for (int ci = first_lm; ci < p.n_cells; ++ci)
        if (int rc = b2asr_cell_step_fwd(&p.cells[ci], stb)) return rc;
      if (int rc = b2asr_linear_lse_fwd(&p.linear[1], stb)) return rc;
      B2_CUDA_TRY(cudaEventRecord(lm->join, lm->stream));
    }
    // chain A: attention, Speller layers, char_trans
    const b2asr_attn_call& q = p.attn;
    if (int rc = b2asr_attn_step_fwd_rows(q.mode, q.s, q.s_stride, q.Wphi, q.keys, q.k_bstride, q.h, q.h_bstride, q.lens,
                                          q.prev_att, q.row_map, q.Wc, q.C, q.Kw, q.Wl, q.wg, q.bg, q.scale, q.att, q.ctx,
                                          q.ctx_stride, nullptr, q.B, q.T, q.D, q.A, q.dec, q.cluster, stream))
      return rc;
    for (int ci = 0; ci < first_lm; ++ci)
      if (int rc = b2asr_cell_step_fwd(&p.cells[ci], stream)) return rc;
    if (p.n_linear > 0)
      if (int rc = b2asr_linear_lse_fwd(&p.linear[0], stream)) return rc;
    if (lm) {
      B2_CUDA_TRY(cudaStreamWaitEvent(st, lm->join, 0));
    } else {
      for (int ci = first_lm; ci < p.n_cells; ++ci)
        if (int rc = b2asr_cell_step_fwd(&p.cells[ci], stream)) return rc;
      if (p.n_linear > 1)
        if (int rc = b2asr_linear_lse_fwd(&p.linear[1], stream)) return rc;
    }
    if (int rc = b2asr_beam_step(p.beam, stream)) return rc;
  }
  return B2ASR_OK;
}
