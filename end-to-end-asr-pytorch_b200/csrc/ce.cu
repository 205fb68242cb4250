// Attention-branch sequence loss, SURVEY.md section 8(f) rank 3: reference src/solver.py:90,149-155
//     seq_loss = CrossEntropyLoss(ignore_index=0, reduction='none')
//     att_loss = seq_loss(att_pred.view(b*t,c), label.view(-1))
//     att_loss = torch.sum(att_loss.view(b,t),dim=-1) / torch.sum(y!=0,dim=-1).float();  att_loss = torch.mean(att_loss)
// plus the argmax that cal_acc / cal_cer take of the same logits (src/postprocess.py:121-146).
//
// The library path is four passes over the [B*L,V] logits (log_softmax, nll, and their two backward kernels) and
// a fifth for the argmax.  Here: ONE streaming launch, one read of the logits and one write of the finished gradient row
//     g[n,v] = prescale / (len_b * B) * (softmax(x_n)_v - [v == label_n])        (zero row when label_n == ignore)
// with the row held in shared memory between the max/argmax pass, the sum pass and the gradient pass; per-row
// losses go to loss_rows, and a one-CTA second launch folds them into the scalar exactly in the reference's order
// (sum over t, divide by the utterance's length, mean over utterances), in a fixed summation order so the scalar
// is bit-reproducible (and the streaming kernel carries no fence or ticket).
// HBM-bound at V=5000: 2 x 4 x V bytes per row.
#include "common.cuh"

namespace b2asr {

struct CeArgs {
  const float* x; long long sx;          // logits [N,V], row stride
  const long long* labels;               // [N]
  long long N; int V;
  int R;                                 // rows per group (decoder steps per utterance)
  const long long* glen;                 // [G] divisor per group (target lengths), nullptr -> 1
  int G;
  long long ignore;
  float prescale;
  float* loss_rows;                      // [N] out
  long long* argmax;                     // [N] out, nullable
  float* grad; long long sg;             // [N,V] out, nullable
  float* loss;                           // [1] out, nullable
};

template <int TPR>
__global__ void __launch_bounds__(256) ce_rows_kernel(CeArgs p, int Vpad) {
  extern __shared__ __align__(16) float sm[];
  constexpr int RPC = 256 / TPR;
  __shared__ float red[2][8];
  __shared__ int redi[8];
  const int g = threadIdx.x / TPR, l = threadIdx.x % TPR;
  const long long row = (long long)blockIdx.x * RPC + g;
  const bool in_range = row < p.N;                          // uniform per warp (TPR >= 32)
  const int V = p.V;

  if (in_range) {
    const long long label = p.labels[row];
    const bool counted = label != p.ignore;
    const int grp = (int)(row / p.R);
    float* cache = sm + (size_t)g * Vpad;
    const float* xr = p.x + row * p.sx;
    float* gr = p.grad ? p.grad + row * p.sg : nullptr;
    const bool vec_in = (V % 4 == 0) && (p.sx % 4 == 0) && (((uintptr_t)p.x & 15) == 0);
    const bool vec_out = gr && (V % 4 == 0) && (p.sg % 4 == 0) && (((uintptr_t)p.grad & 15) == 0);

    // ---- pass 1: row -> shared memory, running max and its first index ------------------------------------
    float m = -INFINITY;
    int mi = 0x7fffffff;
    if (vec_in) {
      for (int i = l; i < V / 4; i += TPR) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(xr) + i);
        reinterpret_cast<float4*>(cache)[i] = v;
        if (v.x > m) { m = v.x; mi = 4 * i; }
        if (v.y > m) { m = v.y; mi = 4 * i + 1; }
        if (v.z > m) { m = v.z; mi = 4 * i + 2; }
        if (v.w > m) { m = v.w; mi = 4 * i + 3; }
      }
    } else {
      for (int i = l; i < V; i += TPR) { const float v = __ldcs(xr + i); cache[i] = v; if (v > m) { m = v; mi = i; } }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {                      // larger value wins, equal values: lower index
      const float om = __shfl_xor_sync(kFull, m, o);
      const int oi = __shfl_xor_sync(kFull, mi, o);
      if (om > m || (om == m && oi < mi)) { m = om; mi = oi; }
    }
    if (TPR == 256) {
      if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = m; redi[threadIdx.x >> 5] = mi; }
      __syncthreads();
      m = red[0][0]; mi = redi[0];
#pragma unroll
      for (int w = 1; w < 8; ++w) {
        const float om = red[0][w];
        const int oi = redi[w];
        if (om > m || (om == m && oi < mi)) { m = om; mi = oi; }
      }
    } else {
      __syncwarp();
    }
    if (mi == 0x7fffffff) mi = 0;
    const float mm = (m == -INFINITY) ? 0.f : m;
    float s = 0.f;
    for (int i = l; i < V; i += TPR) s += ex2((cache[i] - mm) * kLog2e);
    s = warp_sum(s);
    if (TPR == 256) {
      if ((threadIdx.x & 31) == 0) red[1][threadIdx.x >> 5] = s;
      __syncthreads();
      s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[1][w];
    }
    const float lse = mm + logf(s);
    const int lc = (int)min((long long)V - 1, max(0ll, label));
    if (l == 0) {
      p.loss_rows[row] = counted ? lse - cache[lc] : 0.f;
      if (p.argmax) p.argmax[row] = mi;
    }

    // ---- pass 2: finished gradient row ----------------------------------------------------------------------
    if (gr) {
      float gs = 0.f;
      if (counted) gs = p.prescale / ((p.glen ? (float)p.glen[grp] : 1.f) * (float)p.G);
      const float off = lse * kLog2e;
      if (vec_out) {
        for (int i = l; i < V / 4; i += TPR) {
          float4 v = reinterpret_cast<const float4*>(cache)[i];
          v.x = ex2(fmaf(v.x, kLog2e, -off)); v.y = ex2(fmaf(v.y, kLog2e, -off));
          v.z = ex2(fmaf(v.z, kLog2e, -off)); v.w = ex2(fmaf(v.w, kLog2e, -off));
          const int d = lc - 4 * i;
          if (d == 0) v.x -= 1.f; else if (d == 1) v.y -= 1.f; else if (d == 2) v.z -= 1.f; else if (d == 3) v.w -= 1.f;
          v.x *= gs; v.y *= gs; v.z *= gs; v.w *= gs;
          reinterpret_cast<float4*>(gr)[i] = v;
        }
      } else {
        for (int i = l; i < V; i += TPR) gr[i] = (ex2(fmaf(cache[i], kLog2e, -off)) - (i == lc ? 1.f : 0.f)) * gs;
      }
    }
  }
}

// mean over groups of (sum over the group's rows) / len; one CTA of 32 warps, fixed order.  The loads of a group are
// issued four deep: this launch is a pure latency chain (G*R/32 dependent L2 round trips per warp otherwise).
__global__ void __launch_bounds__(1024) ce_scalar_kernel(CeArgs p) {
  __shared__ float red[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc = 0.f;
  for (int q = warp; q < p.G; q += 32) {
    const long long r0 = (long long)q * p.R;
    const int n = (int)min((long long)p.R, p.N - r0);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int t = lane; t < n; t += 128) {
      const float v0 = p.loss_rows[r0 + t];
      const float v1 = t + 32 < n ? p.loss_rows[r0 + t + 32] : 0.f;
      const float v2 = t + 64 < n ? p.loss_rows[r0 + t + 64] : 0.f;
      const float v3 = t + 96 < n ? p.loss_rows[r0 + t + 96] : 0.f;
      s0 += v0; s1 += v1; s2 += v2; s3 += v3;
    }
    const float gsum = warp_sum((s0 + s1) + (s2 + s3));
    acc += gsum / (p.glen ? (float)p.glen[q] : 1.f);
  }
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 32; ++w) tot += red[w];
    *p.loss = tot / (float)p.G;
  }
}

}  // namespace b2asr

using namespace b2asr;

extern "C" int b2asr_ce_loss_fwd_bwd(const float* logits, int64_t stride_row, const int64_t* labels, int64_t N, int V,
                                     int rows_per_group, const int64_t* group_len, int64_t ignore_index, float grad_prescale,
                                     float* loss_rows, int64_t* argmax, float* grad, int64_t grad_stride_row,
                                     float* loss, void* stream) {
  B2_REQUIRE(logits && labels && loss_rows, B2ASR_EINVAL, "ce_loss: null pointer argument");
  B2_REQUIRE(N > 0 && V > 0 && rows_per_group > 0, B2ASR_EINVAL, "ce_loss: bad sizes N=%lld V=%d R=%d", (long long)N, V, rows_per_group);
  B2_REQUIRE(stride_row >= V && (!grad || grad_stride_row >= V), B2ASR_EINVAL, "ce_loss: row stride smaller than V");
  CeArgs a;
  a.x = logits; a.sx = stride_row; a.labels = (const long long*)labels; a.N = N; a.V = V; a.R = rows_per_group;
  a.glen = (const long long*)group_len; a.G = (int)((N + rows_per_group - 1) / rows_per_group);
  a.ignore = ignore_index; a.prescale = grad_prescale; a.loss_rows = loss_rows; a.argmax = (long long*)argmax;
  a.grad = grad; a.sg = grad_stride_row; a.loss = loss;
  const int Vpad = (V + 3) & ~3;
  cudaStream_t st = (cudaStream_t)stream;
  if (V <= 1024) {
    const size_t smem = (size_t)8 * Vpad * sizeof(float);
    ce_rows_kernel<32><<<(unsigned)((N + 7) / 8), 256, smem, st>>>(a, Vpad);
  } else {
    const size_t smem = (size_t)Vpad * sizeof(float);
    B2_REQUIRE(smem <= 227 * 1024, B2ASR_EINVAL, "ce_loss: V=%d does not fit a shared-memory row", V);
    if (smem > 48 * 1024)
      B2_CUDA_TRY(cudaFuncSetAttribute(ce_rows_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ce_rows_kernel<256><<<(unsigned)N, 256, smem, st>>>(a, Vpad);
  }
  if (loss) ce_scalar_kernel<<<1, 1024, 0, st>>>(a);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
