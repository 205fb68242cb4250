// Kernel group (2a), second version: key projection  K = tanh(h . Wpsi^T + b)  (reference src/asr.py:419) as a
// warp-specialised, persistent tcgen05 pipeline.  One CTA per SM walks 128 x 80 output tiles (M2: 400 of them).
//
//   producer warp    one elected lane issues 2-D TMA tensor copies (cp.async.bulk.tensor, SASS UTMALDG) of the RAW fp32
//                    operand tiles -- 128 rows of h and 80 rows of W, 32 columns (128 bytes) each, 128-byte hardware
//                    swizzle -- into a 6-deep shared-memory ring; rows / columns outside the tensors arrive as zeros
//   converter warps  (3 groups of 4, taking the k-blocks in turn) split the operands for the error-compensated product:
//                    lo = x - trunc_tf32(x) (exact in fp32).  The h tile goes row-per-thread from shared memory into
//                    TENSOR memory (tcgen05.st, hi = the raw bits, lo) -- a 4-deep ring of 64 columns; the W tile's lo is
//                    written next to the raw W tile at the same (position-only) swizzled offsets.  kind::tf32 ignores
//                    the low 13 mantissa bits, so the raw words ARE the high halves.
//   MMA warp         one elected lane issues, per 8-wide k-step, TWO tcgen05.mma.kind::tf32 with the A operand from
//                    tensor memory:  a.lo x W.hi (N = 80) and a.hi x [W.hi ; W.lo] (N = 160: the two W tiles are
//                    contiguous 8-row groups, so one descriptor runs through both).  Three accumulators of 80 columns
//                    (hi.hi | hi.lo | lo.hi): the big one sees a third of the tensor core's truncating additions.
//                    tcgen05.commit hands ring stages back to the producer / converters and the tile to the epilogue.
//   epilogue warps   (8: two per tensor-memory lane quarter) tcgen05.ld their 40 columns of the three accumulators into
//                    registers, hand the accumulators straight back (~300 clk per tile, so one set is enough and the
//                    columns buy the deeper h ring), then + bias, tanh, 16-byte stores while the next tile's MMAs run.
// No CTA-wide barrier inside the loop: every hand-off is an mbarrier.
//
// What the measurements said on the way (tools/keys_variants.sh + keys_time.py / keys_trace.py, M2 shape, B200):
//   * issuing tcgen05.mma / TMA under `if (lane == 0)` makes the compiler wrap each one in an election loop (~55 clk per
//     instruction); under elect.sync in a converged warp they issue back to back from uniform registers;
//   * one tcgen05.mma (M = 128, K = 8, tf32) costs ~36 + 0.29 N clk whether A comes from shared or tensor memory, so the
//     pipe is paced by the instruction count: 3 x N=80 per k-step -> 177 clk, N=80 + N=160 -> 141 clk;
//   * the converters are bound by shared-memory bandwidth (per k-block 26 KB in by TMA, 26 KB read, 10 KB written, 30 KB
//     read by the MMAs' B operands), not by ALU work: rounding lo to nearest TF32 or a third converter group changes little.
//   v1 (keys.cu) 94.5 us -> 30.5 us per launch; cuBLAS TF32 (one pass, 1.4e-3 error) + tanh kernel: 31 us.
//
// a.b ~= hi_a.hi_b + hi_a.lo_b + lo_a.hi_b, fp32 accumulation in tensor memory: max abs error against float64 stays
// at fp32 level (tests/test_keys_gpu.py: <= 5e-6 on O(1) outputs; cuBLAS fp32 2.4e-6, single-pass TF32 1.5e-3).
#include "umma.cuh"

namespace b2asr {

namespace {

#ifndef K2_STAGES
#define K2_STAGES 6
#endif
constexpr int kBM = 128, kBN = 80, kBK = 32;
constexpr int kST = K2_STAGES;                                     // shared-memory ring (raw h tile, W tile hi + lo)
constexpr int kTA = 4;                                             // tensor-memory ring of the converted h tile (hi + lo)
#ifndef K2_GROUPS
#define K2_GROUPS 3
#endif
constexpr int kConvGroups = K2_GROUPS, kConvWarps = 4 * kConvGroups, kEpiWarps = 8;
constexpr int kNT = (2 + kConvWarps + kEpiWarps) * 32;             // 704 threads
constexpr unsigned kABytes = kBM * kBK * 4, kWBytes = kBN * kBK * 4;  // 16384, 10240 (multiples of 1024)
constexpr unsigned kStageBytes = kABytes + 2 * kWBytes;            // 36864
static_assert((kWBytes / 16) % 128 == 0, "W tile: whole float4 rounds per converter group");
// tensor memory, 512 columns: accumulators {main, small} at 0 / 80, then kTA x {hi 32, lo 32}.  One accumulator pair is
// enough: the epilogue drains it into registers in ~300 clk per tile and hands it back before its tanh / store work,
// and the columns buy a deeper ring for the h tile, which is what keeps the tensor pipe fed.
constexpr unsigned kTmemA = 3 * kBN, kTmemCols = 512;
static_assert(kTmemA + kTA * 2 * kBK <= kTmemCols, "tensor memory budget");
static_assert(kST > kTA, "the converters' wait for a free tensor-memory slot relies on the deeper shared-memory ring");

struct Keys2Args {
  const float* bias; float* out;
  int M, D, A, n_mtiles, n_ntiles;
  long long* trace;               // -DK2_TRACE builds: per-role clock64 sums of CTA 0 (tools/keys_trace.py)
};
#ifdef K2_TRACE
__device__ long long g_keys_trace[64];
#define K2_T0() long long t_ = clock64()
#define K2_ACC(slot) do { const long long n_ = clock64(); if (blockIdx.x == 0) tr[slot] += n_ - t_; t_ = n_; } while (0)
#else
#define K2_T0() do {} while (0)
#define K2_ACC(slot) do {} while (0)
#endif

// TANH: epilogue tanh (the key projection); otherwise the plain product + optional bias (Listener projections)
template <bool TANH>
__global__ void __launch_bounds__(kNT, 1) attn_keys2_kernel(const __grid_constant__ CUtensorMap map_h,
                                                            const __grid_constant__ CUtensorMap map_w, Keys2Args p) {
  extern __shared__ unsigned char smraw[];
  __shared__ __align__(8) unsigned long long full[kST], conv[kST], empty[kST], acc_full, acc_empty;
  __shared__ unsigned tmem_base_sm;
  __shared__ __align__(16) float bias_sm[kEpiWarps][kBN / 2];                   // each epilogue warp's 40 bias values, zero past A
  unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smraw) + 1023) & ~(uintptr_t)1023);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = p.n_mtiles * p.n_ntiles;
  const int nkb = (p.D + kBK - 1) / kBK;

  if (tid == 0) {
    for (int s = 0; s < kST; ++s) { mbar_init(&full[s], 1); mbar_init(&conv[s], 4); mbar_init(&empty[s], 1); }
    mbar_init(&acc_full, 1); mbar_init(&acc_empty, kEpiWarps);
    mbar_fence_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sm)), "n"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = tmem_base_sm;

  if (warp == 0) {
    // ---------------- producer ----------------
    {
      unsigned it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int m0 = (t / p.n_ntiles) * kBM, n0 = (t % p.n_ntiles) * kBN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const unsigned s = it % kST, ph = (it / kST) & 1u;
          mbar_wait_sleep(&empty[s], ph ^ 1u);                     // the MMAs that read this stage have finished
          unsigned char* st = ring + (size_t)s * kStageBytes;
          if (elect_one()) {
#ifdef K2_NOTMA
            (void)st; (void)m0; (void)n0;
            mbar_arrive_cta(&full[s]);
#else
            mbar_expect_tx(&full[s], kABytes + kWBytes);
            tma_2d(st, &map_h, kb * kBK, m0, &full[s]);            // h tile, raw
            tma_2d(st + kABytes, &map_w, kb * kBK, n0, &full[s]);  // W tile, raw = its high half
#endif
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    {
      // instruction descriptor: D fp32, A/B TF32, both K-major, N = 80, M = 128
      // instruction descriptors: D fp32, A/B TF32, both K-major, M = 128; N = 160 runs through W hi AND W lo, which lie
      // back to back in the stage (one 8-row group every 1024 bytes): a.hi x [W.hi ; W.lo] is ONE instruction.  A
      // tcgen05.mma of this shape costs ~36 + 0.29 N clk (measured), so fewer, wider instructions are what counts.
      const unsigned idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(kBM >> 4) << 24);
      const unsigned idesc80 = idesc_base | ((unsigned)(kBN >> 3) << 17), idesc160 = idesc_base | ((unsigned)(2 * kBN >> 3) << 17);
      unsigned it = 0, j = 0;
#ifdef K2_TRACE
      long long tr[8] = {};
#endif
      K2_T0();
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++j) {
        mbar_wait_sleep(&acc_empty, (j & 1u) ^ 1u);                // the epilogue has drained the accumulators
        tc_fence_after();
        K2_ACC(0);
        const unsigned d_main = tmem, d_lh = d_main + 2 * kBN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const unsigned s = it % kST, ph = (it / kST) & 1u;
          mbar_wait_sleep(&conv[s], ph);                           // h tile split into tensor memory, W low half written
          tc_fence_after();
          K2_ACC(1);
          const unsigned a_hi = tmem + kTmemA + (it % kTA) * 2 * kBK, a_lo = a_hi + kBK;
          const unsigned w_hi = smem_u32(ring + (size_t)s * kStageBytes) + kABytes;   // W lo follows at + kWBytes
          if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBK / 8; ++kk) {
            const unsigned off = (unsigned)kk * 32u;               // 8 tf32 = 32 bytes inside the 128-byte swizzle atom
            const unsigned acc = (kb | kk) ? 1u : 0u;
            umma_tf32_ts(d_lh, a_lo + kk * 8, desc_sw128(w_hi + off), idesc80, acc);      // lo.hi          -> columns 160..239
            umma_tf32_ts(d_main, a_hi + kk * 8, desc_sw128(w_hi + off), idesc160, acc);   // hi.hi | hi.lo  -> columns 0..159
          }
          K2_ACC(2);
          umma_commit(&empty[s]);                                  // shared-memory stage and tensor-memory slot free
          K2_ACC(3);
          }
          __syncwarp();
        }
        if (elect_one()) umma_commit(&acc_full);                   // accumulators complete
        __syncwarp();
      }
#ifdef K2_TRACE
      if (blockIdx.x == 0 && lane == 0) for (int i = 0; i < 4; ++i) p.trace[i] = tr[i];
#endif
    }
  } else if (warp < 2 + kConvWarps) {
    // ---------------- converters ----------------
    // kConvGroups groups of four warps take the k-blocks in turn, so that the latency chain of one block (shared-memory
    // loads -> split -> tcgen05.st -> wait -> fences -> arrive) overlaps the next block's.  h tile: thread = one row (the
    // tensor-memory lane its warp may write: 32 (warp % 4) + lane); the row's 16-byte chunks sit at chunk ^ (row % 8)
    // of its 128-byte line (the copy's swizzle): conflict-free reads.
    const int cw = warp - 2, grp = cw >> 2, ct = (cw & 3) * 32 + lane;   // ct: 0..127 inside the group
    const int row = (warp & 3) * 32 + lane;
    const unsigned lane_addr = (unsigned)((warp & 3) * 32) << 16;
    unsigned it = 0;
#ifdef K2_TRACE
    long long tr[8] = {};
#endif
    K2_T0();
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        if ((int)(it % kConvGroups) != grp) continue;
        const unsigned s = it % kST, ph = (it / kST) & 1u;
        if (it >= (unsigned)kTA) {                                 // tensor-memory slot it % kTA: read by the MMAs of it - kTA
          const unsigned u = it - kTA;
          mbar_wait_sleep(&empty[u % kST], (u / kST) & 1u);
          tc_fence_after();
        }
        K2_ACC(0);
        mbar_wait_sleep(&full[s], ph);
        K2_ACC(1);
        const float4* a_raw = reinterpret_cast<const float4*>(ring + (size_t)s * kStageBytes);
        const float4* w_hi = a_raw + kABytes / 16;
        float4* w_lo = const_cast<float4*>(w_hi) + kWBytes / 16;
#ifndef K2_NOCONV
        const unsigned ta = tmem + lane_addr + kTmemA + (it % kTA) * 2 * kBK;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          unsigned hi[16], lo[16];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float4 v = a_raw[row * 8 + ((half * 4 + c) ^ (row & 7))];
            hi[4 * c + 0] = __float_as_uint(v.x); hi[4 * c + 1] = __float_as_uint(v.y);
            hi[4 * c + 2] = __float_as_uint(v.z); hi[4 * c + 3] = __float_as_uint(v.w);
            lo[4 * c + 0] = __float_as_uint(lo_half(v.x)); lo[4 * c + 1] = __float_as_uint(lo_half(v.y));
            lo[4 * c + 2] = __float_as_uint(lo_half(v.z)); lo[4 * c + 3] = __float_as_uint(lo_half(v.w));
          }
          tmem_st16(ta + (unsigned)half * 16u, hi);
          tmem_st16(ta + kBK + (unsigned)half * 16u, lo);
        }
#pragma unroll
        for (int i = 0; i < (int)(kWBytes / 16) / 128; ++i) {
          const float4 v = w_hi[i * 128 + ct];
          w_lo[i * 128 + ct] = make_float4(lo_half(v.x), lo_half(v.y), lo_half(v.z), lo_half(v.w));
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
#endif
        fence_proxy_async_smem();                                  // generic-proxy stores -> visible to the tensor core
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&conv[s]);
        K2_ACC(2);
      }
    }
#ifdef K2_TRACE
    if (blockIdx.x == 0 && cw == 0 && lane == 0) for (int i = 0; i < 3; ++i) p.trace[8 + i] = tr[i];
#endif
  } else {
    // ---------------- epilogue ----------------
    // two warps per tensor-memory lane quarter, 40 of the tile's 80 columns each
    const int lq = warp & 3, ch = (warp - 2 - kConvWarps) >> 2;
    constexpr int kEC = kBN / 2;                                   // 40 columns per thread
    unsigned j = 0;
#ifdef K2_TRACE
    long long tr[8] = {};
#endif
    K2_T0();
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++j) {
      const int m0 = (t / p.n_ntiles) * kBM, n0 = (t % p.n_ntiles) * kBN + ch * kEC;
      float* bias_w = bias_sm[warp - 2 - kConvWarps];
      __syncwarp();
      for (int i = lane; i < kEC; i += 32) bias_w[i] = (p.bias && n0 + i < p.A) ? __ldg(p.bias + n0 + i) : 0.f;
      __syncwarp();
      mbar_wait_sleep(&acc_full, j & 1u);
      tc_fence_after();
      K2_ACC(0);
      float r[kEC];
      const unsigned tbase = tmem + ((unsigned)(lq * 32) << 16) + (unsigned)(ch * kEC);
#pragma unroll
      for (int c0 = 0; c0 < 32; c0 += 16) {
        unsigned m[16], q[16], u[16];
        tmem_ld16(tbase + c0, m);
        tmem_ld16(tbase + c0 + kBN, q);
        tmem_ld16(tbase + c0 + 2 * kBN, u);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 16; ++i) r[c0 + i] = __uint_as_float(m[i]) + (__uint_as_float(q[i]) + __uint_as_float(u[i]));
      }
      {
        unsigned m[8], q[8], u[8];
        tmem_ld8(tbase + 32, m);
        tmem_ld8(tbase + 32 + kBN, q);
        tmem_ld8(tbase + 32 + 2 * kBN, u);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 8; ++i) r[32 + i] = __uint_as_float(m[i]) + (__uint_as_float(q[i]) + __uint_as_float(u[i]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cta(&acc_empty);                  // the MMA warp may overwrite the accumulators
      K2_ACC(1);
      const int row = m0 + lq * 32 + lane;
#ifdef K2_NOEPI
      if (row < p.M && r[0] == 123.456f) {
#else
      if (row < p.M) {
#endif
        float* o = p.out + (long long)row * p.A + n0;
        const bool vec = (p.A & 3) == 0;
#pragma unroll
        for (int c0 = 0; c0 < kEC; c0 += 4) {
          float v[4];
          const float4 bq = *reinterpret_cast<const float4*>(bias_w + c0);
          v[0] = r[c0 + 0] + bq.x; v[1] = r[c0 + 1] + bq.y; v[2] = r[c0 + 2] + bq.z; v[3] = r[c0 + 3] + bq.w;
          if (TANH) { v[0] = tanh_fast(v[0]); v[1] = tanh_fast(v[1]); v[2] = tanh_fast(v[2]); v[3] = tanh_fast(v[3]); }
          if (vec && n0 + c0 + 4 <= p.A) *reinterpret_cast<float4*>(o + c0) = make_float4(v[0], v[1], v[2], v[3]);
          else {
#pragma unroll
            for (int q = 0; q < 4; ++q) if (n0 + c0 + q < p.A) o[c0 + q] = v[q];
          }
        }
      }
      K2_ACC(2);
    }
#ifdef K2_TRACE
    if (blockIdx.x == 0 && warp == 2 + kConvWarps && lane == 0) for (int i = 0; i < 3; ++i) p.trace[16 + i] = tr[i];
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
}

}  // namespace

// out[M,A] = act(h[M,D] . W[A,D]^T + bias); -> B2ASR_OK, or an error; B2ASR_EUNSUPPORTED = operands the tensor copies
// cannot address (the caller takes another route)
int gemm3x_tn(const float* h, const float* W, const float* bias, float* out, int M, int D, int A, bool tanh_act,
              cudaStream_t st) {
  // tensor copies want 16-byte aligned bases and row pitches
  if ((D & 3) || ((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(W)) & 15)) return B2ASR_EUNSUPPORTED;
  CUtensorMap map_h, map_w;
  if (int rc = make_map(&map_h, h, M, D, kBM)) return rc;
  if (int rc = make_map(&map_w, W, A, D, kBN)) return rc;
  Keys2Args a;
  a.bias = bias; a.out = out; a.M = M; a.D = D; a.A = A;
  a.n_mtiles = (M + kBM - 1) / kBM;
  a.n_ntiles = (A + kBN - 1) / kBN;
  a.trace = nullptr;
#ifdef K2_TRACE
  B2_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&a.trace), g_keys_trace));
#endif
  const size_t smem = (size_t)kST * kStageBytes + 1024;
  static bool configured[64] = {};
  int dev = 0, sms = 0;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B2_CUDA_TRY(cudaFuncSetAttribute(attn_keys2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B2_CUDA_TRY(cudaFuncSetAttribute(attn_keys2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  B2_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long tiles = (long long)a.n_mtiles * a.n_ntiles;
  const int grid = (int)(tiles < sms ? tiles : sms);
  if (tanh_act) attn_keys2_kernel<true><<<grid, kNT, smem, st>>>(map_h, map_w, a);
  else attn_keys2_kernel<false><<<grid, kNT, smem, st>>>(map_h, map_w, a);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

int attn_keys_v2(const float* h, const float* W, const float* bias, float* out, int M, int D, int A, cudaStream_t st) {
  return gemm3x_tn(h, W, bias, out, M, D, A, true, st);
}

}  // namespace b2asr

#ifdef K2_TRACE
extern "C" int b2asr_keys_trace(long long* host64) {
  return cudaMemcpyFromSymbol(host64, b2asr::g_keys_trace, 64 * sizeof(long long)) == cudaSuccess ? 0 : -1;
}
#endif

// Plain product on the same pipeline: c[M,N] = a[M,K] . b[N,K]^T (+ bias[N]); see include/b2asr.h
extern "C" int b2asr_gemm_tn_3xtf32(const float* a, const float* b, const float* bias, float* c, int M, int N, int K,
                                    void* stream) {
  using namespace b2asr;
  B2_REQUIRE(a && b && c, B2ASR_EINVAL, "b2asr_gemm_tn_3xtf32: null pointer");
  B2_REQUIRE(M > 0 && N > 0 && K > 0, B2ASR_EINVAL, "b2asr_gemm_tn_3xtf32: M=%d N=%d K=%d", M, N, K);
  return gemm3x_tn(a, b, bias, c, M, K, N, false, (cudaStream_t)stream);
}
