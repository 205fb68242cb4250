// Weight-gradient products of the Listener (reference src/asr.py:466-501: what autograd computes for nn.LSTM's
// weight_ih / weight_hh):   c[Ma,Nb] = sum_k a[k, 0:Ma]^T b[k, 0:Nb],   k over the Kr = B*T rows of two activation
// matrices -- the reduction index is the SLOW index of both operands ("MN-major").  Same error-compensated TF32
// pipeline as keys2.cu (see there for the roles and the measurements), with the two changes this operand form needs:
//
//   a   raw 32(k) x 128(m) tiles arrive as four unswizzled 32 x 32 TMA boxes; a converter thread owns output row m and
//       reads its COLUMN of the tile (lanes read consecutive words: conflict-free), so the transposition into the
//       tensor-memory operand (lane = m, column = k) is free.  Rows with k % period == period - 1 are zeroed here:
//       dW_hh = sum_t dG_{t+1}^T h_t pairs row r of one matrix with row r + 1 of the other (a pointer offset), and the
//       pairs that straddle two utterances must not count.
//   b   32(k) x 64(n) tiles as two TMA boxes in the one layout the tensor core accepts for MN-major tf32 operands
//       ("128-byte swizzle, 32-byte base": 4 k-rows x 128 bytes per swizzle atom, atoms 512 bytes apart, 32-column
//       groups one box = 4096 bytes apart); lo is written behind hi at the same offsets, so [b.hi ; b.lo] is one
//       N = 128 operand again.
//
// Three 64-column accumulators (hi.hi | hi.lo | lo.hi) + a 5-slot ring of the converted a tile = 512 tensor-memory
// columns.  Split-K: work item = (k-slice, tile); slice s writes its partial product to c + s*Ma*Nb and the caller adds
// the slices (deterministic).  `max_ctas` caps the persistent grid: next to a 96-SM recurrence kernel the launch must
// not take more SMs than are free, or the recurrence's clusters wait for the GEMM's CTAs to finish.
#include "umma.cuh"

namespace b2asr {

namespace {

constexpr int kBM = 128, kBN = 64, kBK = 32;
constexpr int kST = 6;                                             // shared-memory ring
constexpr int kTA = 5;                                             // tensor-memory ring of the converted a tile
constexpr int kChunk = 32;                                         // k-blocks accumulated in tensor memory at a stretch
constexpr int kConvGroups = 3, kConvWarps = 4 * kConvGroups, kEpiWarps = 8;
constexpr int kNT = (2 + kConvWarps + kEpiWarps) * 32;             // 704 threads
constexpr unsigned kBoxBytes = 32 * 32 * 4;                        // one 32 x 32 fp32 box
constexpr unsigned kABytes = 4 * kBoxBytes, kBBytes = 2 * kBoxBytes;   // 16384, 8192
constexpr unsigned kStageBytes = kABytes + 2 * kBBytes;            // 32768
constexpr unsigned kTmemA = 3 * kBN, kTmemCols = 512;
static_assert(kTmemA + kTA * 2 * kBK <= kTmemCols, "tensor memory budget");
static_assert(kST > kTA, "the converters' wait for a free tensor-memory slot relies on the deeper shared-memory ring");

struct NtArgs {
  float* out;                     // [S][Ma][Nb]
  int Ma, Nb, n_mtiles, n_ntiles;
  int nkb, nkb_per, S;            // k-blocks in all, per slice, slices
  int period;                     // 0, or: rows with k % period == period - 1 do not count
};

// MN-major tf32 operands have ONE legal shared-memory layout: "128-byte swizzle, 32-byte base" (layout type 1; TMA's
// SWIZZLE_128B_ATOM_32B): k-rows of 128 bytes (32 columns), atoms of 4 k-rows whose 32-byte chunks are XOR-ed with the
// row number; next atom of 4 k-rows at + 512 (SBO), next 32 columns at + 4096 (LBO = one TMA box)
__device__ __forceinline__ unsigned long long desc_mn_sw128_32b(unsigned saddr) {
  unsigned long long d = 0;
  d |= (unsigned long long)((saddr >> 4) & 0x3fffu);
  d |= (unsigned long long)(4096u >> 4) << 16;
  d |= (unsigned long long)(512u >> 4) << 32;
  d |= 1ull << 46;
  d |= 1ull << 61;
  return d;
}

__global__ void __launch_bounds__(kNT, 1) gemm_nt_kernel(const __grid_constant__ CUtensorMap map_a,
                                                         const __grid_constant__ CUtensorMap map_b, NtArgs p) {
  extern __shared__ unsigned char smraw[];
  __shared__ __align__(8) unsigned long long full[kST], conv[kST], empty[kST], acc_full, acc_empty;
  __shared__ unsigned tmem_base_sm;
  unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smraw) + 1023) & ~(uintptr_t)1023);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = p.n_mtiles * p.n_ntiles, nitems = ntiles * p.S;

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

  // item -> (slice, tile); the k-blocks of slice s are [s * nkb_per, min(nkb, (s + 1) * nkb_per))
#define NT_ITEM(item)                                                                 \
  const int ks = (item) / ntiles, tile = (item) - ks * ntiles;                        \
  const int m0 = (tile / p.n_ntiles) * kBM, n0 = (tile % p.n_ntiles) * kBN;           \
  const int kb0 = ks * p.nkb_per, kb1 = min(p.nkb, kb0 + p.nkb_per);                  \
  (void)m0; (void)n0

  if (warp == 0) {
    // ---------------- producer ----------------
    unsigned it = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      NT_ITEM(item);
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const unsigned s = it % kST, ph = (it / kST) & 1u;
        mbar_wait_sleep(&empty[s], ph ^ 1u);
        unsigned char* st = ring + (size_t)s * kStageBytes;
        if (elect_one()) {
          mbar_expect_tx(&full[s], kABytes + kBBytes);
#pragma unroll
          for (int i = 0; i < 4; ++i) tma_2d(st + i * kBoxBytes, &map_a, m0 + 32 * i, kb * kBK, &full[s]);
#pragma unroll
          for (int i = 0; i < 2; ++i) tma_2d(st + kABytes + i * kBoxBytes, &map_b, n0 + 32 * i, kb * kBK, &full[s]);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    // D fp32, A/B TF32, A K-major (tensor memory), B MN-major (bit 16), M = 128
    const unsigned idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 16) | ((unsigned)(kBM >> 4) << 24);
    const unsigned idesc_n = idesc_base | ((unsigned)(kBN >> 3) << 17), idesc_2n = idesc_base | ((unsigned)(2 * kBN >> 3) << 17);
    unsigned it = 0, j = 0;                                      // j: chunks handed to the epilogue so far
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      NT_ITEM(item);
      const unsigned d_main = tmem, d_lh = d_main + 2 * kBN;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int kc = (kb - kb0) % kChunk;                        // position in the chunk
        if (kc == 0) {
          mbar_wait_sleep(&acc_empty, (j & 1u) ^ 1u);              // the epilogue has taken the previous chunk's sums
          tc_fence_after();
        }
        const unsigned s = it % kST, ph = (it / kST) & 1u;
        mbar_wait_sleep(&conv[s], ph);
        tc_fence_after();
        const unsigned a_hi = tmem + kTmemA + (it % kTA) * 2 * kBK, a_lo = a_hi + kBK;
        const unsigned b_hi = smem_u32(ring + (size_t)s * kStageBytes) + kABytes;   // b lo follows at + kBBytes
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBK / 8; ++kk) {
            const unsigned off = (unsigned)kk * 1024u;             // next group of 8 k-rows
            const unsigned acc = (kc || kk) ? 1u : 0u;
            umma_tf32_ts(d_lh, a_lo + kk * 8, desc_mn_sw128_32b(b_hi + off), idesc_n, acc);       // lo.hi         -> 128..191
            umma_tf32_ts(d_main, a_hi + kk * 8, desc_mn_sw128_32b(b_hi + off), idesc_2n, acc);    // hi.hi | hi.lo -> 0..127
          }
          umma_commit(&empty[s]);
          if (kc == kChunk - 1 || kb == kb1 - 1) umma_commit(&acc_full);   // chunk complete
        }
        __syncwarp();
        if (kc == kChunk - 1 || kb == kb1 - 1) ++j;
      }
    }
  } else if (warp < 2 + kConvWarps) {
    // ---------------- converters ----------------
    const int cw = warp - 2, grp = cw >> 2, ct = (cw & 3) * 32 + lane;
    const unsigned lane_addr = (unsigned)((warp & 3) * 32) << 16;
    unsigned it = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      NT_ITEM(item);
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        if ((int)(it % kConvGroups) != grp) continue;
        const unsigned s = it % kST, ph = (it / kST) & 1u;
        if (it >= (unsigned)kTA) {
          const unsigned u = it - kTA;
          mbar_wait_sleep(&empty[u % kST], (u / kST) & 1u);
          tc_fence_after();
        }
        mbar_wait_sleep(&full[s], ph);
        // this thread's column of the a tile: box (warp % 4), word `lane` of every 128-byte k-row
        const float* a_col = reinterpret_cast<const float*>(ring + (size_t)s * kStageBytes) + (warp & 3) * 1024 + lane;
        const float4* b_hi = reinterpret_cast<const float4*>(ring + (size_t)s * kStageBytes + kABytes);
        float4* b_lo = const_cast<float4*>(b_hi) + kBBytes / 16;
        // the one k of this block (if any) whose row does not count: kmask in [0, 32), or -1
        int kmask = -1;
        if (p.period > 0) {
          const int r = (kb * kBK) % p.period;                     // position of the block's first row in its utterance
          const int d = p.period - 1 - r;                          // rows until the excluded one
          kmask = d < kBK ? d : -1;                                // (period >= 32 is required by the host side)
        }
        const unsigned ta = tmem + lane_addr + kTmemA + (it % kTA) * 2 * kBK;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          unsigned hi[16], lo[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            float v = a_col[(half * 16 + k) * 32];
            if (half * 16 + k == kmask) v = 0.f;
            hi[k] = __float_as_uint(v);
            lo[k] = __float_as_uint(lo_half(v));
          }
          tmem_st16(ta + (unsigned)half * 16u, hi);
          tmem_st16(ta + kBK + (unsigned)half * 16u, lo);
        }
#pragma unroll
        for (int i = 0; i < (int)(kBBytes / 16) / 128; ++i) {
          const float4 v = b_hi[i * 128 + ct];
          b_lo[i * 128 + ct] = make_float4(lo_half(v.x), lo_half(v.y), lo_half(v.z), lo_half(v.w));
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&conv[s]);
      }
    }
  } else {
    // ---------------- epilogue ----------------
    const int lq = warp & 3, ch = (warp - 2 - kConvWarps) >> 2;
    constexpr int kEC = kBN / 2;                                   // 32 columns per thread
    unsigned j = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      NT_ITEM(item);
      // The tensor core's fp32 additions truncate, a bias of ~2^-24 of the accumulator per k-step: sums stay in tensor
      // memory for kChunk k-blocks (128 steps) only and are then added, rounded to nearest, into these registers.
      float r[kEC];
#pragma unroll
      for (int i = 0; i < kEC; ++i) r[i] = 0.f;
      const unsigned tbase = tmem + ((unsigned)(lq * 32) << 16) + (unsigned)(ch * kEC);
      for (int kc0 = kb0; kc0 < kb1; kc0 += kChunk, ++j) {
        mbar_wait_sleep(&acc_full, j & 1u);
        tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < kEC; c0 += 16) {
          unsigned m[16], q[16], u[16];
          tmem_ld16(tbase + c0, m);
          tmem_ld16(tbase + c0 + kBN, q);
          tmem_ld16(tbase + c0 + 2 * kBN, u);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int i = 0; i < 16; ++i) r[c0 + i] += __uint_as_float(m[i]) + (__uint_as_float(q[i]) + __uint_as_float(u[i]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cta(&acc_empty);
      }
      const int row = m0 + lq * 32 + lane, c_first = n0 + ch * kEC;
      if (row < p.Ma) {
        float* o = p.out + ((long long)ks * p.Ma + row) * p.Nb + c_first;
        const bool vec = (p.Nb & 3) == 0;
#pragma unroll
        for (int c0 = 0; c0 < kEC; c0 += 4) {
          if (vec && c_first + c0 + 4 <= p.Nb) *reinterpret_cast<float4*>(o + c0) = make_float4(r[c0], r[c0 + 1], r[c0 + 2], r[c0 + 3]);
          else {
#pragma unroll
            for (int q = 0; q < 4; ++q) if (c_first + c0 + q < p.Nb) o[c0 + q] = r[c0 + q];
          }
        }
      }
    }
  }
#undef NT_ITEM
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
}

}  // namespace

}  // namespace b2asr

extern "C" int b2asr_gemm_nt_3xtf32_slices(int Ma, int Nb, long long Kr, int max_ctas) {
  using namespace b2asr;
  if (Ma <= 0 || Nb <= 0 || Kr <= 0) return 1;
  const long long tiles = (long long)((Ma + kBM - 1) / kBM) * ((Nb + kBN - 1) / kBN);
  const long long nkb = (Kr + kBK - 1) / kBK;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int ctas = max_ctas > 0 && max_ctas < sms ? max_ctas : sms;
  // the persistent grid works through tiles * S items in rounds; an item costs its k-blocks plus ~12 k-blocks' worth of
  // pipeline fill, accumulator hand-over and epilogue: take the S with the cheapest last round (<= 32 slices, >= 16
  // k-blocks each)
  long long S = 1, best = -1;
  for (long long s = 1; s <= 32 && (s == 1 || nkb / s >= 16); ++s) {
    const long long per = (nkb + s - 1) / s;
    if ((s - 1) * per >= nkb) continue;                            // would leave an empty slice
    const long long rounds = (tiles * s + ctas - 1) / ctas;
    const long long cost = rounds * (per + 12);
    if (best < 0 || cost < best) { best = cost; S = s; }
  }
  return (int)S;
}

extern "C" int b2asr_gemm_nt_3xtf32(const float* a, long long pitch_a, const float* b, long long pitch_b, float* c, int Ma,
                                    int Nb, long long Kr, int period, int slices, int max_ctas, void* stream) {
  using namespace b2asr;
  B2_REQUIRE(a && b && c, B2ASR_EINVAL, "b2asr_gemm_nt_3xtf32: null pointer");
  B2_REQUIRE(Ma > 0 && Nb > 0 && Kr > 0 && slices >= 1, B2ASR_EINVAL, "b2asr_gemm_nt_3xtf32: Ma=%d Nb=%d Kr=%lld slices=%d", Ma, Nb, Kr, slices);
  // tensor copies want 16-byte aligned bases and row pitches; the row mask handles one excluded row per k-block
  if ((pitch_a & 3) || (pitch_b & 3) || ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) ||
      (period != 0 && period < kBK) || Kr > 0x7fffffffLL)
    return B2ASR_EUNSUPPORTED;
  CUtensorMap map_a, map_b;
  if (int rc = make_map(&map_a, a, Kr, Ma, 32, pitch_a, kSwzNone)) return rc;
  if (int rc = make_map(&map_b, b, Kr, Nb, 32, pitch_b, kSwz128Atom32)) return rc;
  NtArgs p;
  p.out = c; p.Ma = Ma; p.Nb = Nb;
  p.n_mtiles = (Ma + kBM - 1) / kBM;
  p.n_ntiles = (Nb + kBN - 1) / kBN;
  p.nkb = (int)((Kr + kBK - 1) / kBK);
  p.S = slices;
  p.nkb_per = (p.nkb + slices - 1) / slices;
  B2_REQUIRE((long long)(slices - 1) * p.nkb_per < p.nkb, B2ASR_EINVAL, "b2asr_gemm_nt_3xtf32: %d slices leave empty ones (%d k-blocks)", slices, p.nkb);
  p.period = period;
  const size_t smem = (size_t)kST * kStageBytes + 1024;
  static bool configured[64] = {};
  int dev = 0, sms = 0;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B2_CUDA_TRY(cudaFuncSetAttribute(gemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev & 63] = true;
  }
  B2_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long items = (long long)p.n_mtiles * p.n_ntiles * slices;
  long long grid = items < sms ? items : sms;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  gemm_nt_kernel<<<(int)grid, kNT, smem, (cudaStream_t)stream>>>(map_a, map_b, p);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
