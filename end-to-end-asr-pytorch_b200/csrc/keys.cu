// Kernel group (2a): key projection  K = tanh(h . Wpsi^T + b)  on the tcgen05 tensor cores.
//
// Replaces `self.comp_listener_feature = torch.tanh(self.psi(listener_feature))` of the reference
// (src/asr.py:419), computed once per batch / utterance: a dense [M = B*T', D] x [D, A] contraction.
//
// fp32 parity at 1e-4 rules out a single TF32 pass (10-bit mantissa -> ~1e-3 on a D=640 dot product), so
// the product is error-compensated ("3xTF32"): every operand is split as x = hi + lo with hi = x with the
// mantissa rounded to TF32's 10 bits (exactly a TF32 number) and lo = x - hi (exact in fp32), and
//     a.b  ~=  hi_a.hi_b + hi_a.lo_b + lo_a.hi_b        (the dropped lo.lo term is ~2^-22 relative)
// is accumulated in fp32 in tensor memory by three tcgen05.mma.kind::tf32 per k-step.
//
// CTA = 128 rows x one N tile (<= 160 columns, multiple of 16); grid = (ceil(M/128), ceil(A/160)).
// Operand tiles are split into hi/lo on the way from global to shared memory (generic proxy stores into the
// canonical K-major no-swizzle UMMA layout, then fence.proxy.async); two stages, so the global loads of
// k-block i+1 overlap the MMAs of k-block i; tcgen05.commit on an mbarrier releases a stage.  Epilogue:
// tcgen05.ld (32 lanes x 16 columns per warp), + bias, accurate tanh, 64-byte stores per thread.
#include "common.cuh"

namespace b2asr {

constexpr int kBM = 128;          // rows per CTA == TMEM lanes == UMMA M
constexpr int kBK = 32;           // fp32 per k-block (8 chunks of 16 B); UMMA K for tf32 is 8
constexpr int kBNmax = 160;       // widest N tile (A=300 -> 160 + 144)
constexpr int kKeysNT = 128;      // 4 warps: warp w owns TMEM lanes 32w..32w+31 in the epilogue
constexpr int kTmemCols = 256;    // power of two >= kBNmax

struct KeysArgs {
  const float* h; const float* W; const float* bias; float* out;
  int M, D, A, BN;                // BN = columns of a full N tile (multiple of 16)
};

__host__ __device__ inline size_t keys_stage_bytes(int BN) { return (size_t)(kBM + BN) * kBK * 4 * 2; }   // A hi/lo + W hi/lo

// shared-memory matrix descriptor, K-major, SWIZZLE_NONE (canonical layout ((8,n),2):((16 B, SBO), LBO))
__device__ __forceinline__ unsigned long long umma_desc(unsigned saddr, unsigned lbo_bytes, unsigned sbo_bytes) {
  unsigned long long d = 0;
  d |= (unsigned long long)((saddr >> 4) & 0x3fffu);
  d |= (unsigned long long)((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= (unsigned long long)((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;                                   // descriptor version for sm_100
  return d;                                          // base offset 0, layout type 0 (no swizzle)
}

__device__ __forceinline__ void umma_tf32(unsigned d_tmem, unsigned long long adesc, unsigned long long bdesc,
                                          unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// x = hi + lo with hi = x rounded to the nearest TF32 (10 explicit mantissa bits; exact in the MMA) and
// lo = x - hi (exact in fp32, |lo| <= 2^-11 |x|) itself rounded to the nearest TF32, so the tensor core's
// truncation of the low 13 bits never bites: the operand error left is ~2^-23 relative.
__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void split4(const float4 v, float4& hi, float4& lo) {
  hi.x = tf32_rn(v.x); lo.x = tf32_rn(v.x - hi.x);
  hi.y = tf32_rn(v.y); lo.y = tf32_rn(v.y - hi.y);
  hi.z = tf32_rn(v.z); lo.z = tf32_rn(v.z - hi.z);
  hi.w = tf32_rn(v.w); lo.w = tf32_rn(v.w - hi.w);
}

// Load a [rows x kBK] fp32 tile (row stride ld, rows >= rmax or k >= kmax read as 0) and store its hi/lo
// split into the canonical layout: chunk j (4 floats) of row r at j*LBO + (r/8)*128 + (r%8)*16, LBO = rows*16.
__device__ __forceinline__ void load_split_tile(const float* g, long long ld, int row0, int rmax, int k0, int kmax,
                                                int rows, unsigned char* hi_base, unsigned char* lo_base) {
  const int items = rows * (kBK / 4);
  const unsigned lbo = (unsigned)rows * 16u;
  float4 v[4];
  for (int base = 0; base < items; base += 4 * kKeysNT) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = base + u * kKeysNT + threadIdx.x;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (idx < items) {
        const int r8 = idx & 7, ch = (idx >> 3) & 7, rg = idx >> 6;
        const int r = rg * 8 + r8, k = k0 + ch * 4;
        if (row0 + r < rmax && k < kmax) v[u] = __ldg(reinterpret_cast<const float4*>(g + (long long)(row0 + r) * ld + k));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = base + u * kKeysNT + threadIdx.x;
      if (idx < items) {
        const int r8 = idx & 7, ch = (idx >> 3) & 7, rg = idx >> 6;
        const unsigned off = (unsigned)ch * lbo + (unsigned)rg * 128u + (unsigned)r8 * 16u;
        float4 hi, lo;
        split4(v[u], hi, lo);
        *reinterpret_cast<float4*>(hi_base + off) = hi;
        *reinterpret_cast<float4*>(lo_base + off) = lo;
      }
    }
  }
}

__global__ void __launch_bounds__(kKeysNT, 1) attn_keys_kernel(KeysArgs p) {
  extern __shared__ __align__(128) unsigned char smk[];
  __shared__ __align__(8) unsigned long long bar_free[2];
  __shared__ __align__(8) unsigned long long bar_done;
  __shared__ unsigned tmem_base_sm;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * kBM;
  const int n0 = blockIdx.y * p.BN;
  const int bn = min(p.BN, (p.A - n0 + 15) / 16 * 16);          // this CTA's N tile (multiple of 16)
  const size_t stage = keys_stage_bytes(p.BN);
  const unsigned a_bytes = kBM * kBK * 4, w_bytes = (unsigned)p.BN * kBK * 4;

  if (tid == 0) {
    mbar_init(&bar_free[0], 1); mbar_init(&bar_free[1], 1); mbar_init(&bar_done, 1);
    mbar_fence_init();
  }
  if (warp == 0) {                                             // one warp allocates tensor memory
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sm)), "n"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = tmem_base_sm;

  // instruction descriptor: D fp32, A/B TF32, both K-major, N = bn, M = 128
  const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(bn >> 3) << 17) | ((unsigned)(kBM >> 4) << 24);
  const int nkb = (p.D + kBK - 1) / kBK;

  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb & 1;
    unsigned char* st = smk + (size_t)s * stage;
    unsigned char* a_hi = st, *a_lo = st + a_bytes, *w_hi = st + 2 * a_bytes, *w_lo = st + 2 * a_bytes + w_bytes;
    if (kb >= 2) mbar_wait(&bar_free[s], (unsigned)(((kb - 2) >> 1) & 1));   // MMAs of k-block kb-2 have read this stage
    load_split_tile(p.h, p.D, row0, p.M, kb * kBK, p.D, kBM, a_hi, a_lo);
    load_split_tile(p.W, p.D, n0, p.A, kb * kBK, p.D, bn, w_hi, w_lo);
    fence_proxy_async_smem();                                   // generic-proxy stores -> visible to the tensor core
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const unsigned lbo_a = kBM * 16u, lbo_w = (unsigned)bn * 16u;
#pragma unroll
      for (int kk = 0; kk < kBK / 8; ++kk) {
        const unsigned long long d_ahi = umma_desc(smem_u32(a_hi) + 2u * kk * lbo_a, lbo_a, 128u);
        const unsigned long long d_alo = umma_desc(smem_u32(a_lo) + 2u * kk * lbo_a, lbo_a, 128u);
        const unsigned long long d_whi = umma_desc(smem_u32(w_hi) + 2u * kk * lbo_w, lbo_w, 128u);
        const unsigned long long d_wlo = umma_desc(smem_u32(w_lo) + 2u * kk * lbo_w, lbo_w, 128u);
        umma_tf32(tmem, d_alo, d_whi, idesc, (kb | kk) ? 1u : 0u);     // small terms first
        umma_tf32(tmem, d_ahi, d_wlo, idesc, 1u);
        umma_tf32(tmem, d_ahi, d_whi, idesc, 1u);
      }
      umma_commit(&bar_free[s]);                                // arrives when the MMAs above have finished
    }
  }
  if (tid == 0) umma_commit(&bar_done);
  mbar_wait(&bar_done, 0u);
  tc_fence_after();

  // ---- epilogue: TMEM -> registers -> + bias -> tanh -> global ------------------------------------------
  const int row = row0 + warp * 32 + lane;
  for (int c0 = 0; c0 < bn; c0 += 16) {
    unsigned r[16];
    const unsigned taddr = tmem + ((unsigned)(warp * 32) << 16) + (unsigned)c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    if (row < p.M) {
      float* o = p.out + (long long)row * p.A + n0 + c0;
      const bool full = (n0 + c0 + 16 <= p.A) && ((p.A & 3) == 0);
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int col = n0 + c0 + j;
        const float b = col < p.A ? __ldg(p.bias + col) : 0.f;
        v[j] = tanh_acc(__uint_as_float(r[j]) + b);
      }
      if (full) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) if (n0 + c0 + j < p.A) o[j] = v[j];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
}

}  // namespace b2asr

using namespace b2asr;

extern "C" int b2asr_attn_keys(const float* h, const float* W, const float* bias, float* out,
                               int M, int D, int A, void* stream) {
  B2_REQUIRE(h && W && bias && out, B2ASR_EINVAL, "attn_keys: null pointer argument");
  B2_REQUIRE(M > 0 && D > 0 && A > 0, B2ASR_EINVAL, "attn_keys: bad sizes M=%d D=%d A=%d", M, D, A);
  B2_REQUIRE(D % 4 == 0 && (((uintptr_t)h | (uintptr_t)W) & 15) == 0, B2ASR_EUNSUPPORTED,
             "attn_keys: D=%d must be a multiple of 4 and h/W 16-byte aligned", D);
  KeysArgs a;
  a.h = h; a.W = W; a.bias = bias; a.out = out; a.M = M; a.D = D; a.A = A;
  const int ntiles = (A + kBNmax - 1) / kBNmax;
  a.BN = (((A + ntiles - 1) / ntiles) + 15) / 16 * 16;          // even split, multiple of 16 (A=300 -> 160)
  const size_t smem = 2 * keys_stage_bytes(a.BN);
  B2_CUDA_TRY(cudaFuncSetAttribute(attn_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)((M + kBM - 1) / kBM), (unsigned)ntiles);
  attn_keys_kernel<<<grid, kKeysNT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
