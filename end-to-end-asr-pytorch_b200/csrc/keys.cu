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
// canonical K-major no-swizzle UMMA layout, then fence.proxy.async); three 37 KB stages and register double
// buffering of the global loads, two CTAs per SM; tcgen05.commit on an mbarrier releases a stage.  Epilogue:
// tcgen05.ld (32 lanes x 16 columns per warp), + bias, accurate tanh, 64-byte stores per thread.
#include <stdlib.h>

#include "common.cuh"

namespace b2asr {

constexpr int kBM = 128;          // rows per CTA == TMEM lanes == UMMA M
constexpr int kBK = 16;           // fp32 per k-block (4 chunks of 16 B); UMMA K for tf32 is 8
constexpr int kStages = 3;        // 3 x 36.9 KB: two CTAs fit one SM (and 2 x 256 TMEM columns = all 512)
constexpr int kBNmax = 160;       // widest N tile (A=300 -> 160 + 144)
constexpr int kKeysNT = 256;      // 8 warps stage operands; in the epilogue warp w owns TMEM lanes 32(w%4).. and half of the columns
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

// One k-block of both operands in registers: kBK/4 = 4 chunks (float4) per row, rows*4 items spread over the CTA.
// A tile: 128 rows -> 4 items per thread; W tile: <= 160 rows -> <= 5 items per thread.
constexpr int kItemsA = kBM * (kBK / 4) / kKeysNT;                 // 4
constexpr int kItemsW = (kBNmax * (kBK / 4) + kKeysNT - 1) / kKeysNT;   // 5
struct KTile { float4 a[kItemsA]; float4 w[kItemsW]; };

// item idx of a tile with `rows` rows: row r = (idx>>5)*8 + (idx&7), chunk ch = (idx>>3)&3; a quarter-warp reads one
// 16-byte chunk of 8 consecutive rows, and the same quarter-warp later writes one 128-byte core matrix.
__device__ __forceinline__ void load_ktile(KTile& t, const KeysArgs& p, int row0, int n0, int bn, int k0) {
#pragma unroll
  for (int u = 0; u < kItemsA; ++u) {
    const int idx = u * kKeysNT + threadIdx.x;
    const int r = (idx >> 5) * 8 + (idx & 7), k = k0 + ((idx >> 3) & 3) * 4;
    t.a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < p.M && k < p.D) t.a[u] = __ldg(reinterpret_cast<const float4*>(p.h + (long long)(row0 + r) * p.D + k));
  }
#pragma unroll
  for (int u = 0; u < kItemsW; ++u) {
    const int idx = u * kKeysNT + threadIdx.x;
    const int r = (idx >> 5) * 8 + (idx & 7), k = k0 + ((idx >> 3) & 3) * 4;
    t.w[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < bn && n0 + r < p.A && k < p.D) t.w[u] = __ldg(reinterpret_cast<const float4*>(p.W + (long long)(n0 + r) * p.D + k));
  }
}

// hi/lo split into the canonical layout: chunk ch of row r at ch*LBO + (r/8)*128 + (r%8)*16, LBO = rows*16.
__device__ __forceinline__ void store_ktile(const KTile& t, int bn, unsigned char* a_hi, unsigned char* a_lo,
                                            unsigned char* w_hi, unsigned char* w_lo) {
#pragma unroll
  for (int u = 0; u < kItemsA; ++u) {
    const int idx = u * kKeysNT + threadIdx.x;
    const unsigned off = (unsigned)((idx >> 3) & 3) * (kBM * 16u) + (unsigned)(idx >> 5) * 128u + (unsigned)(idx & 7) * 16u;
    float4 hi, lo;
    split4(t.a[u], hi, lo);
    *reinterpret_cast<float4*>(a_hi + off) = hi;
    *reinterpret_cast<float4*>(a_lo + off) = lo;
  }
#pragma unroll
  for (int u = 0; u < kItemsW; ++u) {
    const int idx = u * kKeysNT + threadIdx.x;
    if (idx < bn * (kBK / 4)) {
      const unsigned off = (unsigned)((idx >> 3) & 3) * ((unsigned)bn * 16u) + (unsigned)(idx >> 5) * 128u + (unsigned)(idx & 7) * 16u;
      float4 hi, lo;
      split4(t.w[u], hi, lo);
      *reinterpret_cast<float4*>(w_hi + off) = hi;
      *reinterpret_cast<float4*>(w_lo + off) = lo;
    }
  }
}

__global__ void __launch_bounds__(kKeysNT, 2) attn_keys_kernel(KeysArgs p) {
  extern __shared__ __align__(128) unsigned char smk[];
  __shared__ __align__(8) unsigned long long bar_free[kStages];
  __shared__ __align__(8) unsigned long long bar_done;
  __shared__ unsigned tmem_base_sm;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * kBM;
  const int n0 = blockIdx.y * p.BN;
  const int bn = min(p.BN, (p.A - n0 + 15) / 16 * 16);          // this CTA's N tile (multiple of 16)
  const size_t stage = keys_stage_bytes(p.BN);
  const unsigned a_bytes = kBM * kBK * 4, w_bytes = (unsigned)p.BN * kBK * 4;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&bar_free[s], 1);
    mbar_init(&bar_done, 1);
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

  // register double buffering: the global loads of k-block kb+1 are in flight while k-block kb is split, stored,
  // and its MMAs issued; the second CTA on the SM fills whatever bubbles remain
  KTile cur, nxt;
  load_ktile(cur, p, row0, n0, bn, 0);
  for (int kb = 0; kb < nkb; ++kb) {
    const int s = kb % kStages;
    unsigned char* st = smk + (size_t)s * stage;
    unsigned char* a_hi = st, *a_lo = st + a_bytes, *w_hi = st + 2 * a_bytes, *w_lo = st + 2 * a_bytes + w_bytes;
    if (kb + 1 < nkb) load_ktile(nxt, p, row0, n0, bn, (kb + 1) * kBK);
    if (kb >= kStages) mbar_wait(&bar_free[s], (unsigned)(((kb - kStages) / kStages) & 1));   // MMAs of k-block kb-kStages have read this stage
    store_ktile(cur, bn, a_hi, a_lo, w_hi, w_lo);
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
    cur = nxt;
  }
  if (tid == 0) umma_commit(&bar_done);
  mbar_wait(&bar_done, 0u);
  tc_fence_after();

  // ---- epilogue: TMEM -> registers -> + bias -> tanh -> global ------------------------------------------
  const int lq = warp & 3;                                      // TMEM lane quarter this warp may read
  const int nch = bn / 16, chalf = (nch + 1) / 2;
  const int cbeg = (warp >> 2) ? chalf * 16 : 0, cend = (warp >> 2) ? bn : chalf * 16;
  const int row = row0 + lq * 32 + lane;
  for (int c0 = cbeg; c0 < cend; c0 += 16) {
    unsigned r[16];
    const unsigned taddr = tmem + ((unsigned)(lq * 32) << 16) + (unsigned)c0;
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

namespace b2asr {
int attn_keys_v2(const float* h, const float* W, const float* bias, float* out, int M, int D, int A, cudaStream_t st);   // keys2.cu
}

using namespace b2asr;

extern "C" int b2asr_attn_keys(const float* h, const float* W, const float* bias, float* out,
                               int M, int D, int A, void* stream) {
  B2_REQUIRE(h && W && bias && out, B2ASR_EINVAL, "attn_keys: null pointer argument");
  B2_REQUIRE(M > 0 && D > 0 && A > 0, B2ASR_EINVAL, "attn_keys: bad sizes M=%d D=%d A=%d", M, D, A);
  B2_REQUIRE(D % 4 == 0 && (((uintptr_t)h | (uintptr_t)W) & 15) == 0, B2ASR_EUNSUPPORTED,
             "attn_keys: D=%d must be a multiple of 4 and h/W 16-byte aligned", D);
  // second generation (keys2.cu: TMA tensor copies, warp-specialised persistent pipeline); B2ASR_KEYS=v1 keeps the
  // first one, which also takes over when the driver cannot encode tensor maps
  static const bool v1 = [] { const char* e = getenv("B2ASR_KEYS"); return e && e[0] == 'v' && e[1] == '1'; }();
  if (!v1) {
    const int rc = attn_keys_v2(h, W, bias, out, M, D, A, (cudaStream_t)stream);
    if (rc != B2ASR_EUNSUPPORTED) return rc;
  }
  KeysArgs a;
  a.h = h; a.W = W; a.bias = bias; a.out = out; a.M = M; a.D = D; a.A = A;
  const int ntiles = (A + kBNmax - 1) / kBNmax;
  a.BN = (((A + ntiles - 1) / ntiles) + 15) / 16 * 16;          // even split, multiple of 16 (A=300 -> 160)
  const size_t smem = kStages * keys_stage_bytes(a.BN);
  B2_CUDA_TRY(cudaFuncSetAttribute(attn_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((unsigned)((M + kBM - 1) / kBM), (unsigned)ntiles);
  attn_keys_kernel<<<grid, kKeysNT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
