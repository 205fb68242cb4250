// Shared device/host helpers for the b2asr sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/b2asr.h"

namespace b2asr {

// ---- error plumbing (host) -------------------------------------------------
void set_error(const char* fmt, ...);
int  cuda_fail(cudaError_t e, const char* what);

#define B2_CUDA_TRY(expr)                                             \
  do {                                                                \
    cudaError_t e__ = (expr);                                         \
    if (e__ != cudaSuccess) return ::b2asr::cuda_fail(e__, #expr);    \
  } while (0)

#define B2_REQUIRE(cond, code, ...)                                   \
  do {                                                                \
    if (!(cond)) { ::b2asr::set_error(__VA_ARGS__); return (code); }  \
  } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- device helpers ----------------------------------------------------------
constexpr float kNeg   = -1.0e30f;           // finite stand-in for log(0): never yields inf-inf
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2   = 0.6931471805599453f;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float ex2(float x) {      // MUFU.EX2, 2 ulp, flushes to 0 for very negative x
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
__device__ __forceinline__ float lg2(float x) {      // MUFU.LG2
  float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// packed fp32: two FMAs per issue slot (SASS FFMA2)
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi)); return d;
}
__device__ __forceinline__ float sum2(f32x2 v) {
  float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); return lo + hi;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}

// cp.async (LDGSTS) 16-byte copy global -> shared, generic proxy.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Programmatic dependent launch (the decoder loop's kernels are launched with the programmatic-serialisation
// attribute): a kernel may be SCHEDULED before its predecessor in the stream has finished; everything it does before
// pdl_wait() must not depend on the predecessor (weights, encoder features, barrier set-up), pdl_wait() returns once
// the predecessor has completed and its writes are visible.  Both are no-ops in a plain launch.
// The predecessor's trigger is left IMPLICIT (its CTAs' exit): measured at M2, triggering early
// (griddepcontrol.launch_dependents right after the wait) made the loop slower (forward 7.4 -> 7.8 ms, backward 14.8
// -> 15.9 ms per batch: the successor's CTAs land on every SM a CTA of the running kernel leaves and their prologue
// competes with the stragglers the clusters are waiting for), the implicit trigger faster (6.5 / 13.4 ms).
// -DB2ASR_PDL_TRIGGER restores the explicit trigger for A/B runs.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#ifdef B2ASR_PDL_TRIGGER
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#else
__device__ __forceinline__ void pdl_launch() {}
#endif
// The forward's decoder-cell kernel (80 CTAs between two attention steps) does trigger, at its very start: the next
// attention kernel's CTAs are then scheduled next to it and run their prologue (weights, key / feature tiles, the
// location conv over the previous alignment -- all complete before the cell kernel could start) while the cell
// computes (M2 forward 6.53 -> 6.28 ms per batch).  The attention kernels themselves never trigger early, and neither
// do the backward's small kernels (measured: 13.4 -> 19.6 ms).
#ifdef B2ASR_PDL_NO_SMALL_TRIGGER
__device__ __forceinline__ void pdl_trigger_small() {}
#else
__device__ __forceinline__ void pdl_trigger_small() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// thread-block cluster helpers
__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ unsigned cluster_nctarank() {
  unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// map a local shared-memory address to the same offset in CTA `rank` of the cluster
__device__ __forceinline__ unsigned dsmem_addr(const void* local_smem, unsigned rank) {
  unsigned a = (unsigned)__cvta_generic_to_shared(local_smem), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ float ld_dsmem_f32(unsigned addr) {
  float v; asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory"); return v;
}
__device__ __forceinline__ float4 ld_dsmem_f32x4(unsigned addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_dsmem_f32(unsigned addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

}  // namespace b2asr

// ---- mbarrier + 1-D bulk (TMA) copy helpers ---------------------------------------------------
namespace b2asr {

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy, completion signalled on `bar` (bytes: multiple of 16; 16-byte aligned both sides)
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// fire-and-forget reduction.  No "memory" clobber on purpose: nothing in the issuing kernel reads the target back,
// and a clobber would pin every surrounding load/store in program order (it serialised the unrolled frames of
// the attention backward's key pass).
__device__ __forceinline__ void red_add_f32(float* gaddr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(gaddr), "f"(v));
}
// predicated form: no branch around the instruction, so unrolled iterations stay one basic block
__device__ __forceinline__ void red_add_f32_if(bool on, float* gaddr, float v) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p red.global.add.f32 [%0], %1;\n\t}" ::"l"(gaddr), "f"(v), "r"((int)on));
}

// accurate tanh from one ex2 and one reciprocal: 1 - 2/(1+e^{2x}); abs error ~1e-7
__device__ __forceinline__ float tanh_acc(float x) {
  const float xc = fminf(fmaxf(x, -15.f), 15.f);
  const float e = ex2(xc * (2.f * kLog2e));
  return 1.f - __fdividef(2.f, 1.f + e);
}

// same formula without the range clamps: e = inf gives 1 - 2/inf = 1, e = 0 gives -1; never NaN for finite x
__device__ __forceinline__ float tanh_fast(float x) {
  const float e = ex2(x * (2.f * kLog2e));
  return 1.f - __fdividef(2.f, 1.f + e);
}

}  // namespace b2asr
