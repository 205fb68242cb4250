// Pieces shared by the tcgen05 kernels (keys2.cu, gemm_nt.cu): mbarrier / TMA tensor-copy / tensor-memory / MMA wrappers
// and the host-side tensor-map encoder.
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace b2asr {

__device__ __forceinline__ void mbar_wait_sleep(unsigned long long* bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void mbar_arrive_cta(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// one lane of a converged warp; code under `if (elect_one())` issues its tensor-core / TMA instructions straight from
// uniform registers (under `if (lane == 0)` the compiler wraps every one of them in an election loop: ~55 clk each)
__device__ __forceinline__ bool elect_one() {
  unsigned pred = 0;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred)::"memory");
  return pred != 0;
}
// shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart (tile base 1024-aligned)
__device__ __forceinline__ unsigned long long desc_sw128(unsigned saddr) {
  unsigned long long d = 0;
  d |= (unsigned long long)((saddr >> 4) & 0x3fffu);
  d |= (unsigned long long)1 << 16;                  // leading byte offset: unused for swizzled K-major
  d |= (unsigned long long)(1024u >> 4) << 32;       // stride byte offset between 8-row groups
  d |= 1ull << 46;                                   // descriptor version (sm_100)
  d |= 2ull << 61;                                   // SWIZZLE_128B
  return d;
}
// A operand from tensor memory (lane = row, one 32-bit column per k), B from shared memory
__device__ __forceinline__ void umma_tf32_ts(unsigned d_tmem, unsigned a_tmem, unsigned long long bdesc, unsigned idesc,
                                             unsigned accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld8(unsigned taddr, unsigned (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st16(unsigned taddr, const unsigned (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld16(unsigned taddr, unsigned (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}

// low half of the split: x - trunc_tf32(x), exact in fp32.  The tensor core keeps its top 10 mantissa bits (|error| <=
// 2^-10 |lo| <= 2^-20 |x|, the size of the dropped lo.lo term); rounding it first would double the converters' ALU work,
// which is what paces the pipeline.
__device__ __forceinline__ float lo_half(float x) {
  return x - __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}

// ---- host ------------------------------------------------------------------------------------------------------
inline PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// [rows, cols] fp32 tensor with `pitch` floats between rows (0 = cols), box = box_rows x 32 columns, swizzled as asked,
// zeros outside the tensor
enum MapSwizzle { kSwzNone = 0, kSwz128 = 1, kSwz128Atom32 = 2 };
inline int make_map(CUtensorMap* map, const float* base, long long rows, int cols, int box_rows, long long pitch = 0,
                    MapSwizzle swizzle = kSwz128) {
  PFN_cuTensorMapEncodeTiled_v12000 enc = encode_fn();
  B2_REQUIRE(enc != nullptr, B2ASR_EUNSUPPORTED, "tensor maps: cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)(pitch ? pitch : cols) * sizeof(float)};
  const cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult rc = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE,
                          swizzle == kSwz128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle == kSwz128Atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B2_REQUIRE(rc == CUDA_SUCCESS, B2ASR_ECUDA, "tensor maps: cuTensorMapEncodeTiled failed (%d)", (int)rc);
  return B2ASR_OK;
}


}  // namespace b2asr
