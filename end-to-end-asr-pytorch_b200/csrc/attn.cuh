// Shared pieces of the per-decoder-step attention kernels (forward and backward).
#pragma once
#include "common.cuh"

namespace b2asr {

constexpr int kAttMaxC = 16;      // location-conv channels supported (reference hard-codes 10, asr.py:395)
constexpr int kAttMaxCluster = 8;
constexpr int kAttMaxIPT = 4;     // float4 context columns per thread

struct AttnArgs {
  int mode;                       // 0 dot, 1 loc
  const float* s; long long s_stride;          // [B,dec] decoder state
  const float* Wphi;                            // [A,dec] or nullptr (proj=False)
  const float* keys; long long k_bstride;       // [*,T,A]
  const float* h; long long h_bstride;          // [*,T,D]
  const int* lens;                              // [B]
  const float* prev_att;                        // [B,T] or nullptr (uniform first step)
  const long long* row_map;                     // forward only, or nullptr: row b reads s and prev_att of row row_map[b]
                                                // (beam decode: the survivor's parent row; saves the state gathers)
  const float* Wc; int C; int Kw;               // [C,2Kw+1]
  const float* Wl;                              // [A,C]
  const float* wg; const float* bg;             // [A],[1]
  float scale;
  float* att; float* ctx;                       // [B,T],[B,D]
  long long ctx_stride, dctx_stride;            // row strides of ctx (forward out) and d_ctx (backward in)
  float* q_out;                                 // [B,A] or nullptr
  int B, T, D, A, dec;
  // launch-derived
  int CS;                                       // cluster size (CTAs per utterance)
  int NT;                                       // threads per CTA
  int chunk_max;                                // ceil(T/CS)
  int TFK, TFH, NS; unsigned stage_bytes;       // ring geometry: frames per key / value tile, stages
  int FS, IPT;                                  // context phase: frame split, items per thread
  // backward only
  const float* q_saved; const float* att_saved; // [B,A], [B,T]
  const float* d_ctx; const float* d_att;       // [B,D], [B,T] or nullptr
  float* d_keys; long long dk_bstride;          // [B,T,A] accumulated in place
  float* d_prev_att;                            // [B,T] out (loc)
  float* d_zq;                                  // [B,A] out: grad wrt phi(s) pre-activation
  float* d_s;                                   // [B,dec] out or nullptr
  float* pgrad; long long pgrad_stride;         // [B*CS, pgrad_stride] per-CTA parameter-gradient slots
  long long* trace;                             // debug: clock64 stamps of CTA 0's phases (nullptr = off)
};

// row stride the NEXT attention launch of this host thread uses for ctx (forward) / d_ctx (backward) instead of D;
// set by the native decoder loop, which keeps [ctx_t | h_{t-1}] side by side for one GEMM; consumed by the launch
extern thread_local long long g_attn_row_stride;

#define B2_TRACE(p, slot) do { if ((p).trace && blockIdx.x == 0 && threadIdx.x == 0) (p).trace[slot] = clock64(); } while (0)
long long* attn_trace_buffer();                // device buffer when B2ASR_ATT_TRACE is set, else nullptr

// shared-memory carve-up, identical arithmetic on host (size) and device (pointers)
struct AttnSmem {
  size_t off_ring, off_bar, off_q, off_s, off_wc, off_pa, off_f, off_e, off_epart, off_ctxp, off_stat, off_red, off_extra, total;
};

__host__ __device__ inline size_t att_align(size_t v, size_t a) { return (v + a - 1) / a * a; }

__host__ __device__ inline AttnSmem attn_smem_layout(const AttnArgs& p, size_t extra_floats) {
  AttnSmem L;
  size_t o = 0;
  const int CP = (p.C + 3) / 4 * 4;
  const int nwarp = p.NT / 32;
  const int tfmax = p.TFK > p.TFH ? p.TFK : p.TFH;
  L.off_ring = o; o += (size_t)p.NS * p.stage_bytes;
  L.off_bar = o; o += att_align(8 * (size_t)p.NS, 16);
  L.off_q = o; o += att_align(4 * (size_t)p.A, 16);
  L.off_s = o; o += att_align(4 * (size_t)p.dec, 16);
  L.off_wc = o; o += att_align(4 * (size_t)(2 * p.Kw + 1) * CP, 16);
  L.off_pa = o; o += att_align(4 * (size_t)(p.chunk_max + 2 * p.Kw + 4), 16);
  L.off_f = o; o += att_align(4 * (size_t)p.chunk_max * CP, 16);
  L.off_e = o; o += att_align(4 * (size_t)(p.chunk_max + 4), 16);
  L.off_epart = o; o += att_align(4 * (size_t)nwarp * tfmax, 16);
  L.off_ctxp = o; o += att_align(4 * (size_t)p.FS * p.D, 16);
  L.off_stat = o; o += att_align(4 * 4 * (size_t)kAttMaxCluster, 16);
  L.off_red = o; o += 4 * 64;
  L.off_extra = o; o += att_align(4 * extra_floats, 16);
  L.total = o;
  return L;
}

// reductions over ALL threads of the CTA (every thread must call)
__device__ __forceinline__ float cta_reduce_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  float r = red[0];
  for (int i = 1; i < nw; ++i) r = fmaxf(r, red[i]);
  return r;
}
__device__ __forceinline__ float cta_reduce_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  float r = 0.f;
  for (int i = 0; i < nw; ++i) r += red[i];
  return r;
}

__device__ __forceinline__ void cluster_or_cta_sync(int CS) {
  if (CS > 1) cluster_sync_all(); else __syncthreads();
}

}  // namespace b2asr
