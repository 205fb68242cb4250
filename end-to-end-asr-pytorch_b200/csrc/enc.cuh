// Shared between the two implementations of the Listener recurrence (enc.cu: any H, exchange through L2;
// enc_cluster.cu: H in {256, 320}, exchange through distributed shared memory inside a 16-CTA cluster).
#pragma once
#include "common.cuh"

namespace b2asr {

struct SeqArgs {
  int B, T, H, ND;
  int Kp;                                    // padded reduction length in shared memory
  float* gates;                              // [B,T,ND,4H]
  const float* w_hh[2];
  const float* b_ih[2];
  const float* b_hh[2];
  const int* lens;
  float* out;                                // [B,T,ND*H]
  float* cells;                              // [B,T,ND,H]
  const float* d_out;                        // [B,T,ND*H]
  float* d_gates;                            // [B,T,ND,4H]
  int* status;                               // [16] device words: [0] 0 ok / 1 exchange timed out, [1] [2] failed polls fwd / bwd, [4-6] [8-10] CTA 0 cycles
};


__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.f, 1.f + ex2(-x * kLog2e)); }
__device__ __forceinline__ float dot4(const float4& a, const float4& b, float acc) {
  acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); return fmaf(a.w, b.w, acc);
}

// enc_cluster.cu: returns B2ASR_EUNSUPPORTED (without setting an error) when the shape or the device cannot take
// the cluster path, so the caller falls through to the L2-exchange kernels.
bool lstm_cluster_supported(int H);
int lstm_cluster_run(const SeqArgs& a, bool backward, cudaStream_t st);

}  // namespace b2asr
