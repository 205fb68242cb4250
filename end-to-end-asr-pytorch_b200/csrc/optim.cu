// Optimizer step of the train step, reference src/solver.py:101-106 (torch.optim.Adadelta(params, lr, eps=1e-8)) and
// solver.py:178-182 (clip -> NaN guard -> asr_opt.step()).  torch's foreach implementation walks ~60 parameter
// tensors in seven multi-tensor passes (0.95 ms per step at M2, launch-bound); here the gradients already live in ONE
// flat buffer (solver.FlatGradients), the two state vectors are laid out the same way, and one launch updates
// everything: a table of chunks maps flat offsets back to the parameter tensors, which stay where torch put them.
//
//   per element (torch/optim/adadelta.py, same order of operations):
//     g   = grad (+ weight_decay * p)
//     sq  = rho * sq + (1 - rho) * g * g
//     d   = sqrt(acc + eps) / sqrt(sq + eps) * g
//     acc = rho * acc + (1 - rho) * d * d
//     p  -= lr * d
//
// HBM-bound: 4 reads + 3 writes of 4 bytes per parameter (M2: 15.0 M parameters, 420 MB, ~70 us).
#include "common.cuh"

namespace b2asr {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void adadelta_one(float& p, float g, float& sq, float& acc, float lr, float rho, float eps,
                                             float wd) {
  if (wd != 0.f) g = fmaf(wd, p, g);
  sq = __fadd_rn(__fmul_rn(sq, rho), __fmul_rn(__fmul_rn(1.f - rho, g), g));
  const float std = sqrtf(sq + eps);
  const float d = __fmul_rn(__fdiv_rn(sqrtf(acc + eps), std), g);
  acc = __fadd_rn(__fmul_rn(acc, rho), __fmul_rn(__fmul_rn(1.f - rho, d), d));
  p = __fadd_rn(p, __fmul_rn(-lr, d));
}

__global__ void __launch_bounds__(kThreads) adadelta_kernel(const b2asr_param_chunk* __restrict__ chunks, int n_chunks,
                                                            const float* __restrict__ grad, float* __restrict__ sq,
                                                            float* __restrict__ acc, float lr, float rho, float eps,
                                                            float wd) {
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const b2asr_param_chunk ch = chunks[c];
    float* p = ch.param;
    const float* g = grad + ch.offset;
    float* s = sq + ch.offset;
    float* a = acc + ch.offset;
    const int n = ch.count;
    const bool vec = ((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)s) | ((uintptr_t)a)) & 15) == 0;
    const int n4 = vec ? n >> 2 : 0;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
      float4 sv = reinterpret_cast<float4*>(s)[i], av = reinterpret_cast<float4*>(a)[i];
      adadelta_one(pv.x, gv.x, sv.x, av.x, lr, rho, eps, wd);
      adadelta_one(pv.y, gv.y, sv.y, av.y, lr, rho, eps, wd);
      adadelta_one(pv.z, gv.z, sv.z, av.z, lr, rho, eps, wd);
      adadelta_one(pv.w, gv.w, sv.w, av.w, lr, rho, eps, wd);
      reinterpret_cast<float4*>(p)[i] = pv;
      reinterpret_cast<float4*>(s)[i] = sv;
      reinterpret_cast<float4*>(a)[i] = av;
    }
    for (int i = 4 * n4 + threadIdx.x; i < n; i += kThreads) {
      float pv = p[i], sv = s[i], av = a[i];
      adadelta_one(pv, g[i], sv, av, lr, rho, eps, wd);
      p[i] = pv; s[i] = sv; a[i] = av;
    }
  }
}

}  // namespace

}  // namespace b2asr

extern "C" size_t b2asr_param_chunk_size(void) { return sizeof(b2asr_param_chunk); }

extern "C" int b2asr_adadelta_step(const b2asr_param_chunk* chunks, int n_chunks, const float* grad, float* square_avg,
                                   float* acc_delta, float lr, float rho, float eps, float weight_decay, void* stream) {
  using namespace b2asr;
  B2_REQUIRE(n_chunks >= 0, B2ASR_EINVAL, "b2asr_adadelta_step: n_chunks %d", n_chunks);
  if (n_chunks == 0) return B2ASR_OK;
  B2_REQUIRE(chunks && grad && square_avg && acc_delta, B2ASR_EINVAL, "b2asr_adadelta_step: null pointer");
  const int grid = n_chunks < 148 * 8 ? n_chunks : 148 * 8;
  adadelta_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(chunks, n_chunks, grad, square_avg, acc_delta, lr, rho, eps,
                                                              weight_decay);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
