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

// GUARD: the clip and the NaN guard of solver.py:178-182 run on the device.  `norm` holds the 2-norm of the flat
// gradient as it stands (the SUM over ranks after the all-reduce); the averaged gradient's norm is norm * pre_scale.
//   NaN norm  -> nothing is updated, *skipped += 1 (the reference prints an error and skips asr_opt.step())
//   otherwise -> g = grad * pre_scale * min(max_norm / (norm + 1e-6), 1)   (torch.nn.utils.clip_grad_norm_), written
//                back so that .grad holds what the reference's .grad holds after clipping, then the update.
template <bool GUARD>
__global__ void __launch_bounds__(kThreads) adadelta_kernel(const b2asr_param_chunk* __restrict__ chunks, int n_chunks,
                                                            float* __restrict__ grad, float* __restrict__ sq,
                                                            float* __restrict__ acc, float lr, float rho, float eps,
                                                            float wd, const float* __restrict__ norm, float pre_scale,
                                                            float max_norm, int* __restrict__ skipped) {
  float coef = 1.f;
  if (GUARD) {
    const float nrm = __fmul_rn(__ldg(norm), pre_scale);
    if (nrm != nrm) {
      if (blockIdx.x == 0 && threadIdx.x == 0 && skipped) atomicAdd(skipped, 1);
      return;
    }
    coef = __fmul_rn(fminf(__fdiv_rn(max_norm, __fadd_rn(nrm, 1e-6f)), 1.f), pre_scale);
  }
  for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const b2asr_param_chunk ch = chunks[c];
    float* p = ch.param;
    float* g = grad + ch.offset;
    float* s = sq + ch.offset;
    float* a = acc + ch.offset;
    const int n = ch.count;
    const bool vec = ((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)s) | ((uintptr_t)a)) & 15) == 0;
    const int n4 = vec ? n >> 2 : 0;
    for (int i = threadIdx.x; i < n4; i += kThreads) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      float4 gv = reinterpret_cast<const float4*>(g)[i];
      if (GUARD) {
        gv.x = __fmul_rn(gv.x, coef); gv.y = __fmul_rn(gv.y, coef); gv.z = __fmul_rn(gv.z, coef); gv.w = __fmul_rn(gv.w, coef);
        reinterpret_cast<float4*>(g)[i] = gv;
      }
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
      float gi = g[i];
      if (GUARD) { gi = __fmul_rn(gi, coef); g[i] = gi; }
      adadelta_one(pv, gi, sv, av, lr, rho, eps, wd);
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
  adadelta_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(chunks, n_chunks, const_cast<float*>(grad), square_avg,
                                                                     acc_delta, lr, rho, eps, weight_decay, nullptr, 1.f,
                                                                     0.f, nullptr);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

extern "C" int b2asr_adadelta_step_clipped(const b2asr_param_chunk* chunks, int n_chunks, float* grad, float* square_avg,
                                           float* acc_delta, float lr, float rho, float eps, float weight_decay,
                                           const float* grad_norm, float pre_scale, float max_norm, int* skipped,
                                           void* stream) {
  using namespace b2asr;
  B2_REQUIRE(n_chunks >= 0, B2ASR_EINVAL, "b2asr_adadelta_step_clipped: n_chunks %d", n_chunks);
  if (n_chunks == 0) return B2ASR_OK;
  B2_REQUIRE(chunks && grad && square_avg && acc_delta && grad_norm, B2ASR_EINVAL, "b2asr_adadelta_step_clipped: null pointer");
  B2_REQUIRE(pre_scale > 0.f && max_norm > 0.f, B2ASR_EINVAL, "b2asr_adadelta_step_clipped: pre_scale and max_norm must be positive");
  const int grid = n_chunks < 148 * 8 ? n_chunks : 148 * 8;
  adadelta_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(chunks, n_chunks, grad, square_avg, acc_delta, lr, rho,
                                                                    eps, weight_decay, grad_norm, pre_scale, max_norm,
                                                                    skipped);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
