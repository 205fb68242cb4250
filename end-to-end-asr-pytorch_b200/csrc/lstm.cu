// Decoder-loop fusion, SURVEY.md section 8(f) rank 1: the pointwise half of the Speller's LSTMCell
// (reference src/asr.py:320-357 steps torch.nn.LSTMCell once per decoder step, asr.py:84-104).
//
// The host side (_dec.py) splits the cell's input projection W_ih [emb | context] so that the embedding part of
// every teacher-forced step is ONE batched GEMM before the loop and the weight gradients are ONE GEMM per
// matrix after the loop; what is left per step is two small GEMMs into `gates` and these kernels:
//   forward : gates [B,4H] (pre-activation sum, PyTorch order i,f,g,o) + both biases -> activations in place,
//             c = f*c_prev + i*g, h = o*tanh(c)
//   backward: (d_h, d_c, activations, c_prev, c) -> d_gates (pre-activation) written straight into the batch's
//             [B,L,4H] buffer slice (row stride) that the deferred weight-gradient GEMMs read, and d_c_prev.
#include "common.cuh"

namespace b2asr {

bool pdl_enabled();                           // attn_fwd.cu

__device__ __forceinline__ float sigmoid_acc(float x) { return __fdividef(1.f, 1.f + ex2(-x * kLog2e)); }

__global__ void __launch_bounds__(256) lstm_cell_fwd_kernel(float* __restrict__ gates, long long g_stride,
                                                            const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                                            const float* __restrict__ c_prev, float* __restrict__ h,
                                                            long long h_stride, float* __restrict__ c, int B, int H,
                                                            float* __restrict__ h2, long long h2_stride) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx - b * H;
  float* g = gates + (long long)b * g_stride;
  float bi = 0.f, bf = 0.f, bg = 0.f, bo = 0.f;
  if (b_ih) { bi += b_ih[u]; bf += b_ih[H + u]; bg += b_ih[2 * H + u]; bo += b_ih[3 * H + u]; }
  if (b_hh) { bi += b_hh[u]; bf += b_hh[H + u]; bg += b_hh[2 * H + u]; bo += b_hh[3 * H + u]; }
  const float i = sigmoid_acc(g[u] + bi);
  const float f = sigmoid_acc(g[H + u] + bf);
  const float gg = tanh_acc(g[2 * H + u] + bg);
  const float o = sigmoid_acc(g[3 * H + u] + bo);
  const float cn = fmaf(f, c_prev[idx], i * gg);
  g[u] = i; g[H + u] = f; g[2 * H + u] = gg; g[3 * H + u] = o;
  c[idx] = cn;
  const float hn = o * tanh_acc(cn);
  h[(long long)b * h_stride + u] = hn;
  if (h2) h2[(long long)b * h2_stride + u] = hn;       // second copy next to the next step's context (decloop.cu)
}

// d_h = d_h[b*dh_stride+u] (+ d_h2 + d_h3 when given: the recurrent and the attention-query paths of the native loop)
__global__ void __launch_bounds__(256) lstm_cell_bwd_kernel(const float* __restrict__ d_h, long long dh_stride,
                                                            const float* __restrict__ d_h2, const float* __restrict__ d_h3,
                                                            const float* __restrict__ d_c, const float* __restrict__ act,
                                                            long long act_stride, const float* __restrict__ c_prev,
                                                            const float* __restrict__ c, float* __restrict__ d_gates,
                                                            long long dg_stride, float* __restrict__ d_c_prev, int B, int H,
                                                            long long dh2_stride) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, u = idx - b * H;
  const float* a = act + (long long)b * act_stride;
  const float i = a[u], f = a[H + u], g = a[2 * H + u], o = a[3 * H + u];      // forward products: no dependence
  const float tc = tanh_acc(c[idx]);
  pdl_wait();                                                                    // on the previous backward kernels
  pdl_launch();
  float dh = d_h ? d_h[(long long)b * dh_stride + u] : 0.f;
  if (d_h2) dh += d_h2[(long long)b * dh2_stride + u];
  if (d_h3) dh += d_h3[idx];
  const float dct = (d_c ? d_c[idx] : 0.f) + dh * o * (1.f - tc * tc);
  float* dg = d_gates + (long long)b * dg_stride;
  dg[u] = dct * g * i * (1.f - i);
  dg[H + u] = dct * c_prev[idx] * f * (1.f - f);
  dg[2 * H + u] = dct * i * (1.f - g * g);
  dg[3 * H + u] = dh * tc * o * (1.f - o);
  d_c_prev[idx] = dct * f;
}

// launchers shared with the native decoder loop (decloop.cu)
void lstm_cell_fwd_launch(float* gates, long long g_stride, const float* b_ih, const float* b_hh, const float* c_prev,
                          float* h, long long h_stride, float* c, int B, int H, cudaStream_t st, float* h2,
                          long long h2_stride) {
  const int n = B * H;
  lstm_cell_fwd_kernel<<<(n + 255) / 256, 256, 0, st>>>(gates, g_stride, b_ih, b_hh, c_prev, h, h_stride, c, B, H, h2,
                                                        h2_stride);
}

void lstm_cell_bwd_launch(const float* d_h, long long dh_stride, const float* d_h2, const float* d_h3, const float* d_c,
                          const float* act, long long act_stride, const float* c_prev, const float* c, float* d_gates,
                          long long dg_stride, float* d_c_prev, int B, int H, cudaStream_t st, long long dh2_stride) {
  const int n = B * H;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((n + 255) / 256)); cfg.blockDim = dim3(256); cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  cudaLaunchKernelEx(&cfg, lstm_cell_bwd_kernel, d_h, dh_stride, d_h2, d_h3, d_c, act, act_stride, c_prev, c, d_gates,
                     dg_stride, d_c_prev, B, H, dh2_stride ? dh2_stride : (long long)H);
}

}  // namespace b2asr

using namespace b2asr;

extern "C" int b2asr_lstm_cell_fwd(float* gates, const float* b_ih, const float* b_hh, const float* c_prev,
                                   float* h, float* c, int B, int H, void* stream) {
  B2_REQUIRE(gates && c_prev && h && c, B2ASR_EINVAL, "lstm_cell_fwd: null pointer argument");
  B2_REQUIRE(B > 0 && H > 0, B2ASR_EINVAL, "lstm_cell_fwd: bad sizes B=%d H=%d", B, H);
  lstm_cell_fwd_launch(gates, 4ll * H, b_ih, b_hh, c_prev, h, H, c, B, H, (cudaStream_t)stream, nullptr, 0);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

extern "C" int b2asr_lstm_cell_bwd(const float* d_h, const float* d_c, const float* act, const float* c_prev, const float* c,
                                   float* d_gates, int64_t dg_stride, float* d_c_prev, int B, int H, void* stream) {
  B2_REQUIRE(act && c_prev && c && d_gates && d_c_prev, B2ASR_EINVAL, "lstm_cell_bwd: null pointer argument");
  B2_REQUIRE(B > 0 && H > 0 && dg_stride >= 4 * (int64_t)H, B2ASR_EINVAL, "lstm_cell_bwd: bad sizes B=%d H=%d stride=%lld", B, H, (long long)dg_stride);
  lstm_cell_bwd_launch(d_h, H, nullptr, nullptr, d_c, act, 4ll * H, c_prev, c, d_gates, dg_stride, d_c_prev, B, H,
                       (cudaStream_t)stream, 0);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}
