// Microbenchmark: issue rate of the warp-level tensor-core path (mma.sync.m16n8k8 tf32 / m16n8k16 bf16) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu ; run: ./mma_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS, bool BF16>
__global__ void k(float* out, int iters, long long* clk) {
  float c[CHAINS][4];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) { c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f; }
  unsigned a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 7, a3 = 11, b0 = 5, b1 = threadIdx.x + 1;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (BF16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int CHAINS, bool BF16>
void run(int warps) {
  float* out; long long* clk; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 8);
  const int iters = 2000;
  k<CHAINS, BF16><<<148, warps * 32>>>(out, iters, clk);
  k<CHAINS, BF16><<<148, warps * 32>>>(out, iters, clk);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  const double per_sm = (double)h / ((double)iters * CHAINS * warps);
  const double mac = BF16 ? 16.0 * 8 * 16 : 16.0 * 8 * 8;
  printf("%s warps %2d chains %d: %.2f clk per MMA per SM -> %.0f MAC/clk/SM (dependent-chain latency if chains=1: %.1f clk)\n",
         BF16 ? "bf16 m16n8k16" : "tf32 m16n8k8 ", warps, CHAINS, per_sm, mac / per_sm, (double)h / iters);
  cudaFree(out); cudaFree(clk);
}

int main() {
  run<1, false>(1);
  run<1, false>(4);
  run<2, false>(4);
  run<4, false>(4);
  run<4, false>(8);
  run<2, false>(10);
  run<4, false>(10);
  run<4, false>(12);
  run<8, false>(16);
  run<1, true>(1);
  run<4, true>(4);
  run<4, true>(8);
  run<4, true>(12);
  run<8, true>(16);
  return 0;
}
