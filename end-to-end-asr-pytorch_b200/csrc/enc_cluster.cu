// Listener recurrence, cluster variant (H = 256 or 320: the shipped timit / libri encoders).  Same contract as
// enc.cu (reference src/asr.py:466-501), but the per-frame exchange never leaves the SMs:
//
//   * one 16-CTA thread-block cluster per (direction, group of BS batch rows); CTA `rank` owns hidden units
//     [rank*H/16, (rank+1)*H/16) and keeps its 4*H/16 rows of W_hh in REGISTERS for the whole sequence
//     (80 floats per thread at H=320), its cell state / cell gradient in registers too;
//   * forward is an all-gather: a CTA multiplies the full h_{t-1} (BS x H, local shared memory) with its gate
//     rows, applies the cell, and sends its BS x H/16 slice of h_t to all 16 CTAs of the cluster;
//   * backward is a reduce-scatter: a CTA multiplies ITS OWN d_gates of the previous frame (BS x 4H/16, local)
//     with its rows of W_hh -- partial dh for all H units -- and sends each owner its BS x H/16 slice; the owner
//     adds the 16 partials in rank order (bit-reproducible).  The d_gates themselves never cross CTAs, so both
//     directions of the exchange move the same bytes per CTA and frame (forward h, backward partial dh);
//   * the exchange is 16 bulk copies (cp.async.bulk shared::cta -> shared::cluster) per CTA and frame, each
//     completing on the DESTINATION's mbarrier: a CTA waits on its own barrier until all 16 slices have landed.
//     No cluster barrier and no thread-issued remote stores inside the loop; staging buffers are double-buffered
//     and the data dependence itself is the flow control (a peer cannot produce frame t+1 before it has consumed
//     everybody's frame t);
//   * the exchange is bound by the ~21 B/clk one SM moves into its peers (15 KB per frame: ~0.4 us), so it travels in
//     two parts: forward, rows [0,BS/3) of a slice go first on their own mbarrier and the receivers start multiplying
//     them while the rest is in flight; backward, the partials of the first two thirds of the rows are forwarded by
//     an extra warp (block = H+32 threads, named barriers) while the compute warps multiply the last third;
//   * BS in {8, 12, 16}: the device holds only ~7 clusters of 16 CTAs at once (cudaOccupancyMaxActiveClusters),
//     so the rows are split into as many groups as fit in one wave;
//   * clusters are independent of each other (no cooperative launch).
// ncu/clock64 evidence and the comparison with the L2-exchange kernels: DESIGN.md section 4, profiles/.
#include <stdlib.h>

#include "enc.cuh"

namespace b2asr {

namespace {

constexpr int NCL = 16;                       // CTAs per cluster (non-portable size, allowed on sm_100)
constexpr unsigned kWaitLimit = 1u << 22;     // mbarrier try_wait attempts before a CTA gives up

__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// local shared memory -> shared memory of CTA `rank` of the cluster, completion on THAT CTA's mbarrier
__device__ __forceinline__ void bulk_s2cluster(void* dst_local_alias, const void* src, unsigned bytes,
                                               unsigned long long* bar_local_alias, unsigned rank) {
  const unsigned d = dsmem_addr(dst_local_alias, rank), m = dsmem_addr(bar_local_alias, rank);
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(d), "r"(smem_u32(src)), "r"(bytes), "r"(m) : "memory");
}
__device__ __forceinline__ bool mbar_wait_bounded(unsigned long long* bar, unsigned parity) {
  const unsigned addr = smem_u32(bar);
  for (unsigned i = 0; i < kWaitLimit; ++i) {
    unsigned ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}

// named barriers: the compute warps arrive (do not block), the copy-issuing warp waits for all of them
__device__ __forceinline__ void named_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void named_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
constexpr int kIssueWarp = 32;                // backward: one extra warp per CTA forwards the first part of the partials

// packed fp32: two FMAs per issue slot (SASS FFMA2); the loops below are issue-bound with 2.5 warps per scheduler
// (f32x2 / ffma2 / pack2 / sum2: common.cuh)

// floats between two ranks' chunks in the gathered h buffer: >= BS*UC and == UC (mod 32), so the 8 float4 a
// quarter warp reads (32 consecutive k, up to three ranks) fall into 8 different bank groups
constexpr int chunk_stride(int UC, int BS) { return UC + 32 * ((BS * UC - UC + 31) / 32); }

// ---------------------------------------------------------------------------------------------------------
// forward.  block = H threads: (unit u = tid/16 of this CTA's H/16, k-slice ks = tid%16)
// ---------------------------------------------------------------------------------------------------------
template <int H, int BS>
__global__ void __launch_bounds__(H, 1) lstm_cl_fwd_kernel(const SeqArgs a) {
  constexpr int UC = H / NCL;                 // units per CTA
  constexpr int KJ = H / 64;                  // float4 per lane and row: lane ks reads k = 4*(ks + 16 j)
  constexpr int RB = BS <= 8 ? 8 : 16;        // rows the 16-lane butterfly resolves (rows >= BS are zeros)
  constexpr int S = chunk_stride(UC, BS);
  constexpr int CH = BS * UC;                 // floats one CTA contributes per frame
  constexpr int R1 = BS / 3;                  // rows sent (and consumed) first: the MACs start on them while the rest travels
  constexpr int CH0 = R1 * UC, CH1 = CH - CH0;
  static_assert(H % 64 == 0 && UC % 4 == 0 && BS % 4 == 0 && BS <= 16, "slices are whole float4s");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long full[2][2];   // [part][buffer]: rows [0,R1) and [R1,BS) of every slice
  __shared__ int lens_s[32];                  // lengths of this cluster's rows (0 past the batch)
  float* h_s = smem;                          // [2][NCL][S]: chunk of rank r = [BS][UC]
  float* stage = smem + 2 * NCL * S;          // [2][BS][UC]
  const int tid = threadIdx.x, ks = tid & 15, u = tid >> 4;
  const unsigned rank = cluster_ctarank();
  // rows are dealt round-robin to the row groups (row r of group y = batch row y + r*groups): with a batch sorted
  // by length every cluster gets the same mix of long and short rows
  const int dir = blockIdx.z, b0 = blockIdx.y, bstep = gridDim.y, u0 = (int)rank * UC;
  const int T = a.T, ND = a.ND;
  const long long ldo = (long long)ND * H;

  ulonglong2 w[4][KJ];                        // float4 of W_hh seen as two packed pairs (k, k+1), (k+2, k+3)
  int koff[KJ];
  {
    const float* W = a.w_hh[dir];
#pragma unroll
    for (int j = 0; j < KJ; ++j) {
      const int k = 4 * (ks + 16 * j);
      koff[j] = (k / UC) * S + (k % UC);
#pragma unroll
      for (int g = 0; g < 4; ++g)
        w[g][j] = __ldg(reinterpret_cast<const ulonglong2*>(W + ((long long)g * H + u0 + u) * H + k));
    }
  }
  for (int idx = tid; idx < NCL * S; idx += H) h_s[idx] = 0.f;
  if (tid < 32) lens_s[tid] = (tid < BS && b0 + tid * bstep < a.B) ? a.lens[b0 + tid * bstep] : 0;
  if (tid == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&full[i >> 1][i & 1], 1);
    mbar_fence_init();
    for (int q = 0; q < 2; ++q) {               // frame 1 reads buffer 1, frame 2 buffer 0
      mbar_expect_tx(&full[0][q], NCL * CH0 * 4);
      mbar_expect_tx(&full[1][q], NCL * CH1 * 4);
    }
  }

  // cell role: RB = 8 -> even lanes own row ks/2; RB = 16 -> every lane owns row ks
  const int row = RB == 8 ? ks >> 1 : ks;
  const bool pw = (RB == 8 ? (ks & 1) == 0 : true) && row < BS;
  const int gb = b0 + row * bstep, unit = u0 + u;
  const bool valid = pw && gb < a.B;
  const int len = valid ? a.lens[gb] : 0;
  float bias[4] = {0.f, 0.f, 0.f, 0.f};
  if (valid) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (a.b_ih[dir]) bias[g] += a.b_ih[dir][g * H + unit];
      if (a.b_hh[dir]) bias[g] += a.b_hh[dir][g * H + unit];
    }
  }
  float c = 0.f;
  unsigned phase = 0;                         // bit 2*part+p = parity to wait for on full[part][p]
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0, cyc_r0 = 0, cyc_r1 = 0, cyc_r2 = 0, cyc_r3 = 0;
  __syncthreads();
  cluster_arrive();
  cluster_wait();                             // every CTA's barriers are armed before the first remote copy

  // pre-activations are fetched one frame ahead (HBM latency ~1 us would otherwise sit on the critical path)
  float pre[4] = {0.f, 0.f, 0.f, 0.f};
  {
    const int t0 = dir ? T - 1 : 0;
    if (valid && t0 < len) {
      const float* g0 = a.gates + (((long long)gb * T + t0) * ND + dir) * 4 * H + unit;
#pragma unroll
      for (int g = 0; g < 4; ++g) pre[g] = g0[(long long)g * H];
    }
  }
  for (int s = 0; s < T; ++s) {
    const int t = dir ? T - 1 - s : s;
    const int p = s & 1;
    const bool active = valid && t < len;
    float* grow = a.gates + (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    float nxt[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int t1 = dir ? t - 1 : t + 1;
      if (s + 1 < T && valid && t1 < len) {
        const float* g1 = a.gates + (((long long)gb * T + t1) * ND + dir) * 4 * H + unit;
#pragma unroll
        for (int g = 0; g < 4; ++g) nxt[g] = g1[(long long)g * H];
      }
    }
    const long long c0 = clock64();
    if (s > 0) {                              // h_{t-1}: rows [0,R1) of all 16 slices have landed in h_s[p]
      if (!mbar_wait_bounded(&full[0][p], (phase >> p) & 1u)) { atomicExch(a.status, 1); return; }
      phase ^= 1u << p;
      if (tid == 0) mbar_expect_tx(&full[0][p], NCL * CH0 * 4);  // re-arm for frame s+2
    }
    const long long c1 = clock64();
    float acc[RB][4];
#pragma unroll
    for (int b = BS; b < RB; ++b)
#pragma unroll
      for (int g = 0; g < 4; ++g) acc[b][g] = 0.f;
    {
      const float* hb = h_s + p * (NCL * S);
      // rows past their length at this frame are skipped (their h_t is written as zero below): CTA-uniform mask
      const unsigned rows_on = __ballot_sync(kFull, t < lens_s[tid & 31]);
#pragma unroll
      for (int b = 0; b < BS; ++b) {
        if (b == R1 && s > 0) {                       // the remaining rows
          if (!mbar_wait_bounded(&full[1][p], (phase >> (2 + p)) & 1u)) { atomicExch(a.status, 1); return; }
          phase ^= 4u << p;
          if (tid == 0) mbar_expect_tx(&full[1][p], NCL * CH1 * 4);
        }
        f32x2 a2[4] = {0ull, 0ull, 0ull, 0ull};      // (even-k, odd-k) partial sums of the row's four gates
        if ((rows_on >> b) & 1u) {
#pragma unroll
        for (int j = 0; j < KJ; ++j) {
          const ulonglong2 hv = *reinterpret_cast<const ulonglong2*>(hb + koff[j] + b * UC);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            a2[g] = ffma2(hv.x, w[g][j].x, a2[g]);
            a2[g] = ffma2(hv.y, w[g][j].y, a2[g]);
          }
        }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) acc[b][g] = sum2(a2[g]);
      }
    }
    // transposing butterfly over the 16 k-slices: RB rows -> one row per lane (pair)
#pragma unroll
    for (int m = 8, n = RB / 2; n >= 1; m >>= 1, n >>= 1) {
      const bool hi = ks & m;
#pragma unroll
      for (int i = 0; i < n; ++i)
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const float send = hi ? acc[i][g] : acc[i + n][g];
          const float keep = hi ? acc[i + n][g] : acc[i][g];
          acc[i][g] = keep + __shfl_xor_sync(kFull, send, m);
        }
    }
    if (RB == 8) {
#pragma unroll
      for (int g = 0; g < 4; ++g) acc[0][g] += __shfl_xor_sync(kFull, acc[0][g], 1);
    }
    const long long c2 = clock64();
    float sv[4] = {0.f, 0.f, 0.f, 0.f}, hval = 0.f;
    if (pw) {
      if (active) {
        const float gi = sigmoid_f(acc[0][0] + pre[0] + bias[0]);
        const float gf = sigmoid_f(acc[0][1] + pre[1] + bias[1]);
        const float gg = tanh_acc(acc[0][2] + pre[2] + bias[2]);
        const float go = sigmoid_f(acc[0][3] + pre[3] + bias[3]);
        c = fmaf(gf, c, gi * gg);
        hval = go * tanh_acc(c);
        sv[0] = gi; sv[1] = gf; sv[2] = gg; sv[3] = go;
      } else {
        c = 0.f;
      }
      stage[p * CH + row * UC + u] = hval;
    }
    if (active) {
      grow[0] = sv[0]; grow[(long long)H] = sv[1]; grow[2ll * H] = sv[2]; grow[3ll * H] = sv[3];
      a.cells[(((long long)gb * T + t) * ND + dir) * H + unit] = c;
    }
    if (valid) a.out[((long long)gb * T + t) * ldo + dir * H + unit] = hval;
    const long long c3 = clock64();
    long long c4 = c3, c5 = c3;
    if (s + 1 < T) {
      fence_proxy_async_smem();               // the cell outputs above are read by the async proxy below
      __syncthreads();
      c4 = clock64();
      // this CTA's BS x UC slice of h_t -> chunk `rank` of h_s[1-p] in every CTA of the cluster
      // (four lanes per warp issue the 32 copies, lanes 0-1 the first rows, lanes 2-3 the rest.  How they are issued
      // does not change the frame time -- one lane per warp looping, distinct lanes, a warp of its own were all
      // measured: the 15 KB a CTA sends per frame share its ~21 B/clk path into the peers, ~0.4 us)
      {
        const int lane = tid & 31, d = (tid >> 5) + (lane & 1) * (H / 32);
        float* dst = h_s + (p ^ 1) * (NCL * S) + (int)rank * S;
        if (lane < 2 && d < NCL) bulk_s2cluster(dst, stage + p * CH, CH0 * 4, &full[0][p ^ 1], (unsigned)d);
        else if (lane < 4 && d < NCL) bulk_s2cluster(dst + CH0, stage + p * CH + CH0, CH1 * 4, &full[1][p ^ 1], (unsigned)d);
      }
      c5 = clock64();
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) pre[g] = nxt[g];
    const long long c6 = clock64();
    cyc_wait += c1 - c0; cyc_mac += c2 - c1; cyc_rest += c6 - c2;
    cyc_r0 += c3 - c2; cyc_r1 += c4 - c3; cyc_r2 += c5 - c4; cyc_r3 += c6 - c5;
  }
  cluster_arrive();
  cluster_wait();                             // nobody leaves while a copy may still read or write its shared memory
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {   // CTA 0's cycle split (clock64)
    a.status[4] = (int)cyc_wait; a.status[5] = (int)cyc_mac; a.status[6] = (int)cyc_rest;
    a.status[11] = (int)cyc_r0; a.status[12] = (int)cyc_r1; a.status[13] = (int)cyc_r2; a.status[14] = (int)cyc_r3;
  }
}

// ---------------------------------------------------------------------------------------------------------
// forward, tensor-core variant (default; B2ASR_ENC_MMA=0 selects the FFMA2 kernel above).
//
// The frame's product G[16 x 4UC] = h_{t-1}[16 x H] . W_slice^T runs on the warp-level tensor-core path
// (mma.sync.m16n8k16, fp16 operands, fp32 accumulation; measured on B200 with tools/micro/mma_rate.cu: 2.0 clk per
// instruction and SM = 1022 MAC/clk/SM, against the 71 MAC/clk/SM the FFMA2 loop reaches) with both operands split in
// two fp16 terms, x = hi + lo (hi = rn16(x), lo = rn16(x - hi): 22 significant bits, absolute error <= 3e-8 for
// |x| <= 1 -- h is in (-1,1), weights are O(1)), and three products per k-step: hi.hi + hi.lo + lo.hi (the dropped
// lo.lo term is 2^-22 of the result; fp16 x fp16 products are exact in the fp32 accumulator).
//   * the CTA's 4UC gate columns are H/32 m16n8 tiles (tile n = the four gates of units 2n, 2n+1).  Warp (pair, half)
//     multiplies the TWO tiles of its pair over ONE half of K (ranks 0-7 or 8-15), so an A fragment feeds six
//     instructions and the CTA reads h from shared memory five times per frame instead of ten (the loads, not the
//     tensor pipe, bound the one-tile-per-warp version: 1.25 us); the two warps of a pair swap one partial tile through
//     shared memory (named barrier of 64 threads) and each finishes one tile.  The W fragments, hi and lo, stay in
//     registers for the whole sequence (4*H/16 = 80 registers at H=320);
//   * h crosses the cluster ALREADY SPLIT: the cell owner writes (hi, lo) of its unit once, 4 bytes per value as before,
//     so the exchange moves the same bytes and the receivers' A fragments are plain 128-bit shared-memory loads
//     without a single conversion instruction.  Inside a rank's chunk the layout is [unit quad][row][hi0 hi1 hi2 hi3 lo0
//     lo1 lo2 lo3] (16 bytes per row and quad);
//   * k-step (m, q) takes quad q of the four ranks 4m..4m+3, one rank per thread-in-group: with the chunk stride
//     == 8 words (mod 32) the 8 lanes of a shared-memory phase (2 rows x 4 ranks) hit 8 different 16-byte bank groups.
//     The order of k inside a tile is free as long as A and B agree, and B is laid out once, in registers;
//   * the accumulator fragment holds (i,f) or (g,o) of a unit for rows g and g+8: one shuffle pair with the neighbour
//     lane gives every lane all four gates of ONE (row, unit) -- 320 lanes = 16 rows x 20 units, every lane a cell.
// ---------------------------------------------------------------------------------------------------------
// words between two ranks' chunks: room for rows up to 15 of the last quad (they are read when BS > 8), == 8 (mod 32)
__host__ __device__ constexpr int mma_chunk_need(int UC, int BS) { return BS > 8 ? ((UC / 4 - 1) * BS + 16) * 4 : BS * UC; }
__host__ __device__ constexpr int mma_chunk_stride(int UC, int BS) { return ((mma_chunk_need(UC, BS) - 8 + 31) / 32) * 32 + 8; }
__device__ __forceinline__ void mma_f16(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ unsigned short f16_bits(float x) {
  unsigned short h; asm("cvt.rn.f16.f32 %0, %1;" : "=h"(h) : "f"(x)); return h;
}
__device__ __forceinline__ float f16_value(unsigned short h) {
  float x; asm("cvt.f32.f16 %0, %1;" : "=f"(x) : "h"(h)); return x;
}
__device__ __forceinline__ void split_f16(float x, unsigned short& hi, unsigned short& lo) {
  hi = f16_bits(x);
  lo = f16_bits(x - f16_value(hi));
}

template <int H, int BS>
__global__ void __launch_bounds__(H, 1) lstm_cl_fwd_mma_kernel(const SeqArgs a) {
  constexpr int UC = H / NCL;                 // units per CTA
  constexpr int QC = UC / 4;                  // unit quads per rank
  constexpr int KS = 4 * QC;                  // k16-steps: (m, q) = quad q of ranks 4m..4m+3
  constexpr int CH = BS * UC;                 // 32-bit words one CTA contributes per frame
  constexpr int S = mma_chunk_stride(UC, BS);
  static_assert(H / 32 == (4 * UC) / 8 && UC % 4 == 0 && BS % 4 == 0 && BS <= 16, "one m16n8 tile per warp");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long full[2];
  unsigned* h_s = reinterpret_cast<unsigned*>(smem);          // [2][NCL][S]
  unsigned* stage = h_s + 2 * NCL * S;                        // [2][CH]
  constexpr int NP = H / 64;                  // warp pairs
  __shared__ float4 xch[H / 32][32];          // the partial tile a warp hands to its partner
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, g = lane >> 2, tig = lane & 3;
  const int np = w % NP, kh = w / NP;         // tiles 2np, 2np+1; ranks [8kh, 8kh+8) of K
  const int tn = 2 * np + kh;                 // the tile this warp finishes
  const unsigned rank = cluster_ctarank();
  const int dir = blockIdx.z, b0 = blockIdx.y, bstep = gridDim.y, u0 = (int)rank * UC;
  const int T = a.T, ND = a.ND;
  const long long ldo = (long long)ND * H;

  // B fragments: column n = g of tile x is gate (g & 3) of unit 2x + (g >> 2)
  unsigned bh[2][KS / 2][2], bl[2][KS / 2][2];
#pragma unroll
  for (int x = 0; x < 2; ++x) {
    const float* Wr = a.w_hh[dir] + ((long long)(g & 3) * H + u0 + 2 * (2 * np + x) + (g >> 2)) * H;
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int q = 0; q < QC; ++q) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(Wr + (4 * (2 * kh + m) + tig) * UC + 4 * q));
        unsigned short h0, h1, h2, h3, l0, l1, l2, l3;
        split_f16(v.x, h0, l0); split_f16(v.y, h1, l1); split_f16(v.z, h2, l2); split_f16(v.w, h3, l3);
        const int kk = m * QC + q;
        bh[x][kk][0] = (unsigned)h0 | ((unsigned)h1 << 16); bh[x][kk][1] = (unsigned)h2 | ((unsigned)h3 << 16);
        bl[x][kk][0] = (unsigned)l0 | ((unsigned)l1 << 16); bl[x][kk][1] = (unsigned)l2 | ((unsigned)l3 << 16);
      }
  }
  for (int idx = tid; idx < 2 * NCL * S; idx += H) h_s[idx] = 0u;
  if (tid == 0) {
    mbar_init(&full[0], 1); mbar_init(&full[1], 1);
    mbar_fence_init();
    mbar_expect_tx(&full[0], NCL * CH * 4);   // frame 1 reads buffer 1, frame 2 buffer 0
    mbar_expect_tx(&full[1], NCL * CH * 4);
  }

  // cell role: even thread-in-group -> row g, odd -> row g + 8; unit 2*tile + tig/2
  const int row = g + 8 * (tig & 1), ul = 2 * tn + (tig >> 1);
  const int gb = b0 + row * bstep, unit = u0 + ul;
  const bool valid = row < BS && gb < a.B;
  const int len = valid ? a.lens[gb] : 0;
  float bias[4] = {0.f, 0.f, 0.f, 0.f};
  if (valid) {
#pragma unroll
    for (int gt = 0; gt < 4; ++gt) {
      if (a.b_ih[dir]) bias[gt] += a.b_ih[dir][gt * H + unit];
      if (a.b_hh[dir]) bias[gt] += a.b_hh[dir][gt * H + unit];
    }
  }
  // where this lane's (hi, lo) halves go inside the CTA's slice: [quad][row][hi x4 | lo x4]
  const int st_half = ((ul >> 2) * BS + row) * 8 + (ul & 3);
  float c = 0.f;
  unsigned phase = 0;
  bool dead = false;
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0;
  __syncthreads();
  cluster_arrive();
  cluster_wait();                             // every CTA's barriers are armed before the first remote copy

  float pre[4] = {0.f, 0.f, 0.f, 0.f};
  {
    const int t0 = dir ? T - 1 : 0;
    if (valid && t0 < len) {
      const float* g0 = a.gates + (((long long)gb * T + t0) * ND + dir) * 4 * H + unit;
#pragma unroll
      for (int gt = 0; gt < 4; ++gt) pre[gt] = g0[(long long)gt * H];
    }
  }
  for (int s = 0; s < T; ++s) {
    const int t = dir ? T - 1 - s : s;
    const int p = s & 1;
    const bool active = valid && t < len;
    float* grow = a.gates + (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    float nxt[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int t1 = dir ? t - 1 : t + 1;
      if (s + 1 < T && valid && t1 < len) {
        const float* g1 = a.gates + (((long long)gb * T + t1) * ND + dir) * 4 * H + unit;
#pragma unroll
        for (int gt = 0; gt < 4; ++gt) nxt[gt] = g1[(long long)gt * H];
      }
    }
    const long long c0 = clock64();
    if (s > 0) {                              // h_{t-1}: all 16 slices have landed in h_s[p]
      // a timed-out wait (status word, checked by the caller) must not leave the partner warp at its named barrier
      // or the peers without this CTA's slices: keep walking on whatever is in the buffer
      if (!dead && !mbar_wait_bounded(&full[p], (phase >> p) & 1u)) { atomicExch(a.status, 1); dead = true; }
      phase ^= 1u << p;
      if (tid == 0) mbar_expect_tx(&full[p], NCL * CH * 4);    // re-arm for frame s+2
    }
    const long long c1 = clock64();
    float acc[2][3][4];
#pragma unroll
    for (int x = 0; x < 2; ++x)
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[x][i][j] = 0.f;
    {
      const uint4* hb = reinterpret_cast<const uint4*>(h_s + p * (NCL * S) + (8 * kh + tig) * S) + g;
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int q = 0; q < QC; ++q) {
          const int kk = m * QC + q;
          const uint4 r0 = hb[(m * 4 * S) / 4 + q * BS];
          uint4 r1 = make_uint4(0u, 0u, 0u, 0u);
          if (BS > 8) r1 = hb[(m * 4 * S) / 4 + q * BS + 8];
          const unsigned ah[4] = {r0.x, r1.x, r0.y, r1.y};
          const unsigned al[4] = {r0.z, r1.z, r0.w, r1.w};
#pragma unroll
          for (int x = 0; x < 2; ++x) {
            mma_f16(acc[x][0], ah, bh[x][kk][0], bh[x][kk][1]);
            mma_f16(acc[x][1], ah, bl[x][kk][0], bl[x][kk][1]);
            mma_f16(acc[x][2], al, bh[x][kk][0], bh[x][kk][1]);
          }
        }
    }
    float gsum[4];
    {
      float v[4], o[4];                       // v: the tile this warp finishes, o: its partner's
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float s0 = acc[0][0][j] + (acc[0][1][j] + acc[0][2][j]), s1 = acc[1][0][j] + (acc[1][1][j] + acc[1][2][j]);
        v[j] = kh ? s1 : s0; o[j] = kh ? s0 : s1;
      }
      xch[w][lane] = make_float4(o[0], o[1], o[2], o[3]);
      named_sync(1 + np, 64);
      const float4 pr = xch[(1 - kh) * NP + np][lane];
      v[0] += pr.x; v[1] += pr.y; v[2] += pr.z; v[3] += pr.w;
      const bool odd = tig & 1;
      const float r0 = __shfl_xor_sync(kFull, odd ? v[0] : v[2], 1);
      const float r1 = __shfl_xor_sync(kFull, odd ? v[1] : v[3], 1);
      gsum[0] = odd ? r0 : v[0]; gsum[1] = odd ? r1 : v[1];
      gsum[2] = odd ? v[2] : r0; gsum[3] = odd ? v[3] : r1;
    }
    const long long c2 = clock64();
    float sv[4] = {0.f, 0.f, 0.f, 0.f}, hval = 0.f;
    if (active) {
      const float gi = sigmoid_f(gsum[0] + pre[0] + bias[0]);
      const float gf = sigmoid_f(gsum[1] + pre[1] + bias[1]);
      const float gg = tanh_acc(gsum[2] + pre[2] + bias[2]);
      const float go = sigmoid_f(gsum[3] + pre[3] + bias[3]);
      c = fmaf(gf, c, gi * gg);
      hval = go * tanh_acc(c);
      sv[0] = gi; sv[1] = gf; sv[2] = gg; sv[3] = go;
    } else {
      c = 0.f;
    }
    if (row < BS) {
      unsigned short hi, lo;
      split_f16(hval, hi, lo);
      unsigned short* sh = reinterpret_cast<unsigned short*>(stage + p * CH) + st_half;
      sh[0] = hi; sh[4] = lo;
    }
    if (active) {
      grow[0] = sv[0]; grow[(long long)H] = sv[1]; grow[2ll * H] = sv[2]; grow[3ll * H] = sv[3];
      a.cells[(((long long)gb * T + t) * ND + dir) * H + unit] = c;
    }
    if (valid) a.out[((long long)gb * T + t) * ldo + dir * H + unit] = hval;
    if (s + 1 < T) {
      fence_proxy_async_smem();               // the cell outputs above are read by the async proxy below
      __syncthreads();
      // this CTA's slice of h_t -> chunk `rank` of h_s[1-p] in every CTA of the cluster (two lanes per warp issue)
      const int d = w + (lane & 1) * (H / 32);
      if (lane < 2 && d < NCL)
        bulk_s2cluster(h_s + (p ^ 1) * (NCL * S) + (int)rank * S, stage + p * CH, CH * 4, &full[p ^ 1], (unsigned)d);
    }
#pragma unroll
    for (int gt = 0; gt < 4; ++gt) pre[gt] = nxt[gt];
    const long long c6 = clock64();
    cyc_wait += c1 - c0; cyc_mac += c2 - c1; cyc_rest += c6 - c2;
  }
  cluster_arrive();
  cluster_wait();                             // nobody leaves while a copy may still read or write its shared memory
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {   // CTA 0's cycle split (clock64)
    a.status[4] = (int)cyc_wait; a.status[5] = (int)cyc_mac; a.status[6] = (int)cyc_rest;
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward.  block = H threads: thread j owns column j of this CTA's rows of W_hh (partial dh of unit j, BS rows);
// the first BS*UC threads double as the cell-backward owners of (row, unit).
// ---------------------------------------------------------------------------------------------------------
template <int H, int BS>
__global__ void __launch_bounds__(H + kIssueWarp, 1) lstm_cl_bwd_kernel(const SeqArgs a) {
  constexpr int UC = H / NCL, LR = 4 * UC, CH = BS * UC;
  static_assert(BS % 4 == 0 && BS * UC <= H, "cell owners fit the block");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long full[2];
  __shared__ int lens_s[32];                  // lengths of this cluster's rows (0 past the batch)
  float* dgl = smem;                          // [LR/2][BS][2]  d_gates this CTA produced last frame: lr = gate*UC+uu -> [lr/2][row][lr%2]
  float* Ps = smem + LR * BS;                 // [2][NCL][BS][UC]  partial dh, grouped by owner rank
  float* recv = Ps + 2 * NCL * CH;            // [2][NCL][BS][UC]  partials of MY units from every rank
  const int tid = threadIdx.x, wc = tid < H ? tid : tid - kIssueWarp;   // threads [H, H+32): the issue warp
  const unsigned rank = cluster_ctarank();
  // rows are dealt round-robin to the row groups (row r of group y = batch row y + r*groups): with a batch sorted
  // by length every cluster gets the same mix of long and short rows
  const int dir = blockIdx.z, b0 = blockIdx.y, bstep = gridDim.y, u0 = (int)rank * UC;
  const int T = a.T, ND = a.ND;
  const long long ldo = (long long)ND * H;

  static_assert(UC % 2 == 0 && BS % 2 == 0, "local gate rows and batch rows pair up");
  f32x2 wr2[LR / 2];                          // column `tid` of local gate rows (2i, 2i+1), row lr = gate*UC + uu
  {
    const float* W = a.w_hh[dir];
#pragma unroll
    for (int i = 0; i < LR / 2; ++i) {
      const int g = (2 * i) / UC, uu = (2 * i) % UC;
      wr2[i] = pack2(__ldg(W + ((long long)g * H + u0 + uu) * H + wc), __ldg(W + ((long long)g * H + u0 + uu + 1) * H + wc));
    }
  }
  for (int idx = tid; idx < LR * BS; idx += H) dgl[idx] = 0.f;
  if (tid < 32) lens_s[tid] = (tid < BS && b0 + tid * bstep < a.B) ? a.lens[b0 + tid * bstep] : 0;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
    mbar_expect_tx(&full[1], NCL * CH * 4);   // frame 1 uses buffer 1, frame 2 buffer 0
    mbar_expect_tx(&full[0], NCL * CH * 4);
  }
  const int pdst = (tid / UC) * CH + (tid % UC);              // where column j's partials go inside a Ps buffer
  const bool pw = tid < CH;
  const int row = tid / UC, uu = tid - row * UC;
  const int gb = b0 + row * bstep, unit = u0 + uu;
  const bool valid = pw && gb < a.B;
  const int len = valid ? a.lens[gb] : 0;
  float dc = 0.f;
  unsigned phase = 0;
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0;
  __syncthreads();
  cluster_arrive();
  cluster_wait();

  bool dead = false;                          // a wait timed out: keep walking (status says so), never hang a peer
  constexpr int Q1 = (BS + 2) / 3, NQ = BS / 2;   // row pairs of the first / of both parts of the reduce-scatter
  if (tid >= H) {
    // The first part of the partials (row pairs [0,Q1)) is forwarded by this warp while the compute warps multiply
    // the rest: a thread that issues a cp.async.bulk into a peer is held until the CTA's ~21 B/clk path has taken the
    // copy, which would stall the MACs of its whole warp.
    const int lane = tid - H;
    for (int s = 1; s < T; ++s) {
      const int p = s & 1;
      named_sync(1, H + kIssueWarp);
      if (lane < NCL)
        bulk_s2cluster(recv + p * (NCL * CH) + (int)rank * CH, Ps + p * (NCL * CH) + lane * CH, 2 * Q1 * UC * 4, &full[p],
                       (unsigned)lane);
    }
  } else {
  for (int s = 0; s < T; ++s) {
    const int t = dir ? s : T - 1 - s;
    const int tcp = dir ? t + 1 : t - 1;      // frame whose cell state entered frame t in the forward walk
    const int p = s & 1;
    const bool active = valid && t < len;
    float gi = 0.f, gf = 0.f, gg = 0.f, go = 0.f, ct = 0.f, cp = 0.f, dh = 0.f;
    const long long grow = (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    if (active) {
      const float* g = a.gates + grow;
      gi = __ldg(g); gf = __ldg(g + H); gg = __ldg(g + 2ll * H); go = __ldg(g + 3ll * H);
      ct = __ldg(a.cells + (((long long)gb * T + t) * ND + dir) * H + unit);
      if (tcp >= 0 && tcp < len) cp = __ldg(a.cells + (((long long)gb * T + tcp) * ND + dir) * H + unit);
      dh = __ldg(a.d_out + ((long long)gb * T + t) * ldo + dir * H + unit);
    }
    const long long c0 = clock64();
    long long c1 = c0, c2 = c0;
    if (s > 0) {
      // dgl rows of batch rows that were past their length at the previous frame are zero: skip the pair
      const int tq = dir ? t - 1 : t + 1;
      const unsigned rows_on = __ballot_sync(kFull, tq < lens_s[tid & 31]);
      float* pb = Ps + p * (NCL * CH) + pdst;
      // reduce-scatter in two parts: the partials of the first Q1 row pairs travel while the remaining pairs are
      // still being multiplied (a CTA moves ~21 B/clk into its peers: 15 KB per frame would otherwise sit between
      // the MACs and the cell); the owner of units [d*UC, +UC) gets this CTA's partials for them in slot `rank`
#pragma unroll
      for (int part = 0; part < 2; ++part) {
        const int qa = part ? Q1 : 0, qb = part ? NQ : Q1;
#pragma unroll
        for (int q = qa; q < qb; ++q) {
          f32x2 e = 0ull, o = 0ull;                  // (even, odd local gate row) partial sums of rows 2q, 2q+1
          if ((rows_on >> (2 * q)) & 3u) {
#pragma unroll
            for (int i = 0; i < LR / 2; ++i) {
              const ulonglong2 d = *reinterpret_cast<const ulonglong2*>(dgl + (i * BS + 2 * q) * 2);
              e = ffma2(d.x, wr2[i], e);
              o = ffma2(d.y, wr2[i], o);
            }
          }
          pb[(2 * q) * UC] = sum2(e);
          pb[(2 * q + 1) * UC] = sum2(o);
        }
        fence_proxy_async_smem();
        if (part == 0) {
          named_arrive(1, H + kIssueWarp);         // the issue warp forwards this part while the MACs go on
        } else {
          named_sync(3, H);                        // the last part: every compute warp issues one or two copies
          c1 = clock64();
          {                                        // two lanes per warp, one copy each (see the forward kernel)
            const int off = 2 * Q1 * UC, bytes = 2 * (NQ - Q1) * UC * 4;
            const int lane = tid & 31, d = (tid >> 5) + lane * (H / 32);
            if (lane < 2 && d < NCL)
              bulk_s2cluster(recv + p * (NCL * CH) + (int)rank * CH + off, Ps + p * (NCL * CH) + d * CH + off, bytes, &full[p],
                             (unsigned)d);
          }
        }
      }
      if (!dead && !mbar_wait_bounded(&full[p], (phase >> p) & 1u)) { atomicExch(a.status, 1); dead = true; }
      phase ^= 1u << p;
      if (tid == 0) mbar_expect_tx(&full[p], NCL * CH * 4);     // re-arm for frame s+2
      c2 = clock64();
      if (pw) {
        const float* rb = recv + p * (NCL * CH) + row * UC + uu;
        float sum = 0.f;
#pragma unroll
        for (int src = 0; src < NCL; ++src) sum += rb[src * CH];
        dh += sum;
      }
    }
    if (pw) {
      float di = 0.f, df = 0.f, dgg = 0.f, dob = 0.f;
      if (active) {
        const float tc = tanh_acc(ct);
        const float dct = dc + dh * go * (1.f - tc * tc);
        di = dct * gg * gi * (1.f - gi);
        df = dct * cp * gf * (1.f - gf);
        dgg = dct * gi * (1.f - gg * gg);
        dob = dh * tc * go * (1.f - go);
        dc = dct * gf;
      } else {
        dc = 0.f;
      }
      {
        float* dq = dgl + ((uu >> 1) * BS + row) * 2 + (uu & 1);          // local gate row lr = gate*UC + uu
        dq[0] = di; dq[(UC / 2) * BS * 2] = df; dq[UC * BS * 2] = dgg; dq[(3 * UC / 2) * BS * 2] = dob;
      }
      if (valid) {
        float* o = a.d_gates + grow;
        o[0] = di; o[(long long)H] = df; o[2ll * H] = dgg; o[3ll * H] = dob;
      }
    }
    named_sync(3, H);                         // compute warps only: dgl is complete before the next frame's MACs
    cyc_mac += c1 - c0; cyc_wait += c2 - c1; cyc_rest += clock64() - c2;
  }
  }
  cluster_arrive();
  cluster_wait();                             // nobody leaves while a copy may still read or write its shared memory
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    a.status[8] = (int)cyc_wait; a.status[9] = (int)cyc_mac; a.status[10] = (int)cyc_rest;
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward, tensor-core variant (default; B2ASR_ENC_MMA=0 selects the FFMA2 kernel above).
//
// partial dh[16 x H] = dG_slice[16 x 4UC] . W_slice[4UC x H] on mma.sync.m16n8k16 (fp16 hi/lo split of both operands,
// three products per k-step, fp32 accumulation -- see the forward kernel).  Warp w owns columns [32w, 32w+32) (four
// n-tiles) over all of K = 4UC; its W fragments stay in registers (80 at H=320).  Gate gradients span many orders of
// magnitude, fp16 does not: every batch row is multiplied by an exact power of two that brings ITS largest |dG| of the
// frame to [2^13, 2^14) before the split (the owners of the row publish that maximum with one shared-memory atomicMax
// each) and the product row is multiplied back, so the result carries 22 bits relative to the row's own scale and no
// row can push another one out of range.  A (16 x 80 fp32, this CTA's own d_gates of the last frame) is converted at
// fragment load -- it is small, ten 128-bit loads per warp and frame.  The exchange is the same reduce-scatter as in
// the FFMA2 kernel, in one part: all rows of a tile finish together, the extra warp issues the 16 copies.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void split_f16x2(float x0, float x1, unsigned& hi2, unsigned& lo2) {
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(hi2) : "f"(x1), "f"(x0));           // lower half = x0
  unsigned short h0, h1;
  asm("mov.b32 {%0, %1}, %2;" : "=h"(h0), "=h"(h1) : "r"(hi2));
  const float l0 = x0 - f16_value(h0), l1 = x1 - f16_value(h1);
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(lo2) : "f"(l1), "f"(l0));
}

template <int H, int BS>
__global__ void __launch_bounds__(H + kIssueWarp, 1) lstm_cl_bwd_mma_kernel(const SeqArgs a) {
  constexpr int UC = H / NCL, LR = 4 * UC, CH = BS * UC;
  constexpr int KSB = LR / 16;                // k16-steps
  constexpr int RSA = LR % 32 == 16 ? LR : LR + 16;   // A row stride (words): rows 2j, 2j+1 of a phase 16 banks apart
  static_assert(BS % 4 == 0 && BS * UC <= H && LR % 16 == 0 && UC % 2 == 0, "cell owners fit the block; whole k-steps");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long full[2];
  __shared__ unsigned rmax[2][16];            // per frame parity: bits of the largest |dG| of every row
  float* dga = smem;                          // [16][RSA]  d_gates this CTA produced last frame, k = gate*UC + uu
  float* Ps = smem + 16 * RSA;                // [2][NCL][BS][UC]  partial dh, grouped by owner rank
  float* recv = Ps + 2 * NCL * CH;            // [2][NCL][BS][UC]  partials of MY units from every rank
  const int tid = threadIdx.x;
  const unsigned rank = cluster_ctarank();
  const int dir = blockIdx.z, b0 = blockIdx.y, bstep = gridDim.y, u0 = (int)rank * UC;
  const int T = a.T, ND = a.ND;
  const long long ldo = (long long)ND * H;
  const int lane = tid & 31, w = (tid >> 5) % (H / 32), g = lane >> 2, tig = lane & 3;

  // B fragments: B[k][n] = W_hh[gate(k)*H + u0 + uu(k)][32w + 8x + n], k-step kk holds k = 16kk + 4tig + {0..3}
  unsigned bh[4][KSB][2], bl[4][KSB][2];
  if (tid < H) {
    const float* W = a.w_hh[dir];
#pragma unroll
    for (int x = 0; x < 4; ++x)
#pragma unroll
      for (int kk = 0; kk < KSB; ++kk) {
        float v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k = 16 * kk + 4 * tig + i;
          v[i] = __ldg(W + ((long long)(k / UC) * H + u0 + (k % UC)) * H + 32 * w + 8 * x + g);
        }
        split_f16x2(v[0], v[1], bh[x][kk][0], bl[x][kk][0]);
        split_f16x2(v[2], v[3], bh[x][kk][1], bl[x][kk][1]);
      }
  }
  for (int idx = tid; idx < 16 * RSA; idx += H + kIssueWarp) dga[idx] = 0.f;
  if (tid < 32) { rmax[0][tid & 15] = 0u; rmax[1][tid & 15] = 0u; }
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
    mbar_expect_tx(&full[1], NCL * CH * 4);   // frame 1 uses buffer 1, frame 2 buffer 0
    mbar_expect_tx(&full[0], NCL * CH * 4);
  }
  const bool pw = tid < CH;
  const int row = tid / UC, uu = tid - row * UC;
  const int gb = b0 + row * bstep, unit = u0 + uu;
  const bool valid = pw && gb < a.B;
  const int len = valid ? a.lens[gb] : 0;
  float dc = 0.f;
  unsigned phase = 0;
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0, cyc_r0 = 0, cyc_r1 = 0, cyc_r2 = 0, cyc_r3 = 0;
  __syncthreads();
  cluster_arrive();
  cluster_wait();

  bool dead = false;                          // a wait timed out: keep walking (status says so), never hang a peer
  if (tid >= H) {
    // the reduce-scatter's copies are issued by this warp: a thread that issues a cp.async.bulk into a peer is held
    // until the CTA's ~21 B/clk path has taken the copy
    for (int s = 1; s < T; ++s) {
      const int p = s & 1;
      named_sync(1, H + kIssueWarp);
      if (lane < NCL)
        bulk_s2cluster(recv + p * (NCL * CH) + (int)rank * CH, Ps + p * (NCL * CH) + lane * CH, CH * 4, &full[p], (unsigned)lane);
    }
  } else {
  for (int s = 0; s < T; ++s) {
    const int t = dir ? s : T - 1 - s;
    const int tcp = dir ? t + 1 : t - 1;      // frame whose cell state entered frame t in the forward walk
    const int p = s & 1;
    const bool active = valid && t < len;
    float gi = 0.f, gf = 0.f, gg = 0.f, go = 0.f, ct = 0.f, cp = 0.f, dh = 0.f;
    const long long grow = (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    if (active) {
      const float* gp = a.gates + grow;
      gi = __ldg(gp); gf = __ldg(gp + H); gg = __ldg(gp + 2ll * H); go = __ldg(gp + 3ll * H);
      ct = __ldg(a.cells + (((long long)gb * T + t) * ND + dir) * H + unit);
      if (tcp >= 0 && tcp < len) cp = __ldg(a.cells + (((long long)gb * T + tcp) * ND + dir) * H + unit);
      dh = __ldg(a.d_out + ((long long)gb * T + t) * ldo + dir * H + unit);
    }
    const long long c0 = clock64();
    long long c1 = c0, c2 = c0;
    if (s > 0) {
      // the rows' scales: 2^(13 - floor(log2 max|dG|)) and its inverse, both exact powers of two
      float sc[2], inv[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const unsigned m = rmax[p ^ 1][g + 8 * h];
        int sf = 267 - (int)(m >> 23);
        sf = m == 0u ? 127 : min(253, max(1, sf));
        sc[h] = __uint_as_float((unsigned)sf << 23);
        inv[h] = __uint_as_float((unsigned)(254 - sf) << 23);
      }
      float acc[4][2][4];
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[x][i][j] = 0.f;
      const float4* ap = reinterpret_cast<const float4*>(dga + g * RSA) + tig;
#pragma unroll
      for (int kk = 0; kk < KSB; ++kk) {
        const float4 r0 = ap[4 * kk];
        float4 r1 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (BS > 8) r1 = ap[4 * kk + 8 * RSA / 4];
        unsigned ah[4], al[4];
        split_f16x2(r0.x * sc[0], r0.y * sc[0], ah[0], al[0]);
        split_f16x2(r1.x * sc[1], r1.y * sc[1], ah[1], al[1]);
        split_f16x2(r0.z * sc[0], r0.w * sc[0], ah[2], al[2]);
        split_f16x2(r1.z * sc[1], r1.w * sc[1], ah[3], al[3]);
#pragma unroll
        for (int x = 0; x < 4; ++x) {
          mma_f16(acc[x][0], ah, bh[x][kk][0], bh[x][kk][1]);
          mma_f16(acc[x][1], ah, bl[x][kk][0], bl[x][kk][1]);
          mma_f16(acc[x][1], al, bh[x][kk][0], bh[x][kk][1]);
        }
      }
      // partials -> Ps, grouped by owner rank: columns (c, c+1) of rows g and g+8
      float* pb = Ps + p * (NCL * CH);
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const int cidx = 32 * w + 8 * x + 2 * tig, d = cidx / UC, cu = cidx - d * UC;
        if (g < BS)
          *reinterpret_cast<float2*>(pb + d * CH + g * UC + cu) =
              make_float2((acc[x][0][0] + acc[x][1][0]) * inv[0], (acc[x][0][1] + acc[x][1][1]) * inv[0]);
        if (g + 8 < BS)
          *reinterpret_cast<float2*>(pb + d * CH + (g + 8) * UC + cu) =
              make_float2((acc[x][0][2] + acc[x][1][2]) * inv[1], (acc[x][0][3] + acc[x][1][3]) * inv[1]);
      }
      fence_proxy_async_smem();
      named_arrive(1, H + kIssueWarp);          // the issue warp sends the 16 slices
      c1 = clock64();
      if (!dead && !mbar_wait_bounded(&full[p], (phase >> p) & 1u)) { atomicExch(a.status, 1); dead = true; }
      phase ^= 1u << p;
      if (tid == 0) mbar_expect_tx(&full[p], NCL * CH * 4);     // re-arm for frame s+2
      c2 = clock64();
      if (pw) {
        const float* rb = recv + p * (NCL * CH) + row * UC + uu;
        float sum = 0.f;
#pragma unroll
        for (int src = 0; src < NCL; ++src) sum += rb[src * CH];
        dh += sum;
      }
    }
    const long long c3 = clock64();
    long long c4 = c3, c5 = c3;
    if (tid < 16) rmax[p ^ 1][tid] = 0u;        // every warp has read it (own copies issued => all MACs of the frame done)
    if (pw) {
      float di = 0.f, df = 0.f, dgg = 0.f, dob = 0.f;
      if (active) {
        const float tc = tanh_acc(ct);
        const float dct = dc + dh * go * (1.f - tc * tc);
        di = dct * gg * gi * (1.f - gi);
        df = dct * cp * gf * (1.f - gf);
        dgg = dct * gi * (1.f - gg * gg);
        dob = dh * tc * go * (1.f - go);
        dc = dct * gf;
      } else {
        dc = 0.f;
      }
      c4 = clock64();
      {
        float* dq = dga + row * RSA + uu;
        dq[0] = di; dq[UC] = df; dq[2 * UC] = dgg; dq[3 * UC] = dob;
        const float mx = fmaxf(fmaxf(fabsf(di), fabsf(df)), fmaxf(fabsf(dgg), fabsf(dob)));
        if (mx > 0.f) atomicMax(&rmax[p][row], __float_as_uint(mx));
      }
      c5 = clock64();
      if (valid) {
        float* o = a.d_gates + grow;
        o[0] = di; o[(long long)H] = df; o[2ll * H] = dgg; o[3ll * H] = dob;
      }
    }
    const long long c6 = clock64();
    named_sync(3, H);                         // compute warps only: dga and the row maxima are complete before the next frame
    const long long c7 = clock64();
    cyc_mac += c1 - c0; cyc_wait += c2 - c1; cyc_rest += c7 - c2;
    cyc_r0 += c4 - c2; cyc_r1 += c5 - c4; cyc_r2 += c6 - c5; cyc_r3 += c7 - c6;
  }
  }
  cluster_arrive();
  cluster_wait();                             // nobody leaves while a copy may still read or write its shared memory
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
    a.status[8] = (int)cyc_wait; a.status[9] = (int)cyc_mac; a.status[10] = (int)cyc_rest;
    a.status[11] = (int)cyc_r0; a.status[12] = (int)cyc_r1; a.status[13] = (int)cyc_r2; a.status[14] = (int)cyc_r3;
  }
}

template <int H, int BS>
int launch(const SeqArgs& a, bool backward, cudaStream_t st, bool probe_only, int* fit_out) {
  constexpr int UC = H / NCL;
  static const bool mma = [] { const char* e = getenv("B2ASR_ENC_MMA"); return !(e && e[0] == '0'); }();
  static const bool mma_bwd = [] { const char* e = getenv("B2ASR_ENC_MMA_BWD"); return !(e && e[0] == '0'); }();
  constexpr int LR = 4 * UC, RSA = LR % 32 == 16 ? LR : LR + 16;
  void (*fn)(const SeqArgs) = backward ? (mma && mma_bwd ? lstm_cl_bwd_mma_kernel<H, BS> : lstm_cl_bwd_kernel<H, BS>)
                              : mma    ? lstm_cl_fwd_mma_kernel<H, BS> : lstm_cl_fwd_kernel<H, BS>;
  const size_t smem = backward ? (mma && mma_bwd ? (size_t)(16 * RSA + 4 * NCL * BS * UC) * sizeof(float)
                                                 : (size_t)(4 * UC * BS + 4 * NCL * BS * UC) * sizeof(float))
                      : mma    ? (size_t)(2 * NCL * mma_chunk_stride(UC, BS) + 2 * BS * UC) * sizeof(float)
                               : (size_t)(2 * NCL * chunk_stride(UC, BS) + 2 * BS * UC) * sizeof(float);
  B2_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  B2_CUDA_TRY(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(NCL, (a.B + BS - 1) / BS, a.ND);
  cfg.blockDim = dim3(backward ? H + kIssueWarp : H);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = NCL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int fit = 0;
  if (cudaOccupancyMaxActiveClusters(&fit, (const void*)fn, &cfg) != cudaSuccess || fit < 1) {
    (void)cudaGetLastError();
    return B2ASR_EUNSUPPORTED;               // e.g. a partition without a 16-SM GPC: caller takes the L2 path
  }
  *fit_out = fit;
  if (probe_only) return B2ASR_OK;
  static const bool dbg = getenv("B2ASR_ENC_DEBUG") != nullptr;
  if (dbg) fprintf(stderr, "lstm_cluster: H=%d BS=%d %s grid=(%d,%d,%d) max active clusters %d\n", H, BS,
                   backward ? "bwd" : "fwd", cfg.gridDim.x, cfg.gridDim.y, cfg.gridDim.z, fit);
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, fn, a));
  return B2ASR_OK;
}

template <int H>
int run_cluster(const SeqArgs& a, bool backward, cudaStream_t st) {
  // rows per cluster: the smallest of {8, 12, 16} whose clusters all fit in one wave (B2ASR_ENC_BS overrides)
  static const int forced = [] { const char* e = getenv("B2ASR_ENC_BS"); return e ? atoi(e) : 0; }();
  int fit = 0;
  int rc = launch<H, 8>(a, backward, st, true, &fit);
  if (rc != B2ASR_OK) return rc;
  int bs = 16;
  for (int cand : {8, 12, 16})
    if (((a.B + cand - 1) / cand) * a.ND <= fit) { bs = cand; break; }
  if (forced == 8 || forced == 12 || forced == 16) bs = forced;
  switch (bs) {
    case 8: return launch<H, 8>(a, backward, st, false, &fit);
    case 12: return launch<H, 12>(a, backward, st, false, &fit);
    default: return launch<H, 16>(a, backward, st, false, &fit);
  }
}

}  // namespace

bool lstm_cluster_supported(int H) { return H == 256 || H == 320; }

int lstm_cluster_run(const SeqArgs& a, bool backward, cudaStream_t st) {
  switch (a.H) {
    case 256: return run_cluster<256>(a, backward, st);
    case 320: return run_cluster<320>(a, backward, st);
    default: return B2ASR_EUNSUPPORTED;
  }
}

}  // namespace b2asr
