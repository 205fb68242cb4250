// Listener recurrence, SURVEY.md section 8(f) "callers either side of the path": the (bi)LSTM layers of the
// encoder that produce the features the attention keys and the CTC head read (reference src/asr.py:466-501
// RNNLayer.forward = pack_padded_sequence -> nn.LSTM -> pad_packed_sequence, stepped by Listener.forward
// asr.py:305-317).  With packed variable-length input cuDNN runs the recurrence as ~8 small launches per frame
// (28 600 launches, 122 ms of the 197 ms M2 train step).  Here the recurrence of one layer -- every frame, both
// directions -- is ONE persistent cooperative launch forward and ONE backward:
//
//   * the input projection x W_ih^T for all frames is a single GEMM on the caller's side (gates buffer
//     [B,T,ND,4H] arrives holding it); weight gradients and dx are single GEMMs over the d_gates buffer afterwards;
//   * W_hh is split by hidden unit over NC = ceil(H/U) CTAs per (direction, group of BS batch rows); a CTA keeps
//     its 4U gate rows (forward) / U columns (backward) in shared memory for the whole sequence and its cell
//     state / cell gradient in registers;
//   * per frame a CTA needs the whole previous h (forward, H floats per row) or the whole previous d_gates
//     (backward, 4H floats per row).  They are exchanged through the output tensors themselves in L2: the call
//     pre-fills them with a sentinel (0xFFFFFFFF, a NaN no arithmetic here produces), owners store their slice
//     with st.global.cg, readers pull rows with 1-D bulk (TMA) copies and pull them again while a sentinel word
//     sits in the staged rows (= some peer's slice has not landed yet; the scan is an unsigned max over the
//     words).  No flags, no atomics, no fences: every 32-bit word is its own ready bit.  Two CTAs of different row groups share an SM, so one group's exchange
//     latency hides under the other's FMAs;
//   * rows past their length write zeros (pad_packed_sequence semantics) and keep a zero state, so the reverse
//     direction starts exactly at each row's last valid frame.
#include <stdlib.h>

#include "enc.cuh"

namespace b2asr {

namespace {

constexpr unsigned kSentinelBits = 0xffffffffu;
constexpr int kSpinLimit = 400000;          // bulk-copy retries per frame before the launch gives up (~0.5 s)

__device__ __forceinline__ void st_cg(float* p, float v) { asm volatile("st.global.cg.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }
__device__ __forceinline__ int ld_status(const int* p) {
  int v; asm volatile("ld.volatile.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
// ---------------------------------------------------------------------------------------------------------
// forward: grid (NC, row groups, ND), block 8 k-slices x BS/4 row quads x U units
// ---------------------------------------------------------------------------------------------------------
template <int U, int BS>
__global__ void __launch_bounds__(8 * (BS / 4) * U) lstm_seq_fwd_kernel(const SeqArgs a) {
  static_assert(BS % 4 == 0 && BS <= 32, "one bulk copy per lane of warp 0");
  constexpr int NBG = BS / 4;
  constexpr int NTHR = 8 * NBG * U;
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long bar;
  const int H = a.H, Kp = a.Kp, T = a.T, ND = a.ND;
  float* W_s = smem;                         // [U*4][Kp]   row = unit*4 + gate
  float* h_s = smem + 4 * U * Kp;            // [BS][Kp]
  const int tid = threadIdx.x, ks = tid & 7, grp = tid >> 3;
  const int u = grp / NBG, bg = grp - u * NBG;
  const int dir = blockIdx.z, u0 = blockIdx.x * U, b0 = blockIdx.y * BS;
  const long long ldo = (long long)ND * H;

  const float* W = a.w_hh[dir];
  for (int idx = tid; idx < 4 * U * Kp; idx += NTHR) {
    const int r = idx / Kp, k = idx - r * Kp, uu = r >> 2, g = r & 3;
    W_s[idx] = (u0 + uu < H && k < H) ? W[((long long)g * H + u0 + uu) * H + k] : 0.f;
  }
  for (int idx = tid; idx < BS * Kp; idx += NTHR) h_s[idx] = 0.f;
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }

  // pointwise role: even k-slice lanes own (row, unit) = (b0 + bg*4 + ks/2, u0 + u)
  const int gb = b0 + bg * 4 + (ks >> 1), unit = u0 + u;
  const bool pw = ((ks & 1) == 0) && gb < a.B && unit < H;
  const int len = pw ? a.lens[gb] : 0;
  float bias[4] = {0.f, 0.f, 0.f, 0.f};
  if (pw) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (a.b_ih[dir]) bias[g] += a.b_ih[dir][g * H + unit];
      if (a.b_hh[dir]) bias[g] += a.b_hh[dir][g * H + unit];
    }
  }
  const int rows = min(BS, a.B - b0);        // valid rows of this group (> 0 by construction of the grid)
  float c = 0.f;
  unsigned phase = 0;
  int retries = 0;
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0;
  __syncthreads();

  for (int s = 0; s < T; ++s) {
    const int t = dir ? T - 1 - s : s;
    const int tp = dir ? t + 1 : t - 1;
    const bool active = pw && t < len;
    float pre[4] = {0.f, 0.f, 0.f, 0.f};
    float* grow = a.gates + (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    if (active) {
#pragma unroll
      for (int g = 0; g < 4; ++g) pre[g] = grow[(long long)g * H];
    }
    float acc[4][4];
    const long long c0 = clock64();
    if (s > 0) {
      int tries = 0;
      while (true) {
        if (tid < 32) {                       // warp 0 posts the transaction count, then one row copy per lane
          if (tid == 0) mbar_expect_tx(&bar, (unsigned)(rows * H * 4));
          __syncwarp();
          if (tid < rows)
            bulk_g2s(h_s + tid * Kp, a.out + ((long long)(b0 + tid) * T + tp) * ldo + dir * H, (unsigned)(H * 4), &bar);
        }
        mbar_wait(&bar, phase);
        phase ^= 1u;
        // a sentinel word anywhere in the staged rows = some peer's slice of h has not landed in L2 yet
        unsigned m = 0u;
        for (int idx = tid; idx < BS * Kp / 4; idx += NTHR) {
          const uint4 q = reinterpret_cast<const uint4*>(h_s)[idx];
          m = max(m, max(max(q.x, q.y), max(q.z, q.w)));
        }
        if (!__syncthreads_or(m == kSentinelBits)) break;
        ++tries;
        if ((tries & 63) == 0) {              // now and then: did a peer give up, or should we?
          const int st = (tid == 0) ? ld_status(a.status) : 0;
          if (__syncthreads_or(st != 0 || tries > kSpinLimit)) {
            if (tid == 0) atomicExch(a.status, 1);
            return;
          }
        }
      }
      retries += tries;
    }
    const long long c1 = clock64();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int g = 0; g < 4; ++g) acc[i][g] = 0.f;
    {
      const float* hrow = h_s + (bg * 4) * Kp + 4 * ks;
      const float* wrow = W_s + (u * 4) * Kp + 4 * ks;
      for (int j = 0; j < Kp; j += 32) {
        float4 hv[4], wv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) hv[i] = *reinterpret_cast<const float4*>(hrow + i * Kp + j);
#pragma unroll
        for (int g = 0; g < 4; ++g) wv[g] = *reinterpret_cast<const float4*>(wrow + g * Kp + j);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int g = 0; g < 4; ++g) acc[i][g] = dot4(hv[i], wv[g], acc[i][g]);
      }
    }
    __syncthreads();                          // every read of h_s is done before the next frame's copy lands
    const long long c2 = clock64();
    // transposing butterfly over the 8 k-slices: 16 -> 8 -> 4 values, then a pair all-reduce
    {
      const bool hi = ks & 4;
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const float send = hi ? acc[i][g] : acc[i + 2][g];
          const float keep = hi ? acc[i + 2][g] : acc[i][g];
          acc[i][g] = keep + __shfl_xor_sync(kFull, send, 4);
        }
    }
    {
      const bool hi = ks & 2;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const float send = hi ? acc[0][g] : acc[1][g];
        const float keep = hi ? acc[1][g] : acc[0][g];
        acc[0][g] = keep + __shfl_xor_sync(kFull, send, 2);
      }
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) acc[0][g] += __shfl_xor_sync(kFull, acc[0][g], 1);
    if (pw) {
      float h = 0.f;
      if (active) {
        const float gi = sigmoid_f(acc[0][0] + pre[0] + bias[0]);
        const float gf = sigmoid_f(acc[0][1] + pre[1] + bias[1]);
        const float gg = tanh_acc(acc[0][2] + pre[2] + bias[2]);
        const float go = sigmoid_f(acc[0][3] + pre[3] + bias[3]);
        c = fmaf(gf, c, gi * gg);
        h = go * tanh_acc(c);
        grow[0] = gi; grow[(long long)H] = gf; grow[2ll * H] = gg; grow[3ll * H] = go;
        a.cells[(((long long)gb * T + t) * ND + dir) * H + unit] = c;
      } else {
        c = 0.f;
      }
      st_cg(a.out + ((long long)gb * T + t) * ldo + dir * H + unit, h);
    }
    cyc_wait += c1 - c0; cyc_mac += c2 - c1; cyc_rest += clock64() - c2;
  }
  if (tid == 0 && retries) atomicAdd(a.status + 1, retries);      // diagnostics: failed exchange polls
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {   // CTA 0's cycle split (clock64)
    a.status[4] = (int)cyc_wait; a.status[5] = (int)cyc_mac; a.status[6] = (int)cyc_rest;
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward: grid (NC, row groups, ND), block = BS/4 warps; warp w owns rows 4w..4w+3 of the group and needs no
// CTA-wide barrier after set-up (its rows of d_gates are staged and read by that warp alone)
// ---------------------------------------------------------------------------------------------------------
template <int U, int BS>
__global__ void __launch_bounds__(32 * (BS / 4)) lstm_seq_bwd_kernel(const SeqArgs a) {
  constexpr int NW = BS / 4;
  constexpr int NTHR = 32 * NW;
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) unsigned long long bars[NW];
  const int H = a.H, Kp = a.Kp, T = a.T, ND = a.ND, H4 = 4 * a.H;
  float* Wt_s = smem;                        // [U][Kp]   Wt_s[uu][r] = W_hh[r][u0+uu]
  float* dg_s = smem + U * Kp;               // [BS][Kp]
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int dir = blockIdx.z, u0 = blockIdx.x * U, b0 = blockIdx.y * BS;
  const long long ldo = (long long)ND * H;

  const float* W = a.w_hh[dir];
  for (int idx = tid; idx < U * Kp; idx += NTHR) {
    const int uu = idx / Kp, r = idx - uu * Kp;
    Wt_s[idx] = (u0 + uu < H && r < H4) ? W[(long long)r * H + u0 + uu] : 0.f;
  }
  for (int idx = tid; idx < BS * Kp; idx += NTHR) dg_s[idx] = 0.f;
  if (lane == 0) mbar_init(&bars[w], 1);
  if (tid == 0) mbar_fence_init();
  __syncthreads();

  const int wrow0 = b0 + 4 * w;                                  // first global row of this warp
  const int wrows = max(0, min(4, a.B - wrow0));                 // valid rows of this warp
  const int bi = lane >> 3, uu = lane & 7;
  const int gb = wrow0 + bi, unit = u0 + uu;
  const bool pw = uu < U && unit < H && gb < a.B;
  const int len = pw ? a.lens[gb] : 0;
  if (wrows == 0) return;                                        // no barrier below: safe
  float dc = 0.f;
  unsigned phase = 0;
  int retries = 0;
  long long cyc_wait = 0, cyc_mac = 0, cyc_rest = 0;
  float* mydg = dg_s + (4 * w) * Kp;

  for (int s = 0; s < T; ++s) {
    const int t = dir ? s : T - 1 - s;
    const int tq = dir ? t - 1 : t + 1;       // frame handled one step earlier in this walk
    const int tcp = dir ? t + 1 : t - 1;      // frame whose cell state entered frame t in the forward walk
    const bool active = pw && t < len;
    float gi = 0.f, gf = 0.f, gg = 0.f, go = 0.f, ct = 0.f, cp = 0.f, dh = 0.f;
    const long long grow = (((long long)gb * T + t) * ND + dir) * 4 * H + unit;
    if (active) {
      const float* g = a.gates + grow;
      gi = __ldg(g); gf = __ldg(g + H); gg = __ldg(g + 2ll * H); go = __ldg(g + 3ll * H);
      ct = __ldg(a.cells + (((long long)gb * T + t) * ND + dir) * H + unit);
      if (tcp >= 0 && tcp < len) cp = __ldg(a.cells + (((long long)gb * T + tcp) * ND + dir) * H + unit);
      dh = __ldg(a.d_out + ((long long)gb * T + t) * ldo + dir * H + unit);
    }
    float v[32];
    const long long c0 = clock64();
    if (s > 0) {
      int tries = 0;
      while (true) {
        if (lane == 0) mbar_expect_tx(&bars[w], (unsigned)(wrows * H4 * 4));
        __syncwarp();
        if (lane < wrows)
          bulk_g2s(mydg + lane * Kp, a.d_gates + (((long long)(wrow0 + lane) * T + tq) * ND + dir) * 4 * H,
                   (unsigned)(H4 * 4), &bars[w]);
        mbar_wait(&bars[w], phase);
        phase ^= 1u;
        unsigned m = 0u;                      // sentinel scan of this warp's four staged rows
        for (int idx = lane; idx < Kp; idx += 32) {
          const uint4 q = reinterpret_cast<const uint4*>(mydg)[idx];
          m = max(m, max(max(q.x, q.y), max(q.z, q.w)));
        }
        if (!__any_sync(kFull, m == kSentinelBits)) break;
        ++tries;
        if ((tries & 63) == 0) {
          if (ld_status(a.status) != 0 || tries > kSpinLimit) {
            if (lane == 0) atomicExch(a.status, 1);
            return;
          }
        }
      }
      retries += tries;
    }
    const long long c1 = clock64();
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0.f;
    for (int j = 4 * lane; j < Kp; j += 128) {
      float4 dv[4], wv[U];
#pragma unroll
      for (int i = 0; i < 4; ++i) dv[i] = *reinterpret_cast<const float4*>(mydg + i * Kp + j);
#pragma unroll
      for (int q = 0; q < U; ++q) wv[q] = *reinterpret_cast<const float4*>(Wt_s + q * Kp + j);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int q = 0; q < U; ++q) v[i * 8 + q] = dot4(dv[i], wv[q], v[i * 8 + q]);
    }
    // transposing butterfly over the 32 lanes: slot (row quad index, unit) = lane afterwards
#pragma unroll
    for (int m = 16, n = 16; m >= 1; m >>= 1, n >>= 1) {
      const bool hi = lane & m;
#pragma unroll
      for (int i = 0; i < n; ++i) {
        if ((i & 7) < U || ((i + n) & 7) < U) {                // slots of units >= U stay zero
          const float send = hi ? v[i] : v[i + n];
          const float keep = hi ? v[i + n] : v[i];
          v[i] = keep + __shfl_xor_sync(kFull, send, m);
        }
      }
    }
    __syncwarp();                             // every lane is done reading the staged rows before the next copy
    const long long c2 = clock64();
    if (pw) {
      float di = 0.f, df = 0.f, dgg = 0.f, dob = 0.f;
      if (active) {
        dh += v[0];
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
      float* o = a.d_gates + grow;
      st_cg(o, di); st_cg(o + H, df); st_cg(o + 2ll * H, dgg); st_cg(o + 3ll * H, dob);
    }
    cyc_wait += c1 - c0; cyc_mac += c2 - c1; cyc_rest += clock64() - c2;
  }
  if (lane == 0 && retries) atomicAdd(a.status + 2, retries);     // diagnostics: failed exchange polls
  if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {   // warp 0 of CTA 0: cycle split
    a.status[8] = (int)cyc_wait; a.status[9] = (int)cyc_mac; a.status[10] = (int)cyc_rest;
  }
}

struct Plan {
  int U, NC, groups_per_launch;
  size_t smem;
};

template <int U, int BS>
int launch_fwd(const SeqArgs& base, int row0, int rows, int groups, size_t smem, cudaStream_t st) {
  SeqArgs a = base;
  a.B = rows;
  a.gates += (long long)row0 * a.T * a.ND * 4 * a.H;
  a.out += (long long)row0 * a.T * a.ND * a.H;
  a.cells += (long long)row0 * a.T * a.ND * a.H;
  a.lens += row0;
  void* params[] = {(void*)&a};
  dim3 grid((a.H + U - 1) / U, groups, a.ND), block(8 * (BS / 4) * U);
  B2_CUDA_TRY(cudaLaunchCooperativeKernel((void*)lstm_seq_fwd_kernel<U, BS>, grid, block, params, smem, st));
  return B2ASR_OK;
}

template <int U, int BS>
int launch_bwd(const SeqArgs& base, int row0, int rows, int groups, size_t smem, cudaStream_t st) {
  SeqArgs a = base;
  a.B = rows;
  a.gates += (long long)row0 * a.T * a.ND * 4 * a.H;
  a.cells += (long long)row0 * a.T * a.ND * a.H;
  a.d_out += (long long)row0 * a.T * a.ND * a.H;
  a.d_gates += (long long)row0 * a.T * a.ND * 4 * a.H;
  a.lens += row0;
  void* params[] = {(void*)&a};
  dim3 grid((a.H + U - 1) / U, groups, a.ND), block(32 * (BS / 4));
  B2_CUDA_TRY(cudaLaunchCooperativeKernel((void*)lstm_seq_bwd_kernel<U, BS>, grid, block, params, smem, st));
  return B2ASR_OK;
}

constexpr int kBS = 16;

template <int U>
int run(const SeqArgs& base, bool backward, cudaStream_t st) {
  int dev = 0, sms = 0, coop = 0;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  B2_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  B2_CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  B2_REQUIRE(coop, B2ASR_EUNSUPPORTED, "lstm_seq: device has no cooperative launch");
  SeqArgs a = base;
  const void* fn = backward ? (const void*)lstm_seq_bwd_kernel<U, kBS> : (const void*)lstm_seq_fwd_kernel<U, kBS>;
  const int threads = backward ? 32 * (kBS / 4) : 8 * (kBS / 4) * U;
  size_t smem;
  if (backward) {
    a.Kp = (int)align_up((size_t)4 * a.H, 128);
    smem = (size_t)(U + kBS) * a.Kp * sizeof(float);
  } else {
    a.Kp = (int)align_up((size_t)a.H, 32);
    smem = (size_t)(4 * U + kBS) * a.Kp * sizeof(float);
  }
  B2_REQUIRE(smem <= 227 * 1024, B2ASR_EUNSUPPORTED, "lstm_seq: H=%d needs %zu bytes of shared memory per CTA", a.H, smem);
  B2_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  B2_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem));
  const int NC = (a.H + U - 1) / U;
  const int capacity = per_sm * sms;
  int groups = capacity / (NC * a.ND);
  B2_REQUIRE(groups >= 1, B2ASR_EUNSUPPORTED, "lstm_seq: H=%d ND=%d needs %d co-resident CTAs, the device holds %d", a.H,
             a.ND, NC * a.ND, capacity);
  for (int row0 = 0; row0 < base.B; row0 += groups * kBS) {
    const int rows = base.B - row0 < groups * kBS ? base.B - row0 : groups * kBS;
    const int g = (rows + kBS - 1) / kBS;
    const int rc = backward ? launch_bwd<U, kBS>(a, row0, rows, g, smem, st) : launch_fwd<U, kBS>(a, row0, rows, g, smem, st);
    if (rc != B2ASR_OK) return rc;
  }
  return B2ASR_OK;
}

int check(const b2asr_lstm_seq* io, bool backward) {
  B2_REQUIRE(io, B2ASR_EINVAL, "lstm_seq: null descriptor");
  B2_REQUIRE(io->B > 0 && io->T > 0 && io->H > 0 && (io->ND == 1 || io->ND == 2), B2ASR_EINVAL,
             "lstm_seq: bad sizes B=%d T=%d H=%d ND=%d", io->B, io->T, io->H, io->ND);
  B2_REQUIRE(io->H % 4 == 0, B2ASR_EUNSUPPORTED, "lstm_seq: H=%d is not a multiple of 4", io->H);
  B2_REQUIRE(io->gates && io->lens && io->cells && io->status && io->w_hh[0] && (io->ND == 1 || io->w_hh[1]), B2ASR_EINVAL,
             "lstm_seq: null pointer argument");
  if (backward) B2_REQUIRE(io->d_out && io->d_gates, B2ASR_EINVAL, "lstm_seq_bwd: null gradient pointer");
  else B2_REQUIRE(io->out, B2ASR_EINVAL, "lstm_seq_fwd: null output pointer");
  return B2ASR_OK;
}

SeqArgs to_args(const b2asr_lstm_seq* io) {
  SeqArgs a;
  a.B = io->B; a.T = io->T; a.H = io->H; a.ND = io->ND; a.Kp = 0;
  a.gates = io->gates;
  for (int d = 0; d < 2; ++d) { a.w_hh[d] = io->w_hh[d]; a.b_ih[d] = io->b_ih[d]; a.b_hh[d] = io->b_hh[d]; }
  a.lens = io->lens; a.out = io->out; a.cells = io->cells; a.d_out = io->d_out; a.d_gates = io->d_gates;
  a.status = io->status;
  return a;
}

int pick_units(int H) {                       // units per CTA: ~64 CTAs per (direction, row group)
  if (H <= 4 * 64) return 4;
  if (H <= 5 * 64) return 5;
  return 8;
}

int dispatch(const b2asr_lstm_seq* io, bool backward, cudaStream_t st) {
  const SeqArgs a = to_args(io);
  switch (pick_units(io->H)) {
    case 4: return run<4>(a, backward, st);
    case 5: return run<5>(a, backward, st);
    default: return run<8>(a, backward, st);
  }
}

// H in {256, 320} takes the cluster kernels (enc_cluster.cu) unless the L2-exchange path is forced
// (b2asr_lstm_seq_path(1), or B2ASR_ENC_PATH=l2 in the environment at load time): A/B runs and tests
int g_path = [] { const char* e = getenv("B2ASR_ENC_PATH"); return (e && e[0] == 'l') ? 1 : 0; }();
bool use_cluster(int H) { return g_path == 0 && lstm_cluster_supported(H); }

}  // namespace
}  // namespace b2asr

using namespace b2asr;

extern "C" size_t b2asr_lstm_seq_size(void) { return sizeof(b2asr_lstm_seq); }

extern "C" int b2asr_lstm_seq_path(int path) {
  const int prev = g_path;
  if (path == 0 || path == 1) g_path = path;
  return prev;
}

extern "C" int b2asr_lstm_seq_fwd(const b2asr_lstm_seq* io, void* stream) {
  const int rc = check(io, false);
  if (rc != B2ASR_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (use_cluster(io->H)) {
    const int rc2 = lstm_cluster_run(to_args(io), false, st);
    if (rc2 != B2ASR_EUNSUPPORTED) return rc2;
  }
  B2_CUDA_TRY(cudaMemsetAsync(io->out, 0xff, (size_t)io->B * io->T * io->ND * io->H * sizeof(float), st));
  return dispatch(io, false, st);
}

extern "C" int b2asr_lstm_seq_bwd(const b2asr_lstm_seq* io, void* stream) {
  const int rc = check(io, true);
  if (rc != B2ASR_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (use_cluster(io->H)) {
    const int rc2 = lstm_cluster_run(to_args(io), true, st);
    if (rc2 != B2ASR_EUNSUPPORTED) return rc2;
  }
  B2_CUDA_TRY(cudaMemsetAsync(io->d_gates, 0xff, (size_t)io->B * io->T * io->ND * 4 * io->H * sizeof(float), st));
  return dispatch(io, true, st);
}
