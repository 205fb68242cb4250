// Kernel group (1): CTC loss + gradient for sm_100a.
//
// Replaces torch.nn.CTCLoss(blank=0,'mean') (+ the fused F.log_softmax) as called by the
// reference at src/solver.py:93,160,253-254.  Arithmetic: Graves 2006 forward-backward in
// log2-space over the blank-interleaved label lattice, ATen gradient convention
// (softmax - occupancy), see DESIGN.md section "CTC".
//
// Three launches per call (two when only the loss is wanted):
//   ctc_rows_kernel     wide, HBM-bound.  One pass over the [T,B,V] logits: row log-sum-exp,
//                       grad row = prescale*scale_b*softmax (rows past in_len zeroed) and the
//                       per-frame table of the log2-probs the lattice will need
//                       (whole vocabulary when V is small, gathered at the label string
//                       otherwise) -> 1 read + 1 write of the big tensors.
//   ctc_lattice_kernel  narrow, latency-bound.  One 2-CTA cluster per utterance: CTA 0 walks alpha forward in
//                       time over all frames, CTA 1 walks beta backward; every renormalised row goes to
//                       the workspace.  Label states are striped across the warps of a CTA, K consecutive
//                       states per thread; neighbours travel by warp shuffle and a one-float
//                       shared-memory hand-off per warp boundary.  Values are renormalised every chunk
//                       so fp32 keeps ~1e-6 relative accuracy at T=3000.
//   ctc_occupancy_kernel wide.  One warp per (utterance, frame): occupancies from the stored alpha/beta rows,
//                       normalised by the frame's own sum (= P(l|x) measured at that frame), subtracted from the
//                       gradient row as sparse fire-and-forget reductions; entries too small to change the fp32
//                       gradient are skipped, so only the band of the lattice around the alignments is touched.
//
// Schedules (B2ASR_CTC_OVERLAP; V > 1024 and gradient wanted, otherwise 0):
//   0  the three kernels one after the other in the caller's stream;
//   2  (default) rows pass, then the walks on a second, high-priority stream and the occupancy pass in the caller's
//      stream CONCURRENTLY: the occupancy pass visits the frames from the middle outwards -- the order in which both
//      walks have passed them -- and each of its CTAs waits on the walks' published progress (`prog`, device flags);
//   3  all three concurrently: additionally the rows pass visits the frames in the order the walks consume them
//      (16-frame blocks alternately from the front and the back) and counts finished rows per block (`rdy`), and a
//      producer warp per walk CTA waits for the blocks of the next chunk before issuing its bulk copy.  Measured
//      slower than 2 at every M5 point: next to the HBM-streaming rows pass the walks' one latency-critical warp per
//      scheduler loses its issue slots, and the rows pass loses bandwidth to the walks' table reads.
// For 2 and 3 the walk CTAs must be resident before the wide kernels flood the SMs (gate kernel + `started` counter),
// every kernel is loaded before the first overlapped call (lazy module loading would otherwise synchronise the
// context under a spinning kernel) and all kernels ask for the maximum shared-memory carve-out so that they co-reside.
#include "common.cuh"
#include <atomic>
#include <mutex>

namespace b2asr {

struct CtcArgs {
  const float* x; long long sx_t, sx_b;
  const long long* targets; long long tgt_stride;
  const long long* in_len; const long long* tgt_len;
  int T, B, V, Lmax, blank, from_logits, reduction;
  float prescale;
  float* nll; float* loss;
  float* grad; long long sg_t, sg_b;
  // workspace carve-up
  unsigned* ticket;
  float* tab; int W; int tab_gather;      // table row width (multiple of 4); 1 = gathered at labels
  float* scr; int Spad;                   // [B][2][T][Spad] stored lattice rows (global state coords)
  float* poff;                            // [B][2][T] offset folded away inside the row's chunk, up to and including the row
  double* cbase; int nchunk_max;          // [B][2][nchunk_max] offset folded away before each chunk
  double* logp2; int* feasible;           // [B] log2 P(l|x) and whether any alignment exists
  int* cls;                               // [B][Lmax] label classes clamped to [0,V)
  int* chain;                             // [B][Lmax] bits 0..29: 1 + next position with the same class (0 = none); bit 30: not the first of its class
  int RBlog, NST;                         // ring: log2(rows per stage), stages
  int pair;                               // 1 = 2-CTA cluster per utterance (meet in the middle)
  // overlapped schedule (see the header of this file); all zeroed by the call before the first launch
  int overlap, nblkT, pub;                // overlap: bit 0 = rows pass feeds the walks by flags, bit 1 = occupancy pass
                                          // follows the walks by flags; nblkT = ceil(T/16); pub = chunks between publications
  unsigned* rdy;                          // [B][nblkT] valid rows of each 16-frame block whose table row is written
  unsigned* prog;                         // [B][2] steps of the alpha / beta walk whose stored rows are published
  int* err;                               // [1] set to 1 when a wait gave up (results are then undefined)
  unsigned* started;                      // [1] walk CTAs that are resident (the rows pass is held back until then)
};

#ifdef B2ASR_CTC_TRACE
// one-off instrumentation: %globaltimer stamps in the workspace header (words 8.. of the first 256 bytes)
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
#define CTC_TRACE_MIN(slot) atomicMin(reinterpret_cast<unsigned long long*>(p.ticket) + 8 + (slot), gtime())
#define CTC_TRACE_MAX(slot) atomicMax(reinterpret_cast<unsigned long long*>(p.ticket) + 8 + (slot), gtime())
#else
#define CTC_TRACE_MIN(slot)
#define CTC_TRACE_MAX(slot)
#endif
constexpr int kBlkLog = 4;                // rows-pass / walk hand-off granularity: 16 frames
constexpr long long kPollLimit = 1ll << 26;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() {
  asm volatile("fence.proxy.async.global;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ float row_scale(const CtcArgs& p, int L) {
  return p.reduction == B2ASR_REDUCE_MEAN ? p.prescale / ((float)p.B * (float)max(L, 1)) : p.prescale;
}

// Overlapped schedule: one thread in the rows pass's stream, ahead of the rows pass.  Both kernels become eligible at
// the same moment (the walks' stream waits for the same event), and if the 10^5 small CTAs of the rows pass are
// dispatched first, the large walk CTAs find no room until they have drained.  This holds the rows pass back until
// `want` walk CTAs are resident (or gives up: the schedule then degrades to the serial one, it does not deadlock).
__global__ void ctc_gate_kernel(const unsigned* started, unsigned want) {
  for (int i = 0; i < (1 << 16) && ld_acquire_u32(started) < want; ++i) __nanosleep(100);
}

// --------------------------------------------------------------------------------------
// Wide pass.  TPR threads cooperate on one row; a 256-thread CTA handles 256/TPR rows.
// --------------------------------------------------------------------------------------
// One row of the wide pass (TPR threads; the whole CTA when TPR == 256).  `staged`: the row is already on its way
// into `cache` by a bulk copy that completes on `rbar`.
template <int TPR>
__device__ __forceinline__ void ctc_row(const CtcArgs& p, int Vpad, int b, int t, bool in_range, float* cache,
                                        float (*red)[8], unsigned long long* rbar, bool staged) {
  const int l = threadIdx.x % TPR;
  const int len = in_range ? (int)min((long long)p.T, max(0ll, p.in_len[b])) : 0;
  const int L = in_range ? (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b])) : 0;
  const bool valid = in_range && t < len;
  const float* xr = p.x + (long long)t * p.sx_t + (long long)b * p.sx_b;
  float* gr = p.grad ? p.grad + (long long)t * p.sg_t + (long long)b * p.sg_b : nullptr;
  const int V = p.V;
  const bool vec_in = (V % 4 == 0) && (p.sx_t % 4 == 0) && (p.sx_b % 4 == 0) && (((uintptr_t)p.x & 15) == 0);
  const bool vec_out = gr && (V % 4 == 0) && (p.sg_t % 4 == 0) && (p.sg_b % 4 == 0) && (((uintptr_t)p.grad & 15) == 0);

  if (TPR < 256 && !valid) {           // warp-per-row: padded / out-of-range rows finish here
    if (in_range && gr) {
      if (vec_out) for (int i = l; i < V / 4; i += TPR) reinterpret_cast<float4*>(gr)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      else for (int i = l; i < V; i += TPR) gr[i] = 0.f;
    }
    return;
  }
  if (TPR == 256 && !valid) {          // uniform for the whole CTA
    if (in_range && gr) {
      if (vec_out) for (int i = l; i < V / 4; i += TPR) reinterpret_cast<float4*>(gr)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      else for (int i = l; i < V; i += TPR) gr[i] = 0.f;
    }
    return;
  }

  // ---- pass 1: stream the row into shared memory, running max ------------------------------
  float m = -INFINITY;
  if (TPR == 256 && staged) {
    mbar_wait(rbar, 0u);
    for (int i = l; i < V / 4; i += TPR) {
      const float4 v = reinterpret_cast<const float4*>(cache)[i];
      m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    }
  } else if (vec_in) {
    for (int i = l; i < V / 4; i += TPR) {
      float4 v = __ldcs(reinterpret_cast<const float4*>(xr) + i);
      reinterpret_cast<float4*>(cache)[i] = v;
      m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    }
  } else {
    for (int i = l; i < V; i += TPR) { float v = __ldcs(xr + i); cache[i] = v; m = fmaxf(m, v); }
  }
  m = warp_max(m);
  if (TPR == 256) {
    if ((threadIdx.x & 31) == 0) red[0][threadIdx.x >> 5] = m;
    __syncthreads();
    m = red[0][0];
#pragma unroll
    for (int w = 1; w < 8; ++w) m = fmaxf(m, red[0][w]);
  } else {
    __syncwarp();
  }
  float lse = 0.f;
  if (p.from_logits) {
    const float mm = (m == -INFINITY) ? 0.f : m;
    float s = 0.f;
    for (int i = l; i < V; i += TPR) s += ex2((cache[i] - mm) * kLog2e);
    s = warp_sum(s);
    if (TPR == 256) {
      if ((threadIdx.x & 31) == 0) red[1][threadIdx.x >> 5] = s;
      __syncthreads();
      s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[1][w];
    }
    lse = mm + logf(s);
  }

  // ---- table of log2-probs for the lattice, relative to the row's largest entry tm ---------
  // gathered layout: [0,L) = labels in string order; dense layout: [0,V) = classes.  Both: slot W-2 = blank,
  // slot W-1 = tm (so the lattice fetches {blank, tm} with one 64-bit broadcast load).
  float* trow = p.tab + ((long long)b * p.T + t) * p.W;
  float tm;
  if (p.tab_gather) {
    const long long* lab = p.targets + (long long)b * p.tgt_stride;
    tm = kNeg;
    for (int i = l; i <= L; i += TPR) {
      long long c = (i == L) ? (long long)p.blank : lab[i];
      c = min((long long)V - 1, max(0ll, c));
      tm = fmaxf(tm, fmaxf((cache[c] - lse) * kLog2e, kNeg));
    }
    tm = warp_max(tm);
    if (TPR == 256) {
      __syncthreads();
      if ((threadIdx.x & 31) == 0) red[0][threadIdx.x >> 5] = tm;
      __syncthreads();
      tm = red[0][0];
#pragma unroll
      for (int w = 1; w < 8; ++w) tm = fmaxf(tm, red[0][w]);
    }
    if (!(tm > -1.0e29f)) tm = 0.f;
    for (int i = l; i <= L; i += TPR) {
      long long c = (i == L) ? (long long)p.blank : lab[i];
      c = min((long long)V - 1, max(0ll, c));
      trow[i == L ? p.W - 2 : i] = fmaxf(fmaxf((cache[c] - lse) * kLog2e, kNeg) - tm, kNeg);
    }
  } else {
    tm = fmaxf((m - lse) * kLog2e, kNeg);
    if (!(tm > -1.0e29f)) tm = 0.f;
    for (int i = l; i < V; i += TPR) trow[i] = fmaxf(fmaxf((cache[i] - lse) * kLog2e, kNeg) - tm, kNeg);
    if (l == 0) trow[p.W - 2] = fmaxf(fmaxf((cache[p.blank] - lse) * kLog2e, kNeg) - tm, kNeg);
  }
  if (l == 0) trow[p.W - 1] = tm;
  // (the table goes out BEFORE the gradient row: the fence below then waits for a few hundred bytes of table stores,
  // not for the 20 KB row, whose stores stay fire-and-forget)
  if (TPR == 256 && (p.overlap & 1)) {    // this frame's table row is complete: count it for the walks
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(p.rdy + (long long)b * p.nblkT + (t >> kBlkLog), 1u);
  }

  // ---- pass 2 (last: nothing waits for it): grad row = prescale * scale_b * softmax ----------------
  if (gr) {
    const float gs = row_scale(p, L);
    const float off = lse * kLog2e;
    if (vec_out) {
      for (int i = l; i < V / 4; i += TPR) {
        float4 v = reinterpret_cast<const float4*>(cache)[i];
        v.x = ex2(fmaf(v.x, kLog2e, -off)) * gs; v.y = ex2(fmaf(v.y, kLog2e, -off)) * gs;
        v.z = ex2(fmaf(v.z, kLog2e, -off)) * gs; v.w = ex2(fmaf(v.w, kLog2e, -off)) * gs;
        reinterpret_cast<float4*>(gr)[i] = v;
      }
    } else {
      for (int i = l; i < V; i += TPR) gr[i] = ex2(fmaf(cache[i], kLog2e, -off)) * gs;
    }
  }
  if (threadIdx.x == 0) { CTC_TRACE_MAX(1); }
}

// The wide pass.  TPR == 32: a warp per row, 8 rows per CTA.  TPR == 256: a CTA works on `rpc` rows one after the
// other, all of them requested up front -- each row is ONE bulk (TMA) copy into shared memory, so its bytes are in
// flight whatever the number of resident CTAs and no load occupies a register.
template <int TPR, int NR>
__global__ void __launch_bounds__(256, 8) ctc_rows_kernel(CtcArgs p, int Vpad) {
  constexpr int rpc = NR;
  extern __shared__ __align__(16) float sm[];
  __shared__ float red[2][8];
  __shared__ __align__(8) unsigned long long rbar[4];
  if (threadIdx.x == 0) { CTC_TRACE_MIN(0); }
  if (TPR < 256) {
    const int g = threadIdx.x / TPR;
    const long long row = (long long)blockIdx.x * (256 / TPR) + g;
    const bool in_range = row < (long long)p.T * p.B;
    ctc_row<TPR>(p, Vpad, in_range ? (int)(row / p.T) : 0, in_range ? (int)(row % p.T) : 0, in_range,
                 sm + (size_t)g * Vpad, red, nullptr, false);
    return;
  }
  const bool vec_in = (p.V % 4 == 0) && (p.sx_t % 4 == 0) && (p.sx_b % 4 == 0) && (((uintptr_t)p.x & 15) == 0);
  int bb[NR], tt[NR];
  bool inr[NR], stg[NR];
#pragma unroll
  for (int rr = 0; rr < rpc; ++rr) {
    int b, t;
    bool in_range;
    if (p.overlap & 1) {
      // consumption order: round r hands every utterance one 16-frame block, alternately its next block from the
      // front and from the back (the order the two walks need them); blocks past the utterance's length come last
      const int per = (1 << kBlkLog) / rpc;                          // CTAs per block
      const long long q = (long long)blockIdx.x / per;
      b = (int)(q % p.B);
      const int r = (int)(q / p.B);
      const int len_b = (int)min((long long)p.T, max(0ll, p.in_len[b]));
      const int nblk = (len_b + (1 << kBlkLog) - 1) >> kBlkLog;
      const int blk = r < nblk ? ((r & 1) ? nblk - 1 - (r >> 1) : (r >> 1)) : r;
      t = (blk << kBlkLog) + (int)(blockIdx.x % per) * rpc + rr;
      in_range = t < p.T;
    } else {
      const long long row = (long long)blockIdx.x * rpc + rr;
      in_range = row < (long long)p.T * p.B;
      b = in_range ? (int)(row / p.T) : 0;
      t = in_range ? (int)(row % p.T) : 0;
    }
    if (!in_range) { b = 0; t = 0; }
    bb[rr] = b; tt[rr] = t; inr[rr] = in_range;
    stg[rr] = vec_in && in_range && t < (int)min((long long)p.T, max(0ll, p.in_len[b]));
  }
  if (threadIdx.x == 0) {
#pragma unroll
    for (int rr = 0; rr < rpc; ++rr) mbar_init(rbar + rr, 1);
    mbar_fence_init();
#pragma unroll
    for (int rr = 0; rr < rpc; ++rr)
      if (stg[rr]) {
        mbar_expect_tx(rbar + rr, (unsigned)p.V * 4u);
        bulk_g2s(sm + (size_t)rr * Vpad, p.x + (long long)tt[rr] * p.sx_t + (long long)bb[rr] * p.sx_b, (unsigned)p.V * 4u, rbar + rr);
      }
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < rpc; ++rr) {
    if (rr > 0) __syncthreads();                                     // `red` is reused from row to row
    ctc_row<256>(p, Vpad, bb[rr], tt[rr], inr[rr], sm + (size_t)rr * Vpad, red, rbar + rr, stg[rr]);
  }
}

// --------------------------------------------------------------------------------------
// Lattice pass.
// --------------------------------------------------------------------------------------
#ifdef B2ASR_CTC_ACCURATE
#define LAT_EX2(x) exp2f(x)
#define LAT_LG2(x) log2f(x)
#else
#define LAT_EX2(x) ex2(x)
#define LAT_LG2(x) lg2(x)
#endif
__device__ __forceinline__ float lse2_log2(float a, float b) {
  const float m = fmaxf(a, b), d = fminf(a, b) - m;
  return m + LAT_LG2(1.f + LAT_EX2(d));
}
__device__ __forceinline__ float lse3_log2(float a, float b, float c) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  const float m = fmaxf(hi, c), o = fminf(hi, c);
  return m + LAT_LG2(1.f + LAT_EX2(lo - m) + LAT_EX2(o - m));
}

__host__ __device__ inline size_t lattice_smem_bytes(int W, int Lmax, int RB, int NST) {
  size_t f = (size_t)NST * RB * W;                 // table ring
  f += (size_t)((Lmax + 4) / 4 * 4);               // label string (int)
  f += 64 + 64 + 4;                                // hand-off slots [2][32], warp maxima [2][32], final pair
  return f * 4 + 8 * 8;                            // + ring mbarriers
}

template <int K>
constexpr int lattice_max_threads() { return K <= 4 ? 1024 : 512; }

__device__ __forceinline__ void bar_sync_n(int nthreads) {
  asm volatile("bar.sync 0, %0;" ::"r"(nthreads) : "memory");
}
// order-preserving float -> int key, so a warp maximum is one redux.sync
__device__ __forceinline__ int f2ord(float f) {
  const int i = __float_as_int(f);
  return i ^ ((i >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ord2f(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

// One direction of the lattice walk per CTA (pair mode: cluster rank 0 = alpha forward in time, rank 1 = beta
// backward) over ALL frames of the utterance.  A thread owns K/2 consecutive (blank, label) state pairs in the
// walk's LOCAL coordinates (local state j = global state j forward, S-1-j backward: the reversed lattice has the
// same shape).  Per step and thread: one 64-bit broadcast load {blank, row max}, K/2 label loads, one shuffle,
// K log-sum-exps, one vector store of the row, one CTA barrier.  Renormalisation never touches the registers:
// the CTA maximum measured at the end of a chunk is folded into the frame offset of the next chunk's first step.
// Overlapped schedule, one lane of the producer warp: before a chunk's bulk copy, wait until the rows pass has
// finished every frame of it, then make those generic-proxy writes visible to the async proxy.  The compute warps
// hand a stage back through `empty` when they are done with it.
__device__ __noinline__ void lattice_producer(const CtcArgs& p, int b, int dir, int T, int nchunks, float* ring,
                                              unsigned long long* full, unsigned long long* empty) {
  const int W = p.W, RB = 1 << p.RBlog, NST = p.NST;
  const float* tab_b = p.tab + (long long)b * p.T * W;
  const unsigned* rdy_b = p.rdy + (long long)b * p.nblkT;
  long long polls = 0;
  for (int c = 0; c < nchunks; ++c) {
    if (c >= NST) mbar_wait(empty + (c % NST), (unsigned)(((c / NST) - 1) & 1));
    const int k0 = c << p.RBlog;
    const int rows = min(RB, T - k0);
    const int t_lo = dir ? (T - k0 - rows) : k0;
    for (int blk = t_lo >> kBlkLog; blk <= ((t_lo + rows - 1) >> kBlkLog); ++blk) {
      const unsigned need = (unsigned)min(1 << kBlkLog, T - (blk << kBlkLog));
      while (ld_acquire_u32(rdy_b + blk) < need) {
        __nanosleep(64);
        if (++polls > kPollLimit) { *p.err = 1; break; }
      }
    }
    fence_proxy_async_global();
    const unsigned bytes = (unsigned)rows * (unsigned)W * 4u;
    unsigned long long* bar = full + (c % NST);
    mbar_expect_tx(bar, bytes);
    bulk_g2s(ring + (size_t)(c % NST) * RB * W, tab_b + (long long)t_lo * W, bytes, bar);
  }
}

template <int K, bool GRAD, bool OVL>
__global__ void __launch_bounds__(lattice_max_threads<K>(), 1) ctc_lattice_kernel(CtcArgs p) {
  extern __shared__ __align__(16) unsigned char smraw[];
  constexpr int H = K / 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = p.pair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int dir = p.pair ? (int)cluster_ctarank() : 0;
  const int T = (int)min((long long)p.T, max(0ll, p.in_len[b]));
  const int L = (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b]));
  const int S = 2 * L + 1;
  const int W = p.W, RB = 1 << p.RBlog, NST = p.NST;

  float* ring = reinterpret_cast<float*>(smraw);
  int* labs = reinterpret_cast<int*>(ring + (size_t)NST * RB * W);
  float* xch = reinterpret_cast<float*>(labs + (p.Lmax + 4) / 4 * 4);
  int* wred = reinterpret_cast<int*>(xch + 64);
  float* fin = reinterpret_cast<float*>(wred + 64);
  unsigned long long* full = reinterpret_cast<unsigned long long*>(fin + 4);
  unsigned long long* empty = full + 4;               // OVL: stage handed back to the producer warp

  // warps that hold no state of THIS utterance leave after the set-up; the rest synchronise among themselves
  const int nact = ((S + K - 1) / K + 31) >> 5;
  const int nsync = nact * 32;

  const long long* lab_g = p.targets + (long long)b * p.tgt_stride;
  for (int i = tid; i < L; i += blockDim.x) labs[i] = (int)min((long long)p.V - 1, max(0ll, lab_g[i]));
  if (tid < 64) xch[tid] = kNeg;                      // step 0 reads the virtual row before the first frame
  if (tid == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(full + s, 1); if (OVL) mbar_init(empty + s, 1); }
    mbar_fence_init();
  }
  __syncthreads();

  const int nchunks = (T + RB - 1) >> p.RBlog;
  const float* tab_b = p.tab + (long long)b * p.T * W;
  auto issue = [&](int c) {                                  // thread 0: one bulk copy per chunk (rows are contiguous)
    if (c < nchunks) {
      const int k0 = c << p.RBlog;
      const int rows = min(RB, T - k0);
      const int t_lo = dir ? (T - k0 - rows) : k0;
      const unsigned bytes = (unsigned)rows * (unsigned)W * 4u;
      unsigned long long* bar = full + (c % NST);
      mbar_expect_tx(bar, bytes);
      bulk_g2s(ring + (size_t)(c % NST) * RB * W, tab_b + (long long)t_lo * W, bytes, bar);
    }
  };
  if (!OVL && tid == 0) for (int c = 0; c < NST - 1; ++c) issue(c);

  // same-class chains of the label string, for the occupancy pass (encoding at CtcArgs::chain): the first position
  // of a class sums the occupancies of the whole chain, so each (frame, class) receives exactly ONE reduction
  // and the gradient is bit-reproducible.
  if (GRAD && dir == 0) {
    const int Lp = (p.Lmax + 4) / 4 * 4;               // 16-byte aligned rows: the occupancy pass bulk-copies them
    int* ch = p.chain + (long long)b * Lp;
    int* cl = p.cls + (long long)b * Lp;
    for (int i = tid; i < L; i += blockDim.x) {
      const int c = labs[i];
      int nx = -1, first = 1;
      for (int j = i + 1; j < L; ++j) if (labs[j] == c) { nx = j; break; }
      for (int j = i - 1; j >= 0; --j) if (labs[j] == c) { first = 0; break; }
      ch[i] = (nx + 1) | (first ? 0 : 0x40000000);
      cl[i] = c;
    }
  }
  if (OVL && warp == (int)(blockDim.x >> 5) - 1) {
    // producer warp (never a compute warp: the launch adds it on top of the widest lattice); out of line so that
    // its registers do not count against the walk's
    if (lane == 0) lattice_producer(p, b, dir, T, nchunks, ring, full, empty);
    return;
  }
  if (warp >= nact) return;

  // ---- static description of my K/2 (blank, label) pairs ----------------------------------------
  int lix[H];                       // table slot of the label state
  bool skip[H], vB[H], vL[H];
#pragma unroll
  for (int j = 0; j < H; ++j) {
    const int lj = tid * H + j;                       // local label index
    vB[j] = lj <= L;
    vL[j] = lj < L;
    const int li = dir ? (L - 1 - lj) : lj;
    const int cls = vL[j] ? labs[li] : p.blank;
    lix[j] = vL[j] ? (p.tab_gather ? li : cls) : (W - 2);     // padding states read the blank slot: finite
    skip[j] = false;
    if (vL[j] && lj >= 1) skip[j] = labs[dir ? li + 1 : li - 1] != cls;
  }

  // virtual row before the first frame: all mass on local state 0, so the uniform step yields the first row
  float aB[H], aL[H];
#pragma unroll
  for (int j = 0; j < H; ++j) { aB[j] = kNeg; aL[j] = kNeg; }
  if (tid == 0) aB[0] = 0.f;

  double C = 0.0;                                     // offsets folded away before the current chunk
  float shift = 0.f;                                  // CTA maximum of the previous chunk's last row
  float drift = 0.f;                                  // running estimate of the best path's per-frame slope
  float* srow = p.scr + ((long long)b * 2 + dir) * p.T * p.Spad + (long long)tid * K;
  unsigned xw = (unsigned)warp * 4u;                  // byte offset of my hand-off slot, parity toggles bit 7
  const unsigned xr = (unsigned)(warp > 0 ? warp - 1 : 0) * 4u;
  const char* xchb = reinterpret_cast<const char*>(xch);

  int k = 0;
  if (p.overlap && tid == 0) atomicAdd(p.started, 1u);
  if (tid == 0) { CTC_TRACE_MIN(2); }
  for (int c = 0; c < nchunks; ++c) {
    if (tid == 0 && c == nchunks / 2) { CTC_TRACE_MIN(3); CTC_TRACE_MAX(4); }
    if (!OVL && tid == 0) issue(c + NST - 1);
    mbar_wait(full + (c % NST), (unsigned)((c / NST) & 1));
    const int rows = min(RB, T - (c << p.RBlog));
    const float* trow = ring + (size_t)(c % NST) * RB * W + (dir ? (size_t)(rows - 1) * W : 0);
    const int tstep = dir ? -W : W;
    float part = 0.f;                                 // offsets added inside this chunk (fp32: <= RB terms)
    float off = drift + shift;                        // first step also absorbs the previous chunk's maximum
    for (int r = 0; r < rows; ++r, ++k, trow += tstep) {
      const float2 bt = *reinterpret_cast<const float2*>(trow + W - 2);      // {blank, row max}
      float el[H];
#pragma unroll
      for (int j = 0; j < H; ++j) el[j] = trow[lix[j]] - off;
      const float eb = bt.x - off;
      part += bt.y + off;
      off = drift;
      float l1 = __shfl_up_sync(kFull, aL[H - 1], 1);
      if (lane == 0) l1 = warp > 0 ? *reinterpret_cast<const float*>(xchb + ((xw & 128u) ^ 128u) + xr) : kNeg;
#pragma unroll
      for (int j = H - 1; j >= 0; --j) {
        const float pl = j > 0 ? aL[j - 1] : l1;       // label state of the pair to my left (old row)
        const float nL = lse3_log2(aL[j], aB[j], skip[j] ? pl : kNeg);
        const float nB = lse2_log2(aB[j], pl);
        aL[j] = fmaxf(nL + el[j], kNeg);
        aB[j] = fmaxf(nB + eb, kNeg);
      }
      if (lane == 31) *reinterpret_cast<float*>(const_cast<char*>(xchb) + xw) = aL[H - 1];
      xw ^= 128u;
      if (GRAD) {
        if (K == 2) *reinterpret_cast<float2*>(srow) = make_float2(aB[0], aL[0]);
        else {
#pragma unroll
          for (int j = 0; j < H; j += 2) *reinterpret_cast<float4*>(srow + 2 * j) = make_float4(aB[j], aL[j], aB[j + 1], aL[j + 1]);
        }
        srow += p.Spad;
      }
      if (r + 1 == rows) {                            // publish my warp's maximum before the chunk's last barrier
        float mx = kNeg;
#pragma unroll
        for (int j = 0; j < H; ++j) mx = fmaxf(mx, fmaxf(vB[j] ? aB[j] : kNeg, vL[j] ? aL[j] : kNeg));
        const int km = __reduce_max_sync(kFull, f2ord(mx));
        if (lane == 0) wred[(c & 1) * 32 + warp] = km;
      }
      bar_sync_n(nsync);
    }
    if (OVL && tid == 0) mbar_arrive(empty + (c % NST));       // every compute thread is past the chunk's last barrier
    if (GRAD && (p.overlap & 2) && (((c + 1) % p.pub) == 0 || c + 1 == nchunks)) {
      __threadfence();                                         // my stored rows, then ...
      bar_sync_n(nsync);
      if (tid == 0) st_release_u32(p.prog + (long long)b * 2 + dir, (unsigned)k);     // ... k steps are published
    }
    C += (double)part;
    const int km = __reduce_max_sync(kFull, lane < nact ? wred[(c & 1) * 32 + lane] : f2ord(kNeg));
    const float mx = ord2f(km);
    shift = 0.f;
    if (mx > -1.0e29f) { shift = mx; drift += mx / (float)rows; }
  }
  if (tid == 0) { CTC_TRACE_MIN(5); CTC_TRACE_MAX(6); }
  // the last chunk's maximum was measured but never subtracted: the stored rows do not include it either
  // ---- forward direction: log P = C + log2(2^a[S-1] + 2^a[S-2]) off the last row ---------------
  if (dir == 0) {
#pragma unroll
    for (int j = 0; j < H; ++j) {
      const int lj = tid * H + j;
      if (lj == L) fin[2] = aB[j];
      if (lj == L - 1) fin[1] = aL[j];
    }
    if (tid == 0 && L == 0) fin[1] = kNeg;
    bar_sync_n(nsync);
    if (tid == 0) {
      float r; double lp2 = 0.0; int ok = 0;
      if (T == 0) r = (L == 0) ? 0.f : INFINITY;
      else {
        const float v = lse2_log2(fin[1], fin[2]);
        ok = v > -1.0e29f;
        lp2 = C + (double)v;
        r = ok ? (float)(-lp2 * (double)kLn2) : INFINITY;
      }
      p.nll[b] = r;
      if (GRAD) { p.logp2[b] = lp2; p.feasible[b] = ok; }
    }
  }

  // ---- batch reduction by the last CTA to finish (deterministic order) ----------------------
  if (p.loss != nullptr && dir == 0 && tid == 0) {
    __threadfence();
    if (atomicAdd(p.ticket, 1u) == (unsigned)p.B - 1u) {
      __threadfence();
      float acc = 0.f;
      for (int i = 0; i < p.B; ++i) {
        const float v = ((volatile float*)p.nll)[i];
        const int Li = (int)min((long long)p.Lmax, max(0ll, p.tgt_len[i]));
        acc += p.reduction == B2ASR_REDUCE_MEAN ? v / (float)max(Li, 1) : v;
      }
      *p.loss = p.reduction == B2ASR_REDUCE_MEAN ? acc / (float)p.B : acc;
    }
  }
}

// --------------------------------------------------------------------------------------
// Occupancy pass: wide.  A group of warps per (utterance, frame), consecutive frames of one utterance per CTA (they
// share the label classes and chains in shared memory).  For every lattice state e_s = alpha_t(s) + beta_t(s) -
// lp_t(s) (log2; both walks include the frame's emission).  The stored rows are relative to per-row offsets that
// are the same for every state of the frame, so the occupancies are formed as 2^(e_s - max) / sum_s 2^(e_s - max):
// each frame is normalised by ITS OWN measurement of P(l|x) (in exact arithmetic sum_s alpha_t(s) beta_t(s) / y_t(s)
// is P at every t), which needs nothing from the end of the walks -- so the pass can follow the walks frame by
// frame -- and cancels the fp32 drift the two walks accumulate over thousands of steps.  A frame whose states are
// all at log zero has no alignment through it (then no frame of the utterance has one): NaN row, as ATen does.
// grad[t,b,class(s)] -= scale_b * occupancy: blank states are reduced in the warp (one reduction per row); label
// states are staged in shared memory and the first position of every class sums its chain, so each (frame, class)
// receives at most one fire-and-forget reduction (bit-reproducible gradient).  A class whose occupancy is below
// 2^-26 of its softmax term could not change the fp32 entry and is skipped: only the band of the lattice around
// the alignments touches the gradient row again.
// Overlapped schedule: CTAs are ordered from the middle frames outwards and wait for both walks (`prog`).
// --------------------------------------------------------------------------------------
// Persistent CTAs (256 threads) take units of kOccRun consecutive frames of one utterance, round-robin.  A frame is
// worked on by WPF warps (1..8, a power of two chosen from the label length: ~100 state pairs per warp), so a CTA has
// FPC = 8 / WPF frames in progress; each of these groups runs a two-stage ring: while it works on one frame, the next
// frame's alpha, beta and table rows arrive by bulk (TMA) copies.  The utterance's class / chain arrays are bulk-copied
// once per unit.
constexpr int kOccRun = 16;

__global__ void __launch_bounds__(256) ctc_occupancy_kernel(CtcArgs p, int WPF, int nunits_per_utt) {
  // cls[Lp] | chain[Lp] | 17 mbarriers ([group][stage], [16] = cls/chain) | red[8][4] | per group: 2 x { A[Sp] B[Sp] Tb[W] }
  extern __shared__ __align__(16) int osm_i[];
  const int FPC = 8 / WPF;
  const int Lp = (p.Lmax + 4) / 4 * 4;
  const int Sp = (2 * p.Lmax + 1 + 3) / 4 * 4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int fr = warp / WPF;                               // my group within the CTA
  const int wl = warp % WPF;                               // my warp within the group
  const int gt = wl * 32 + lane, gn = WPF * 32;            // my index / number of threads in the group
  int* cls = osm_i;
  int* chn = osm_i + Lp;
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(osm_i + 2 * Lp);
  float* red = reinterpret_cast<float*>(bars + 18);                                      // [8 warps][4]
  const int fstride = 2 * Sp + p.W;                                                      // floats per staged frame
  float* stage0 = red + 32 + (size_t)fr * 2 * fstride;
  if (tid == 0) {
    for (int i = 0; i < 17; ++i) mbar_init(bars + i, 1);
    mbar_fence_init();
  }
  __syncthreads();
  auto group_sync = [&]() {
    if (WPF == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(1 + fr), "r"(gn) : "memory");
  };
  unsigned use[2] = {0u, 0u};                              // completed uses of my group's two stages
  unsigned cuse = 0u;                                      // completed uses of the cls/chain barrier
  if (tid == 0) { CTC_TRACE_MIN(7); }

  const long long nunits = (long long)nunits_per_utt * p.B;
  for (long long u = blockIdx.x; u < nunits; u += gridDim.x) {
    const int b = (int)(u % p.B);
    int grp = (int)(u / p.B);
    const int T = (int)min((long long)p.T, max(0ll, p.in_len[b]));
    const int ngrp = (T + kOccRun - 1) / kOccRun;
    if (grp >= ngrp) continue;                             // uniform for the CTA
    if (p.overlap & 2) {                                       // rank -> run of frames, from the middle outwards
      const int mid = ngrp >> 1, j = grp >> 1;
      grp = (grp & 1) ? mid - 1 - j : mid + j;
    }
    const int t0 = grp * kOccRun;
    const int t1 = min(T, t0 + kOccRun);
    const int L = (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b]));
    const int S = 2 * L + 1;
    const float gs = row_scale(p, L);
    if (tid == 0) {
      if (p.overlap & 2) {                                     // alpha must have stored frame t1-1, beta frame t0
        const unsigned n0 = (unsigned)t1, n1 = (unsigned)(T - t0);
        const unsigned* pr = p.prog + (long long)b * 2;
        long long polls = 0;
        while (ld_acquire_u32(pr) < n0 || ld_acquire_u32(pr + 1) < n1) {
          __nanosleep(256);
          if (++polls > kPollLimit) { *p.err = 1; break; }
        }
        fence_proxy_async_global();
      }
      fence_proxy_async_smem();
      const unsigned cb = (unsigned)Lp * 4u;               // class / chain arrays of the utterance
      mbar_expect_tx(bars + 16, 2u * cb);
      bulk_g2s(cls, p.cls + (long long)b * Lp, cb, bars + 16);
      bulk_g2s(chn, p.chain + (long long)b * Lp, cb, bars + 16);
    }
    __syncthreads();                                       // the unit's frames are published (overlap)
    const unsigned sb = (unsigned)((S + 3) / 4 * 4) * 4u, tb_bytes = (unsigned)p.W * 4u;
    auto issue = [&](int t, int st) {                      // group leader: stage frame t
      if (t < t1) {
        float* A = stage0 + (size_t)st * fstride;
        if (p.overlap & 2) fence_proxy_async_global();
        fence_proxy_async_smem();
        mbar_expect_tx(bars + fr * 2 + st, 2u * sb + tb_bytes);
        bulk_g2s(A, p.scr + (((long long)b * 2 + 0) * p.T + t) * p.Spad, sb, bars + fr * 2 + st);
        bulk_g2s(A + Sp, p.scr + (((long long)b * 2 + 1) * p.T + (T - 1 - t)) * p.Spad, sb, bars + fr * 2 + st);
        bulk_g2s(A + 2 * Sp, p.tab + ((long long)b * p.T + t) * p.W, tb_bytes, bars + fr * 2 + st);
      }
    };
    if (gt == 0) { issue(t0 + fr, 0); issue(t0 + fr + FPC, 1); }
    mbar_wait(bars + 16, cuse & 1u);
    ++cuse;
    int st = 0;
    for (int t = t0 + fr; t < t1; t += FPC, st ^= 1) {
      float* A = stage0 + (size_t)st * fstride;
      const float* Bv = A + Sp;
      const float* Tb = A + 2 * Sp;
      float* grow = p.grad + (long long)t * p.sg_t + (long long)b * p.sg_b;
      mbar_wait(bars + fr * 2 + st, use[st] & 1u);
      ++use[st];
      const float tb = Tb[p.W - 2], tm = Tb[p.W - 1];
      // rows are stored in each walk's local coordinates: alpha state s at A[s], beta state s at Bv[S-1-s]
      float m = -INFINITY;
      for (int i = gt; i <= L; i += gn) {                 // pair i = blank state 2i and label state 2i+1
        float2 av = *reinterpret_cast<const float2*>(A + 2 * i);
        av.x = av.x + Bv[2 * (L - i)] - tb;
        m = fmaxf(m, av.x);
        if (i < L) {
          av.y = av.y + Bv[2 * (L - i) - 1] - Tb[p.tab_gather ? i : cls[i]];
          m = fmaxf(m, av.y);
        }
        *reinterpret_cast<float2*>(A + 2 * i) = av;
      }
      m = warp_max(m);
      if (WPF > 1) {
        if (lane == 0) red[warp * 4] = m;
        group_sync();
        for (int w = 0; w < WPF; ++w) m = fmaxf(m, red[(fr * WPF + w) * 4]);
      }
      if (!(m > -1.0e29f)) {                               // no alignment: NaN on the valid rows, as ATen does
        for (int v = gt; v < p.V; v += gn) grow[v] = __int_as_float(0x7fc00000);
      } else {
        float sum = 0.f, bsum = 0.f;
        for (int i = gt; i <= L; i += gn) {
          const float2 e = *reinterpret_cast<const float2*>(A + 2 * i);
          bsum += ex2(e.x - m);
          if (i < L) { const float uu = ex2(e.y - m); A[2 * i + 1] = uu; sum += uu; }
        }
        bsum = warp_sum(bsum);
        sum = warp_sum(sum);
        if (WPF > 1) {
          if (lane == 0) { red[warp * 4 + 1] = bsum; red[warp * 4 + 2] = sum; }
          group_sync();
          bsum = 0.f; sum = 0.f;
          for (int w = 0; w < WPF; ++w) { bsum += red[(fr * WPF + w) * 4 + 1]; sum += red[(fr * WPF + w) * 4 + 2]; }
        } else {
          __syncwarp();
        }
        const float inv = 1.f / (sum + bsum);
        for (int i = gt; i < L; i += gn) {                // the first position of every class sums its chain
          const int v = chn[i];
          if (v & 0x40000000) continue;
          float acc = A[2 * i + 1];
          for (int nx = (v & 0x3fffffff) - 1; nx >= 0; nx = (chn[nx] & 0x3fffffff) - 1) acc += A[2 * nx + 1];
          acc *= inv;
          if (acc < 1.4901161e-8f * ex2(Tb[p.tab_gather ? i : cls[i]] + tm)) continue;   // below half an ulp of the entry
          red_add_f32(grow + cls[i], -gs * acc);
        }
        if (gt == 0) red_add_f32(grow + p.blank, -gs * bsum * inv);
      }
      group_sync();                                        // everybody is done with this stage: refill it
      if (gt == 0) issue(t + 2 * FPC, st);
    }
    __syncthreads();                                       // cls / chain are reused by the next unit
  }
  if (tid == 0) { CTC_TRACE_MAX(8); }
}

// --------------------------------------------------------------------------------------
// Rescale helper for autograd backward.
// --------------------------------------------------------------------------------------
__global__ void rescale_if_needed_kernel(float* buf, long long n, const float* upstream, float assumed) {
  const float u = *upstream;
  if (u == assumed) return;
  const float r = u / assumed;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) buf[i] *= r;
}

// --------------------------------------------------------------------------------------
// Host side: plan + launch.
// --------------------------------------------------------------------------------------
struct CtcPlan {
  int W, tab_gather, Spad, K, nthr, RBlog, NST, pair, overlap, nblkT, pub;
  size_t smem;
  size_t off_ticket, off_prog, off_rdy, zero_bytes, off_tab, off_scr, off_lp, off_feas, off_chain, off_cls, total;
};

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s && *s) ? atoi(s) : dflt;
}

// overlap: the caller allows the concurrent schedule (it needs a second stream, created once per device)
static int make_plan(int T, int B, int V, int Lmax, bool want_grad, bool allow_overlap, CtcPlan& pl) {
  const int S = 2 * Lmax + 1;
  pl.tab_gather = (V > 2 * (Lmax + 1)) ? 1 : 0;
  pl.W = (int)align_up((size_t)(pl.tab_gather ? Lmax : V) + 2, 4);       // + blank slot + per-frame offset slot
  // schedule: 0 serial; 2 = the occupancy pass follows the walks by flags (default for large vocabularies);
  // 3 = additionally the walks follow the rows pass (measured slower: next to the walks' CTAs fewer CTAs of the rows
  // pass fit on an SM, and the rows pass loses more than the overlap gains; kept behind B2ASR_CTC_OVERLAP=3)
  pl.overlap = 0;
  if (allow_overlap && want_grad && V > 1024) {
    const int m = env_int("B2ASR_CTC_OVERLAP", 2);
    pl.overlap = (m == 3 || m == 1) ? 3 : (m == 2 ? 2 : 0);
  }
  // K (even) states per thread.  The walk is bound by per-warp issue latency, not by throughput, so the best
  // split is the finest one: K=2 while the lattice fits 1024 threads, then 4, 8, 12.  The overlapped schedule
  // adds a producer warp on top.
  const int extra = (pl.overlap & 1) ? 32 : 0;
  int bestK = -1;
  for (int K : {2, 4, 8, 12}) {
    if ((int)align_up((size_t)((S + K - 1) / K), 32) + extra <= (K <= 4 ? 1024 : 512)) { bestK = K; break; }
  }
  const int forcedK = env_int("B2ASR_CTC_K", 0);
  if ((forcedK == 2 || forcedK == 4 || forcedK == 8 || forcedK == 12) && forcedK > bestK && bestK > 0) bestK = forcedK;
  if (bestK < 0) { set_error("ctc: label length %d too long (max %d)", Lmax, (12 * 480 - 1) / 2); return B2ASR_EUNSUPPORTED; }
  pl.K = bestK;
  pl.nthr = (int)align_up((size_t)((S + bestK - 1) / bestK), 32);
  pl.Spad = pl.nthr * pl.K;                                        // one aligned vector store per thread and row
  pl.nthr += extra;
  pl.pair = want_grad ? 1 : 0;
  // table ring: NST stages of 2^RBlog rows, each stage one bulk copy; ~64 rows in flight hide the copy latency.
  // Overlapped: a smaller ring, so that CTAs of the rows pass fit next to a walk CTA on the same SM.
  pl.NST = 4;
  pl.RBlog = 4;
  const size_t cap = (pl.overlap ? 96 : 160) * 1024;
  while (pl.RBlog > 0 && lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST) > cap) --pl.RBlog;
  if (lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST) > cap) pl.NST = 2;
  pl.smem = lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST);
  if (pl.smem > 227 * 1024) { set_error("ctc: lattice needs %zu bytes of shared memory", pl.smem); return B2ASR_EUNSUPPORTED; }
  pl.nblkT = (T + (1 << kBlkLog) - 1) >> kBlkLog;
  pl.pub = 64 >> pl.RBlog;                                          // progress published every 64 frames
  if (pl.pub < 1) pl.pub = 1;
  size_t off = 0;
  pl.off_ticket = off; off += 256;                                  // [0] ticket, [1] error word
  pl.off_prog = off; off += align_up((size_t)B * 2 * sizeof(unsigned), 256);
  pl.off_rdy = off; if (pl.overlap & 1) off += align_up((size_t)B * pl.nblkT * sizeof(unsigned), 256);
  pl.zero_bytes = off;                                              // cleared by the call before the first launch
  pl.off_tab = off; off += align_up((size_t)B * T * pl.W * sizeof(float), 256);
  pl.off_scr = off; if (want_grad) off += align_up((size_t)B * 2 * T * pl.Spad * sizeof(float), 256);
  pl.off_lp = off; off += align_up((size_t)B * sizeof(double), 256);
  pl.off_feas = off; off += align_up((size_t)B * sizeof(int), 256);
  pl.off_chain = off; off += align_up((size_t)B * ((Lmax + 4) / 4 * 4) * sizeof(int), 256);
  pl.off_cls = off; off += align_up((size_t)B * ((Lmax + 4) / 4 * 4) * sizeof(int), 256);
  pl.total = off;
  return B2ASR_OK;
}

template <int K, bool GRAD, bool OVL>
static int launch_lattice_t(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<K, GRAD, OVL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.pair ? 2 * a.B : a.B);
  cfg.blockDim = dim3(pl.nthr);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = a.pair ? 2 : 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, ctc_lattice_kernel<K, GRAD, OVL>, a));
  return B2ASR_OK;
}

template <int K>
static int launch_lattice(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  if (a.overlap & 1) return launch_lattice_t<K, true, true>(a, pl, st);
  return a.grad != nullptr ? launch_lattice_t<K, true, false>(a, pl, st) : launch_lattice_t<K, false, false>(a, pl, st);
}

static int launch_lattice_k(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  switch (pl.K) {
    case 2: return launch_lattice<2>(a, pl, st);
    case 4: return launch_lattice<4>(a, pl, st);
    case 8: return launch_lattice<8>(a, pl, st);
    case 12: return launch_lattice<12>(a, pl, st);
  }
  set_error("ctc: internal: K=%d", pl.K);
  return B2ASR_EUNSUPPORTED;
}

// The overlapped schedule's second stream and its fork / join events: one set per device, created on first use
// (host-side driver objects; no device memory).
struct SideStream { cudaStream_t stream; cudaEvent_t fork, join; std::atomic<int> ready; };
static SideStream g_side[64];
static std::mutex g_side_mu;

static int side_stream(SideStream** out) {
  int dev = -1;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < 64, B2ASR_EUNSUPPORTED, "ctc: device ordinal %d outside [0,64)", dev);
  SideStream& s = g_side[dev];
  if (!s.ready.load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> lk(g_side_mu);
    if (!s.ready.load(std::memory_order_relaxed)) {
      int lo = 0, hi = 0;
      B2_CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      B2_CUDA_TRY(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, hi));
      B2_CUDA_TRY(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
      B2_CUDA_TRY(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
      // With lazy module loading the FIRST launch of a kernel may need a context synchronisation, i.e. it can wait
      // for every running kernel -- and the walks spin until the rows pass has run.  Load the kernels that are
      // launched behind the walks now, while nothing of this call is in flight.
      // Kernels of different launches share an SM only inside the shared-memory carve-out the SM was configured
      // with by the first of them: ask for the largest one, or a walk CTA (40-100 KB) leaves no room for the
      // 20 KB CTAs of the rows pass (measured: the rows pass then runs on the 20 walk-free SMs only, 6.6x slower).
      const int maxsh = (int)cudaSharedmemCarveoutMaxShared;
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_rows_kernel<256, 1>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_rows_kernel<256, 2>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_occupancy_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_gate_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<2, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<4, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<8, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<12, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<2, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<4, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<8, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<12, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, maxsh));
      cudaFuncAttributes fa;
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<2, true, false>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<4, true, false>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<8, true, false>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<12, true, false>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_rows_kernel<256, 1>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_rows_kernel<256, 2>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_occupancy_kernel));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_gate_kernel));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<2, true, true>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<4, true, true>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<8, true, true>));
      B2_CUDA_TRY(cudaFuncGetAttributes(&fa, ctc_lattice_kernel<12, true, true>));
      s.ready.store(1, std::memory_order_release);
    }
  }
  *out = &s;
  return B2ASR_OK;
}

static int ctc_call(const float* x, int64_t sx_t, int64_t sx_b, const int64_t* targets, int64_t tgt_stride,
                    const int64_t* in_len, const int64_t* tgt_len, int T, int B, int V, int Lmax, int blank,
                    int from_logits, int reduction, float prescale, float* nll, float* loss,
                    float* grad, int64_t sg_t, int64_t sg_b, void* ws, size_t ws_bytes, void* stream, bool allow_overlap) {
  B2_REQUIRE(x && targets && in_len && tgt_len && nll && ws, B2ASR_EINVAL, "ctc: null pointer argument");
  B2_REQUIRE(T > 0 && B > 0 && V > 1 && Lmax >= 0, B2ASR_EINVAL, "ctc: bad sizes T=%d B=%d V=%d Lmax=%d", T, B, V, Lmax);
  B2_REQUIRE(blank >= 0 && blank < V, B2ASR_EINVAL, "ctc: blank %d outside [0,%d)", blank, V);
  B2_REQUIRE(reduction >= 0 && reduction <= 2, B2ASR_EINVAL, "ctc: bad reduction %d", reduction);
  B2_REQUIRE(((uintptr_t)ws & 255) == 0, B2ASR_EINVAL, "ctc: workspace must be 256-byte aligned");
  const bool want_grad = grad != nullptr;
  CtcPlan pl;
  int rc = make_plan(T, B, V, Lmax, want_grad, allow_overlap, pl);
  if (rc != B2ASR_OK) return rc;
  if (ws_bytes < pl.total && pl.overlap) {             // a workspace sized for the serial schedule still works
    rc = make_plan(T, B, V, Lmax, want_grad, false, pl);
    if (rc != B2ASR_OK) return rc;
  }
  B2_REQUIRE(ws_bytes >= pl.total, B2ASR_EWORKSPACE, "ctc: workspace %zu < required %zu", ws_bytes, pl.total);
  cudaStream_t st = (cudaStream_t)stream;
  if (env_int("B2ASR_CTC_DEBUG", 0))
    fprintf(stderr, "ctc plan: T=%d B=%d V=%d Lmax=%d overlap=%d K=%d nthr=%d RB=%d NST=%d smem=%zu W=%d Spad=%d total=%zu ws=%zu\n",
            T, B, V, Lmax, pl.overlap, pl.K, pl.nthr, 1 << pl.RBlog, pl.NST, pl.smem, pl.W, pl.Spad, pl.total, ws_bytes);

  CtcArgs a;
  a.x = x; a.sx_t = sx_t; a.sx_b = sx_b;
  a.targets = (const long long*)targets; a.tgt_stride = tgt_stride;
  a.in_len = (const long long*)in_len; a.tgt_len = (const long long*)tgt_len;
  a.T = T; a.B = B; a.V = V; a.Lmax = Lmax; a.blank = blank; a.from_logits = from_logits; a.reduction = reduction;
  a.prescale = prescale;
  a.nll = nll; a.loss = loss; a.grad = grad; a.sg_t = sg_t; a.sg_b = sg_b;
  unsigned char* w = (unsigned char*)ws;
  a.ticket = (unsigned*)(w + pl.off_ticket);
  a.err = (int*)(w + pl.off_ticket) + 1;
  a.started = (unsigned*)(w + pl.off_ticket) + 2;
  a.prog = (unsigned*)(w + pl.off_prog); a.rdy = (unsigned*)(w + pl.off_rdy);
  a.overlap = pl.overlap; a.nblkT = pl.nblkT; a.pub = pl.pub;
  a.tab = (float*)(w + pl.off_tab); a.W = pl.W; a.tab_gather = pl.tab_gather;
  a.scr = (float*)(w + pl.off_scr); a.Spad = pl.Spad;
  a.logp2 = (double*)(w + pl.off_lp); a.feasible = (int*)(w + pl.off_feas); a.chain = (int*)(w + pl.off_chain); a.cls = (int*)(w + pl.off_cls);
  a.RBlog = pl.RBlog; a.NST = pl.NST; a.pair = pl.pair;

  B2_CUDA_TRY(cudaMemsetAsync(w, 0, pl.zero_bytes, st));          // ticket, error word, progress and ready counters
#ifdef B2ASR_CTC_TRACE
  B2_CUDA_TRY(cudaMemsetAsync(w + 64, 0xff, 8 * 8, st));          // atomicMin slots 0..7 start at ~0 (slot 8+: max)
  B2_CUDA_TRY(cudaMemsetAsync(w + 64 + 8, 0, 8, st));             // slot 1 max
  B2_CUDA_TRY(cudaMemsetAsync(w + 64 + 4 * 8, 0, 8, st));         // slot 4 max
  B2_CUDA_TRY(cudaMemsetAsync(w + 64 + 6 * 8, 0, 8, st));         // slot 6 max
#endif

  // ---- overlapped schedule: the walks go first, on the second stream, and wait for the rows pass by flags --------
  SideStream* side = nullptr;
  const unsigned walkers = (unsigned)(a.pair ? 2 * B : B);
  auto fork_walks = [&]() -> int {                     // walks on the second stream, then hold `st` until they are resident
    B2_CUDA_TRY(cudaEventRecord(side->fork, st));
    B2_CUDA_TRY(cudaStreamWaitEvent(side->stream, side->fork, 0));
    int r = launch_lattice_k(a, pl, side->stream);
    if (r != B2ASR_OK) return r;
    B2_CUDA_TRY(cudaEventRecord(side->join, side->stream));
    ctc_gate_kernel<<<1, 1, 0, st>>>(a.started, walkers < 128u ? walkers : 128u);
    B2_CUDA_TRY(cudaGetLastError());
    return B2ASR_OK;
  };
  if (pl.overlap) {
    rc = side_stream(&side);
    if (rc != B2ASR_OK) return rc;
  }
  if (pl.overlap & 1) {
    rc = fork_walks();
    if (rc != B2ASR_OK) return rc;
  }

  // ---- wide pass --------------------------------------------------------------------------
  const long long rows = (long long)T * B;
  if (V <= 1024) {
    const int Vpad = (int)align_up((size_t)V, 4);
    const size_t smem = (size_t)8 * Vpad * sizeof(float);
    const long long blocks = (rows + 7) / 8;
    ctc_rows_kernel<32, 1><<<(unsigned)blocks, 256, smem, st>>>(a, Vpad);
  } else {
    const int Vpad = (int)align_up((size_t)V, 4);
    int rpc = env_int("B2ASR_CTC_RPC", (pl.overlap & 1) ? 2 : 1);      // rows per CTA (1 or 2)
    if (rpc != 2 || (size_t)2 * Vpad * sizeof(float) > 100 * 1024) rpc = 1;
    const size_t smem = (size_t)rpc * Vpad * sizeof(float);
    B2_REQUIRE(smem <= 200 * 1024, B2ASR_EUNSUPPORTED, "ctc: vocabulary %d too large for the row cache", V);
    auto kern = rpc == 2 ? ctc_rows_kernel<256, 2> : ctc_rows_kernel<256, 1>;
    if (smem > 48 * 1024)
      B2_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long blocks = (pl.overlap & 1) ? (((long long)pl.nblkT * B) << kBlkLog) / rpc : (rows + rpc - 1) / rpc;
    kern<<<(unsigned)blocks, 256, smem, st>>>(a, Vpad);
  }
  B2_CUDA_TRY(cudaGetLastError());

  // ---- lattice pass -------------------------------------------------------------------------
  if (pl.overlap == 2) rc = fork_walks();              // after the rows pass, next to the occupancy pass
  else if (!pl.overlap) rc = launch_lattice_k(a, pl, st);
  if (rc != B2ASR_OK) return rc;

  // ---- occupancy pass (gradient only) ---------------------------------------------------------
  if (want_grad) {
    const int Lp = (Lmax + 4) / 4 * 4, Sp = (2 * Lmax + 1 + 3) / 4 * 4;
    // ~10 state pairs per thread: with fewer, the per-frame fixed cost (barriers, reductions, pointer set-up) made
    // the pass instruction-bound (measured at L=600: 8 warps per frame 807 us, 2 warps per frame see DESIGN.md)
    const int ppw = env_int("B2ASR_CTC_OCC_PPW", 320);
    int WPF = 1;
    while (WPF < 8 && (Lmax + 1) > ppw * WPF) WPF <<= 1;
    auto occ_smem = [&](int fpc) { return (size_t)2 * Lp * 4 + 144 + 128 + (size_t)fpc * 2 * (2 * Sp + pl.W) * 4; };
    while (WPF < 8 && occ_smem(8 / WPF) > (size_t)(pl.overlap ? 56 : 104) * 1024) WPF <<= 1;
    const int FPC = 8 / WPF;
    const size_t osm = occ_smem(FPC);
    B2_REQUIRE(osm <= 200 * 1024, B2ASR_EUNSUPPORTED, "ctc: label length %d too long for the occupancy pass", Lmax);
    if (osm > 48 * 1024)
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_occupancy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osm));
    const int nunits_per_utt = (T + kOccRun - 1) / kOccRun;
    int per_sm = 0;
    B2_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ctc_occupancy_kernel, 256, osm));
    int dev = 0, sms = 0;
    B2_CUDA_TRY(cudaGetDevice(&dev));
    B2_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    long long nblk = (long long)(per_sm > 0 ? per_sm : 1) * sms;
    if (nblk > (long long)nunits_per_utt * B) nblk = (long long)nunits_per_utt * B;
    ctc_occupancy_kernel<<<(unsigned)nblk, 256, osm, st>>>(a, WPF, nunits_per_utt);
    B2_CUDA_TRY(cudaGetLastError());
  }
  if (pl.overlap) B2_CUDA_TRY(cudaStreamWaitEvent(st, side->join, 0));     // nll / loss come from the walks
  return B2ASR_OK;
}

}  // namespace b2asr

using namespace b2asr;

extern "C" size_t b2asr_ctc_workspace_bytes(int T, int B, int V, int Lmax, int want_grad) {
  CtcPlan pl;
  if (T <= 0 || B <= 0 || V <= 0 || Lmax < 0) return 0;
  if (make_plan(T, B, V, Lmax, want_grad != 0, true, pl) != B2ASR_OK) return 0;
  size_t need = pl.total;
  if (make_plan(T, B, V, Lmax, want_grad != 0, false, pl) == B2ASR_OK && pl.total > need) need = pl.total;
  return need;
}

extern "C" int b2asr_ctc_loss_fwd_bwd(const float* x, int64_t sx_t, int64_t sx_b,
                                      const int64_t* targets, int64_t tgt_stride,
                                      const int64_t* in_len, const int64_t* tgt_len,
                                      int T, int B, int V, int Lmax, int blank,
                                      int from_logits, int reduction, float prescale,
                                      float* nll, float* loss,
                                      float* grad, int64_t sg_t, int64_t sg_b,
                                      void* ws, size_t ws_bytes, void* stream) {
  return ctc_call(x, sx_t, sx_b, targets, tgt_stride, in_len, tgt_len, T, B, V, Lmax, blank, from_logits, reduction,
                  prescale, nll, loss, grad, sg_t, sg_b, ws, ws_bytes, stream, true);
}

extern "C" int b2asr_rescale_if_needed(float* buf, int64_t n, const float* upstream, float assumed, void* stream) {
  B2_REQUIRE(buf && upstream && n >= 0, B2ASR_EINVAL, "rescale: bad arguments");
  B2_REQUIRE(assumed != 0.f, B2ASR_EINVAL, "rescale: assumed upstream must be non-zero");
  if (n == 0) return B2ASR_OK;
  const int blocks = (int)((n + 256 * 8 - 1) / (256 * 8) < 148 * 8 ? (n + 256 * 8 - 1) / (256 * 8) : 148 * 8);
  rescale_if_needed_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(buf, n, upstream, assumed);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

// --------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference).
// Callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the
// same three launches with the workspace in a file-scope device array instead of a caller buffer.  Every GPU that
// loads the library holds B2ASR_CTC_STATIC_WS_BYTES of it until exit.  Host code cannot use `g_ctc_static_ws` as a
// pointer (the name is a host-side stand-in for the symbol): the device address is looked up once per device with
// cudaGetSymbolAddress, with that device current, and cached.
// --------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_ctc_static_ws[B2ASR_CTC_STATIC_WS_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<void*> g_static_ws_addr[kMaxDevices];

int static_ws_for_current_device(void** out) {
  int dev = -1;
  B2_CUDA_TRY(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < kMaxDevices, B2ASR_EUNSUPPORTED, "ctc static: device ordinal %d outside [0,%d)", dev, kMaxDevices);
  void* p = g_static_ws_addr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    B2_CUDA_TRY(cudaGetSymbolAddress(&p, g_ctc_static_ws));     // the current device's copy of the array
    B2_REQUIRE(p != nullptr && ((uintptr_t)p & 255) == 0, B2ASR_ECUDA, "ctc static: bad symbol address %p", p);
    g_static_ws_addr[dev].store(p, std::memory_order_release);  // racing threads look up the same address
  }
  *out = p;
  return B2ASR_OK;
}
}  // namespace

extern "C" int b2asr_ctc_loss_fwd_bwd_static(const float* x, int64_t sx_t, int64_t sx_b,
                                             const int64_t* targets, int64_t tgt_stride,
                                             const int64_t* in_len, const int64_t* tgt_len,
                                             int T, int B, int V, int Lmax, int blank,
                                             int from_logits, int reduction, float prescale,
                                             float* nll, float* loss,
                                             float* grad, int64_t sg_t, int64_t sg_b, void* stream) {
  B2_REQUIRE(T > 0 && B > 0 && V > 1 && Lmax >= 0, B2ASR_EINVAL, "ctc static: bad sizes T=%d B=%d V=%d Lmax=%d", T, B, V, Lmax);
  CtcPlan pl;
  int rc = make_plan(T, B, V, Lmax, grad != nullptr, false, pl);     // serial schedule: no second stream to create
  if (rc != B2ASR_OK) return rc;
  B2_REQUIRE(pl.total <= (size_t)B2ASR_CTC_STATIC_WS_BYTES, B2ASR_EWORKSPACE,
             "ctc static: T=%d B=%d V=%d Lmax=%d needs %zu bytes of scratch, the built-in array holds %zu "
             "(use b2asr_ctc_loss_fwd_bwd with a caller workspace)", T, B, V, Lmax, pl.total, (size_t)B2ASR_CTC_STATIC_WS_BYTES);
  void* ws = nullptr;
  rc = static_ws_for_current_device(&ws);
  if (rc != B2ASR_OK) return rc;
  // The call clears the words that must start at zero (ticket, error word, progress counters) itself; every table,
  // row and offset the later passes read is written earlier in the same call, so the previous call's contents do
  // not matter.
  return ctc_call(x, sx_t, sx_b, targets, tgt_stride, in_len, tgt_len, T, B, V, Lmax, blank, from_logits,
                  reduction, prescale, nll, loss, grad, sg_t, sg_b, ws, (size_t)B2ASR_CTC_STATIC_WS_BYTES, stream, false);
}
