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
//   ctc_occupancy_kernel wide.  One CTA per (utterance, frame): occupancies from the stored alpha/beta rows,
//                       subtracted from the gradient row as sparse fire-and-forget reductions.  (An
//                       earlier version emitted them from inside the lattice walk after a meet-in-the-
//                       middle; ncu showed the walk paying ~340 warp-instructions per step for it.)
#include "common.cuh"
#include <atomic>

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
};

__device__ __forceinline__ float row_scale(const CtcArgs& p, int L) {
  return p.reduction == B2ASR_REDUCE_MEAN ? p.prescale / ((float)p.B * (float)max(L, 1)) : p.prescale;
}

// --------------------------------------------------------------------------------------
// Wide pass.  TPR threads cooperate on one row; a 256-thread CTA handles 256/TPR rows.
// --------------------------------------------------------------------------------------
template <int TPR>
__global__ void __launch_bounds__(256) ctc_rows_kernel(CtcArgs p, int Vpad) {
  extern __shared__ __align__(16) float sm[];
  constexpr int RPC = 256 / TPR;
  __shared__ float red[2][8];
  const int g = threadIdx.x / TPR, l = threadIdx.x % TPR;
  if (blockIdx.x == 0 && threadIdx.x == 0) *p.ticket = 0u;

  const long long row = (long long)blockIdx.x * RPC + g;
  const bool in_range = row < (long long)p.T * p.B;      // uniform per group (per CTA when TPR==256)
  const int b = in_range ? (int)(row / p.T) : 0;
  const int t = in_range ? (int)(row % p.T) : 0;
  const int len = in_range ? (int)min((long long)p.T, max(0ll, p.in_len[b])) : 0;
  const int L = in_range ? (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b])) : 0;
  const bool valid = in_range && t < len;
  float* cache = sm + (size_t)g * Vpad;
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
  if (vec_in) {
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

  // ---- pass 2: grad row = prescale * scale_b * softmax -------------------------------------
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
template <int K, bool GRAD>
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

  // warps that hold no state of THIS utterance leave after the set-up; the rest synchronise among themselves
  const int nact = ((S + K - 1) / K + 31) >> 5;
  const int nsync = nact * 32;

  const long long* lab_g = p.targets + (long long)b * p.tgt_stride;
  for (int i = tid; i < L; i += blockDim.x) labs[i] = (int)min((long long)p.V - 1, max(0ll, lab_g[i]));
  if (tid < 64) xch[tid] = kNeg;                      // step 0 reads the virtual row before the first frame
  if (tid == 0) {
    for (int s = 0; s < NST; ++s) mbar_init(full + s, 1);
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
  if (tid == 0) for (int c = 0; c < NST - 1; ++c) issue(c);

  // same-class chains of the label string, for the occupancy pass (encoding at CtcArgs::chain): the first position
  // of a class sums the occupancies of the whole chain, so each (frame, class) receives exactly ONE reduction
  // and the gradient is bit-reproducible.
  if (GRAD && dir == 0) {
    int* ch = p.chain + (long long)b * p.Lmax;
    int* cl = p.cls + (long long)b * p.Lmax;
    for (int i = tid; i < L; i += blockDim.x) {
      const int c = labs[i];
      int nx = -1, first = 1;
      for (int j = i + 1; j < L; ++j) if (labs[j] == c) { nx = j; break; }
      for (int j = i - 1; j >= 0; --j) if (labs[j] == c) { first = 0; break; }
      ch[i] = (nx + 1) | (first ? 0 : 0x40000000);
      cl[i] = c;
    }
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
  float* prow = p.poff + ((long long)b * 2 + dir) * p.T;
  double* cb = p.cbase + ((long long)b * 2 + dir) * p.nchunk_max;
  unsigned xw = (unsigned)warp * 4u;                  // byte offset of my hand-off slot, parity toggles bit 7
  const unsigned xr = (unsigned)(warp > 0 ? warp - 1 : 0) * 4u;
  const char* xchb = reinterpret_cast<const char*>(xch);

  int k = 0;
  for (int c = 0; c < nchunks; ++c) {
    if (tid == 0) { issue(c + NST - 1); if (GRAD) cb[c] = C; }
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
        if (tid == 0) prow[k] = part;
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
    C += (double)part;
    const int km = __reduce_max_sync(kFull, lane < nact ? wred[(c & 1) * 32 + lane] : f2ord(kNeg));
    const float mx = ord2f(km);
    shift = 0.f;
    if (mx > -1.0e29f) { shift = mx; drift += mx / (float)rows; }
  }
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
// Occupancy pass: wide.  One WARP per (utterance, frame), FPC consecutive frames of one utterance per CTA (they
// share the label classes and chains in shared memory): occupancy of every lattice state from the stored
// alpha / beta rows, o_s = 2^(alpha_t(s) + beta_t(s) - lp_t(s) + C_a + C_b - log2 P), then
// grad[t,b,class(s)] -= scale_b * o_s.  Blank states are reduced in the warp (one reduction per row); label
// states are staged in shared memory and the first position of every class sums its chain, so each
// (frame, class) receives exactly one fire-and-forget reduction (bit-reproducible gradient).
// --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ctc_occupancy_kernel(CtcArgs p, int FPC) {
  extern __shared__ __align__(16) int osm_i[];             // cls[Lp] | chain[Lp] | occ[FPC][Lp]
  const int b = blockIdx.y;
  const int T = (int)min((long long)p.T, max(0ll, p.in_len[b]));
  const int t0 = blockIdx.x * FPC;
  if (t0 >= T) return;
  const int L = (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b]));
  const int S = 2 * L + 1;
  const int Lp = (p.Lmax + 4) / 4 * 4;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int* cls = osm_i;
  int* chn = osm_i + Lp;
  float* occ = reinterpret_cast<float*>(osm_i + 2 * Lp) + (size_t)warp * Lp;
  const bool feasible = p.feasible[b] != 0;
  if (feasible) {
    const int* cg = p.cls + (long long)b * p.Lmax;
    const int* hg = p.chain + (long long)b * p.Lmax;
    for (int i = tid; i < L; i += blockDim.x) { cls[i] = cg[i]; chn[i] = hg[i]; }
  }
  __syncthreads();
  const int t = t0 + warp;
  if (t >= T) return;
  float* grow = p.grad + (long long)t * p.sg_t + (long long)b * p.sg_b;
  if (!feasible) {                                       // no alignment: NaN on the valid rows, as ATen does
    for (int v = lane; v < p.V; v += 32) grow[v] = __int_as_float(0x7fc00000);
    return;
  }
  const float gs = row_scale(p, L);
  const int kb = T - 1 - t;                                // the backward walk's step index of this frame
  const float* arow = p.scr + (((long long)b * 2 + 0) * p.T + t) * p.Spad;
  const float* brow = p.scr + (((long long)b * 2 + 1) * p.T + kb) * p.Spad;
  const float* trow = p.tab + ((long long)b * p.T + t) * p.W;
  const double ca = p.cbase[((long long)b * 2 + 0) * p.nchunk_max + (t >> p.RBlog)] + (double)p.poff[((long long)b * 2 + 0) * p.T + t];
  const double cbk = p.cbase[((long long)b * 2 + 1) * p.nchunk_max + (kb >> p.RBlog)] + (double)p.poff[((long long)b * 2 + 1) * p.T + kb];
  const float cst = (float)(ca + cbk - p.logp2[b] - (double)trow[p.W - 1]);   // table entries are relative to the row max
  const float cb = cst - trow[p.W - 2];
  // rows are stored in each walk's local coordinates: alpha state s at arow[s], beta state s at brow[S-1-s]
  float bsum = 0.f;
  for (int i = lane; i <= L; i += 32) {                   // pair i = blank state 2i and label state 2i+1
    const float2 av = *reinterpret_cast<const float2*>(arow + 2 * i);
    const float bb = brow[2 * (L - i)];
    bsum += ex2(fminf(av.x + bb + cb, 0.f));
    if (i < L) {
      const float bl = brow[2 * (L - i) - 1];
      const float lp = trow[p.tab_gather ? i : cls[i]];
      occ[i] = ex2(fminf(av.y + bl - lp + cst, 0.f));
    }
  }
  bsum = warp_sum(bsum);
  __syncwarp();
  for (int i = lane; i < L; i += 32) {                    // the first position of every class sums its chain
    const int v = chn[i];
    if (v & 0x40000000) continue;
    float sum = occ[i];
    for (int nx = (v & 0x3fffffff) - 1; nx >= 0; nx = (chn[nx] & 0x3fffffff) - 1) sum += occ[nx];
    red_add_f32(grow + cls[i], -gs * sum);
  }
  if (lane == 0) red_add_f32(grow + p.blank, -gs * bsum);
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
  int W, tab_gather, Spad, K, nthr, RBlog, NST, pair, nchunk_max;
  size_t smem;
  size_t off_ticket, off_tab, off_scr, off_poff, off_cbase, off_lp, off_feas, off_chain, off_cls, total;
};

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s && *s) ? atoi(s) : dflt;
}

static int make_plan(int T, int B, int V, int Lmax, bool want_grad, CtcPlan& pl) {
  const int S = 2 * Lmax + 1;
  pl.tab_gather = (V > 2 * (Lmax + 1)) ? 1 : 0;
  pl.W = (int)align_up((size_t)(pl.tab_gather ? Lmax : V) + 2, 4);       // + blank slot + per-frame offset slot
  pl.Spad = (int)align_up((size_t)S, 4);
  // K (even) states per thread.  The walk is bound by per-warp issue latency, not by throughput, so the best
  // split is the finest one: K=2 while the lattice fits 1024 threads, then 4, 8, 12.
  int bestK = -1;
  for (int K : {2, 4, 8, 12}) {
    if ((S + K - 1) / K <= (K <= 4 ? 1024 : 512)) { bestK = K; break; }
  }
  const int forcedK = env_int("B2ASR_CTC_K", 0);
  if ((forcedK == 2 || forcedK == 4 || forcedK == 8 || forcedK == 12) && forcedK > bestK && bestK > 0) bestK = forcedK;
  if (bestK < 0) { set_error("ctc: label length %d too long (max %d)", Lmax, (12 * 512 - 1) / 2); return B2ASR_EUNSUPPORTED; }
  pl.K = bestK;
  pl.nthr = (int)align_up((size_t)((S + bestK - 1) / bestK), 32);
  pl.Spad = pl.nthr * pl.K;                                        // one aligned vector store per thread and row
  pl.pair = want_grad ? 1 : 0;
  // table ring: NST stages of 2^RBlog rows, each stage one bulk copy; ~64 rows in flight hide the copy latency
  pl.NST = 4;
  pl.RBlog = 4;
  const size_t cap = 160 * 1024;
  while (pl.RBlog > 0 && lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST) > cap) --pl.RBlog;
  if (lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST) > cap) pl.NST = 2;
  pl.smem = lattice_smem_bytes(pl.W, Lmax, 1 << pl.RBlog, pl.NST);
  if (pl.smem > 227 * 1024) { set_error("ctc: lattice needs %zu bytes of shared memory", pl.smem); return B2ASR_EUNSUPPORTED; }
  pl.nchunk_max = (T + (1 << pl.RBlog) - 1) >> pl.RBlog;
  size_t off = 0;
  pl.off_ticket = off; off += 256;
  pl.off_tab = off; off += align_up((size_t)B * T * pl.W * sizeof(float), 256);
  pl.off_scr = off; if (want_grad) off += align_up((size_t)B * 2 * T * pl.Spad * sizeof(float), 256);
  pl.off_poff = off; if (want_grad) off += align_up((size_t)B * 2 * T * sizeof(float), 256);
  pl.off_cbase = off; if (want_grad) off += align_up((size_t)B * 2 * pl.nchunk_max * sizeof(double), 256);
  pl.off_lp = off; off += align_up((size_t)B * sizeof(double), 256);
  pl.off_feas = off; off += align_up((size_t)B * sizeof(int), 256);
  pl.off_chain = off; off += align_up((size_t)B * (Lmax + 1) * sizeof(int), 256);
  pl.off_cls = off; off += align_up((size_t)B * (Lmax + 1) * sizeof(int), 256);
  pl.total = off;
  return B2ASR_OK;
}

template <int K, bool GRAD>
static int launch_lattice_t(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<K, GRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
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
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, ctc_lattice_kernel<K, GRAD>, a));
  return B2ASR_OK;
}

template <int K>
static int launch_lattice(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  return a.grad != nullptr ? launch_lattice_t<K, true>(a, pl, st) : launch_lattice_t<K, false>(a, pl, st);
}

}  // namespace b2asr

using namespace b2asr;

extern "C" size_t b2asr_ctc_workspace_bytes(int T, int B, int V, int Lmax, int want_grad) {
  CtcPlan pl;
  if (T <= 0 || B <= 0 || V <= 0 || Lmax < 0) return 0;
  if (make_plan(T, B, V, Lmax, want_grad != 0, pl) != B2ASR_OK) return 0;
  return pl.total;
}

extern "C" int b2asr_ctc_loss_fwd_bwd(const float* x, int64_t sx_t, int64_t sx_b,
                                      const int64_t* targets, int64_t tgt_stride,
                                      const int64_t* in_len, const int64_t* tgt_len,
                                      int T, int B, int V, int Lmax, int blank,
                                      int from_logits, int reduction, float prescale,
                                      float* nll, float* loss,
                                      float* grad, int64_t sg_t, int64_t sg_b,
                                      void* ws, size_t ws_bytes, void* stream) {
  B2_REQUIRE(x && targets && in_len && tgt_len && nll && ws, B2ASR_EINVAL, "ctc: null pointer argument");
  B2_REQUIRE(T > 0 && B > 0 && V > 1 && Lmax >= 0, B2ASR_EINVAL, "ctc: bad sizes T=%d B=%d V=%d Lmax=%d", T, B, V, Lmax);
  B2_REQUIRE(blank >= 0 && blank < V, B2ASR_EINVAL, "ctc: blank %d outside [0,%d)", blank, V);
  B2_REQUIRE(reduction >= 0 && reduction <= 2, B2ASR_EINVAL, "ctc: bad reduction %d", reduction);
  B2_REQUIRE(((uintptr_t)ws & 255) == 0, B2ASR_EINVAL, "ctc: workspace must be 256-byte aligned");
  const bool want_grad = grad != nullptr;
  CtcPlan pl;
  int rc = make_plan(T, B, V, Lmax, want_grad, pl);
  if (rc != B2ASR_OK) return rc;
  B2_REQUIRE(ws_bytes >= pl.total, B2ASR_EWORKSPACE, "ctc: workspace %zu < required %zu", ws_bytes, pl.total);
  cudaStream_t st = (cudaStream_t)stream;

  CtcArgs a;
  a.x = x; a.sx_t = sx_t; a.sx_b = sx_b;
  a.targets = (const long long*)targets; a.tgt_stride = tgt_stride;
  a.in_len = (const long long*)in_len; a.tgt_len = (const long long*)tgt_len;
  a.T = T; a.B = B; a.V = V; a.Lmax = Lmax; a.blank = blank; a.from_logits = from_logits; a.reduction = reduction;
  a.prescale = prescale;
  a.nll = nll; a.loss = loss; a.grad = grad; a.sg_t = sg_t; a.sg_b = sg_b;
  unsigned char* w = (unsigned char*)ws;
  a.ticket = (unsigned*)(w + pl.off_ticket);
  a.tab = (float*)(w + pl.off_tab); a.W = pl.W; a.tab_gather = pl.tab_gather;
  a.scr = (float*)(w + pl.off_scr); a.Spad = pl.Spad;
  a.poff = (float*)(w + pl.off_poff); a.cbase = (double*)(w + pl.off_cbase); a.nchunk_max = pl.nchunk_max;
  a.logp2 = (double*)(w + pl.off_lp); a.feasible = (int*)(w + pl.off_feas); a.chain = (int*)(w + pl.off_chain); a.cls = (int*)(w + pl.off_cls);
  a.RBlog = pl.RBlog; a.NST = pl.NST; a.pair = pl.pair;

  // ---- wide pass --------------------------------------------------------------------------
  const long long rows = (long long)T * B;
  if (V <= 1024) {
    const int Vpad = (int)align_up((size_t)V, 4);
    const size_t smem = (size_t)8 * Vpad * sizeof(float);
    const long long blocks = (rows + 7) / 8;
    ctc_rows_kernel<32><<<(unsigned)blocks, 256, smem, st>>>(a, Vpad);
  } else {
    const int Vpad = (int)align_up((size_t)V, 4);
    const size_t smem = (size_t)Vpad * sizeof(float);
    B2_REQUIRE(smem <= 200 * 1024, B2ASR_EUNSUPPORTED, "ctc: vocabulary %d too large for the row cache", V);
    if (smem > 48 * 1024)
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_rows_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ctc_rows_kernel<256><<<(unsigned)rows, 256, smem, st>>>(a, Vpad);
  }
  B2_CUDA_TRY(cudaGetLastError());

  // ---- lattice pass -------------------------------------------------------------------------
  switch (pl.K) {
    case 2: rc = launch_lattice<2>(a, pl, st); break;
    case 4: rc = launch_lattice<4>(a, pl, st); break;
    case 8: rc = launch_lattice<8>(a, pl, st); break;
    case 12: rc = launch_lattice<12>(a, pl, st); break;
    default: set_error("ctc: internal: K=%d", pl.K); return B2ASR_EUNSUPPORTED;
  }
  if (rc != B2ASR_OK) return rc;

  // ---- occupancy pass (gradient only) ---------------------------------------------------------
  if (want_grad) {
    const int Lp = (Lmax + 4) / 4 * 4;
    int FPC = 8;
    while (FPC > 1 && (size_t)(FPC + 2) * Lp * 4 > 96 * 1024) FPC >>= 1;
    const size_t osm = (size_t)(FPC + 2) * Lp * 4;
    B2_REQUIRE(osm <= 200 * 1024, B2ASR_EUNSUPPORTED, "ctc: label length %d too long for the occupancy pass", Lmax);
    if (osm > 48 * 1024)
      B2_CUDA_TRY(cudaFuncSetAttribute(ctc_occupancy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osm));
    ctc_occupancy_kernel<<<dim3((unsigned)((T + FPC - 1) / FPC), (unsigned)B), 32 * FPC, osm, st>>>(a, FPC);
    B2_CUDA_TRY(cudaGetLastError());
  }
  return B2ASR_OK;
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
  int rc = make_plan(T, B, V, Lmax, grad != nullptr, pl);
  if (rc != B2ASR_OK) return rc;
  B2_REQUIRE(pl.total <= (size_t)B2ASR_CTC_STATIC_WS_BYTES, B2ASR_EWORKSPACE,
             "ctc static: T=%d B=%d V=%d Lmax=%d needs %zu bytes of scratch, the built-in array holds %zu "
             "(use b2asr_ctc_loss_fwd_bwd with a caller workspace)", T, B, V, Lmax, pl.total, (size_t)B2ASR_CTC_STATIC_WS_BYTES);
  void* ws = nullptr;
  rc = static_ws_for_current_device(&ws);
  if (rc != B2ASR_OK) return rc;
  // Nothing in the workspace has to start at zero: the rows pass re-arms the ticket word and every table, row and
  // offset the later passes read is written earlier in the same call, so the previous call's contents do not matter.
  return b2asr_ctc_loss_fwd_bwd(x, sx_t, sx_b, targets, tgt_stride, in_len, tgt_len, T, B, V, Lmax, blank, from_logits,
                                reduction, prescale, nll, loss, grad, sg_t, sg_b, ws, (size_t)B2ASR_CTC_STATIC_WS_BYTES, stream);
}
