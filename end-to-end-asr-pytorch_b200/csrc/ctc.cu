// Kernel group (1): CTC loss + gradient for sm_100a.
//
// Replaces torch.nn.CTCLoss(blank=0,'mean') (+ the fused F.log_softmax) as called by the
// reference at src/solver.py:93,160,253-254.  Arithmetic: Graves 2006 forward-backward in
// log2-space over the blank-interleaved label lattice, ATen gradient convention
// (softmax - occupancy), see DESIGN.md section "CTC".
//
// Two launches per call:
//   ctc_rows_kernel     wide, HBM-bound.  One pass over the [T,B,V] logits: row log-sum-exp,
//                       grad row = prescale*scale_b*softmax (rows past in_len zeroed) and the
//                       per-frame table of the log2-probs the lattice will need
//                       (whole vocabulary when V is small, gathered at the label string
//                       otherwise) -> 1 read + 1 write of the big tensors.
//   ctc_lattice_kernel  narrow, latency-bound.  One 2-CTA cluster per utterance: CTA 0 walks
//                       alpha forward in time, CTA 1 walks beta backward; they meet in the
//                       middle (one cluster barrier), after which each side finishes the
//                       other's half, emitting occupancies as sparse corrections onto grad.
//                       Label states are striped across the warps of a CTA, K consecutive
//                       states per thread; neighbours travel by warp shuffle and a one-float
//                       shared-memory hand-off per warp boundary.  Values are renormalised
//                       every chunk so fp32 keeps ~1e-6 relative accuracy at T=3000.
#include "common.cuh"

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
  double* cs;                             // [B][2][T] renormalisation offsets per stored row
  int RB, NST;                            // ring: rows per stage, stages
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

  // ---- table of log2-probs for the lattice; last slot = a per-frame offset (max entry) ------
  float* trow = p.tab + ((long long)b * p.T + t) * p.W;
  float tm;
  if (p.tab_gather) {
    const long long* lab = p.targets + (long long)b * p.tgt_stride;
    tm = kNeg;
    for (int i = l; i <= L; i += TPR) {
      long long c = (i == 0) ? (long long)p.blank : lab[i - 1];
      c = min((long long)V - 1, max(0ll, c));
      const float val = fmaxf((cache[c] - lse) * kLog2e, kNeg);
      trow[i] = val;
      tm = fmaxf(tm, val);
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
  } else {
    for (int i = l; i < V; i += TPR) trow[i] = fmaxf((cache[i] - lse) * kLog2e, kNeg);
    tm = fmaxf((m - lse) * kLog2e, kNeg);
  }
  if (l == 0) trow[p.W - 1] = (tm > -1.0e29f) ? tm : 0.f;
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

struct LatticeSmem {
  float* tabring; float* othring; float* obuf; int* nxt; int* labs; int* prv; float* xch; float* wred; float* bsum;
};

__host__ __device__ inline size_t lattice_smem_bytes(int W, int Spad, int Lmax, int RB, int NST, bool want_grad) {
  size_t f = (size_t)NST * RB * W;                 // tabring
  if (want_grad) f += (size_t)NST * RB * Spad;     // othring
  f += 2 * (size_t)Spad;                            // obuf
  f += (size_t)Spad;                                // nxt (int)
  f += 2 * (size_t)((Lmax + 4) / 4 * 4);            // labs, prv (int)
  f += 64 + 32 + 64;                                // xch, wred, bsum
  return f * 4;
}

template <int K>
constexpr int lattice_max_threads() { return K <= 4 ? 1024 : 512; }

template <int K>
__global__ void __launch_bounds__(lattice_max_threads<K>(), 1) ctc_lattice_kernel(CtcArgs p) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
  const bool want_grad = p.grad != nullptr;
  const int b = p.pair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int T = (int)min((long long)p.T, max(0ll, p.in_len[b]));
  const int L = (int)min((long long)p.Lmax, max(0ll, p.tgt_len[b]));
  const int S = 2 * L + 1;
  const int W = p.W, Spad = p.Spad, RB = p.RB, NST = p.NST;

  LatticeSmem sm;
  {
    float* f = reinterpret_cast<float*>(smraw);
    sm.tabring = f; f += (size_t)NST * RB * W;
    sm.othring = f; if (want_grad) f += (size_t)NST * RB * Spad;
    sm.obuf = f; f += 2 * (size_t)Spad;
    sm.nxt = reinterpret_cast<int*>(f); f += Spad;
    sm.labs = reinterpret_cast<int*>(f); f += (p.Lmax + 4) / 4 * 4;
    sm.prv = reinterpret_cast<int*>(f); f += (p.Lmax + 4) / 4 * 4;
    sm.xch = f; f += 64;
    sm.wred = f; f += 32;
    sm.bsum = f;
  }

  // ---- labels + same-label chains (so each (t,class) gets exactly one correction) ----------
  const long long* lab_g = p.targets + (long long)b * p.tgt_stride;
  for (int i = tid; i < L; i += nthr) sm.labs[i] = (int)min((long long)p.V - 1, max(0ll, lab_g[i]));
  for (int i = tid; i < Spad; i += nthr) sm.nxt[i] = -1;
  __syncthreads();
  // prv[i] = nearest earlier label position with the same class (-1: i heads its chain);
  // nxt[2i+1] = next lattice state with the same class (-1: end of chain).
  for (int i = tid; i < L; i += nthr) {
    const int c = sm.labs[i];
    int j = i - 1;
    while (j >= 0 && sm.labs[j] != c) --j;
    sm.prv[i] = j;
  }
  __syncthreads();
  for (int i = tid; i < L; i += nthr) {
    const int pj = sm.prv[i];
    if (pj >= 0) sm.nxt[2 * pj + 1] = 2 * i + 1;       // one writer per slot: a label has at most one successor
  }
  __syncthreads();
  unsigned headmask = 0;                                // bit kk: my state kk is the head of its class chain

  const float gs = row_scale(p, L);
  const int mid = T / 2;
  const int npass = (p.pair || !want_grad) ? 1 : 2;
  double logP = 0.0;
  bool feasible = true;

  for (int pass = 0; pass < npass; ++pass) {
    const int dir = p.pair ? (int)cluster_ctarank() : pass;
    int n1;                                            // steps [0,n1) are stored, [n1,T) are emitted
    if (!want_grad) n1 = T;
    else if (p.pair) n1 = dir == 0 ? mid : T - mid;
    else n1 = pass == 0 ? T : 0;

    // ---- per-thread static description of my K local states ---------------------------------
    int sglob[K], tix[K], cls[K];
    bool valid[K], skip[K];
    headmask = 0;
#pragma unroll
    for (int kk = 0; kk < K; ++kk) {
      const int jl = tid * K + kk;
      valid[kk] = jl < S;
      const int s = dir ? (S - 1 - jl) : jl;
      sglob[kk] = s;
      const bool is_lab = valid[kk] && (s & 1);
      const int li = (s - 1) >> 1;
      cls[kk] = is_lab ? sm.labs[li] : p.blank;
      tix[kk] = p.tab_gather ? (is_lab ? li + 1 : 0) : cls[kk];
      skip[kk] = false;
      if (is_lab && jl >= 2) {
        const int li2 = dir ? li + 1 : li - 1;
        skip[kk] = sm.labs[li2] != cls[kk];
      }
      if (is_lab && sm.prv[li] < 0) headmask |= 1u << kk;
    }

    float a[K];
#pragma unroll
    for (int kk = 0; kk < K; ++kk) a[kk] = kNeg;
    double C = 0.0;
    int pend_k = -1;                                    // emission waiting for its barrier
    float xshift = 0.f;                                 // renormalisation applied after xch was written
    float drift = 0.f;                                  // running estimate of the best path's per-frame slope

    for (int phase = 0; phase < 2; ++phase) {
      const int ks = phase == 0 ? 0 : n1;
      const int ke = phase == 0 ? n1 : T;
      if (phase == 1) {
        if (!want_grad) break;
        __threadfence();
        if (p.pair) cluster_sync_all(); else __syncthreads();
      }
      const int nsteps = ke - ks;
      if (nsteps <= 0) continue;
      const int nchunks = (nsteps + RB - 1) / RB;
      const int wq = W >> 2, sq = Spad >> 2;

      auto issue = [&](int c) {
        if (c < nchunks) {
          const int st = c % NST;
          const int k0 = ks + c * RB;
          const int rows = min(RB, ke - k0);
          float* tdst = sm.tabring + (size_t)st * RB * W;
          for (int i = tid; i < rows * wq; i += nthr) {
            const int r = i / wq, q = i - r * wq;
            const int k = k0 + r;
            const int t = dir ? (T - 1 - k) : k;
            cp_async16(tdst + (size_t)r * W + 4 * q, p.tab + ((long long)b * p.T + t) * W + 4 * q);
          }
          if (phase == 1) {
            float* odst = sm.othring + (size_t)st * RB * Spad;
            for (int i = tid; i < rows * sq; i += nthr) {
              const int r = i / sq, q = i - r * sq;
              const int k = k0 + r;
              const int ko = T - 1 - k;
              cp_async16(odst + (size_t)r * Spad + 4 * q,
                         p.scr + (((long long)b * 2 + (1 - dir)) * p.T + ko) * Spad + 4 * q);
            }
          }
        }
        cp_async_commit();
      };

      for (int c = 0; c < NST - 1; ++c) issue(c);

      for (int c = 0; c < nchunks; ++c) {
        // wait for stage c (all but the NST-2 most recent groups), make it CTA-visible
        if (NST == 2) cp_async_wait<0>(); else if (NST == 3) cp_async_wait<1>(); else cp_async_wait<2>();
        __syncthreads();
        issue(c + NST - 1);

        const int st = c % NST;
        const int k0 = ks + c * RB;
        const int rows = min(RB, ke - k0);
        for (int r = 0; r < rows; ++r) {
          const int k = k0 + r;
          const int t = dir ? (T - 1 - k) : k;
          const float* trow = sm.tabring + ((size_t)st * RB + r) * W;

          // ---- finish the previous step's emission (its obuf/bsum are visible now) ----------
          if (pend_k >= 0) {
            const int tp = dir ? (T - 1 - pend_k) : pend_k;
            const float* ob = sm.obuf + (size_t)(pend_k & 1) * Spad;
            float* grow = p.grad + (long long)tp * p.sg_t + (long long)b * p.sg_b;
#pragma unroll
            for (int kk = 1; kk < K; kk += 2) {
              if (headmask & (1u << kk)) {
                float sum = ob[sglob[kk]];
                int nx = sm.nxt[sglob[kk]];
                while (nx >= 0) { sum += ob[nx]; nx = sm.nxt[nx]; }
                atomicAdd(grow + cls[kk], -gs * sum);
              }
            }
            if (tid == 0) {
              float bs = 0.f;
              for (int w = 0; w < nwarp; ++w) bs += sm.bsum[(pend_k & 1) * 32 + w];
              atomicAdd(grow + p.blank, -gs * bs);
            }
            pend_k = -1;
          }

          // ---- lattice step ---------------------------------------------------------------
          float lp[K];
#pragma unroll
          for (int kk = 0; kk < K; ++kk) lp[kk] = valid[kk] ? trow[tix[kk]] : 0.f;
          // per-frame offset (table max + learned slope): keeps the best path near 0, where fp32 is dense
          const float rmax = trow[W - 1] + drift;
          C += (double)rmax;
          if (k == 0) {
#pragma unroll
            for (int kk = 0; kk < K; ++kk) {
              const int jl = tid * K + kk;
              a[kk] = (jl < 2 && valid[kk]) ? fmaxf(lp[kk] - rmax, kNeg) : kNeg;
            }
          } else {
            float l1 = __shfl_up_sync(kFull, a[K - 1], 1);
            if (lane == 0) l1 = warp > 0 ? fmaxf(sm.xch[((k - 1) & 1) * 32 + warp - 1] - xshift, kNeg) : kNeg;
            xshift = 0.f;
#pragma unroll
            for (int kk = K - 1; kk >= 0; --kk) {
              const float p0 = a[kk];
              const float p1 = kk > 0 ? a[kk - 1] : l1;
              float v;
              if (kk & 1) {
                const float p2 = kk > 1 ? a[kk - 2] : l1;   // kk==1: two back is the left neighbour
                v = lse3_log2(p0, p1, skip[kk] ? p2 : kNeg);
              } else {
                v = lse2_log2(p0, p1);
              }
              a[kk] = fmaxf(v + (lp[kk] - rmax), kNeg);
            }
          }
          if (lane == 31) sm.xch[(k & 1) * 32 + warp] = a[K - 1];

          if (phase == 0) {
            if (want_grad) {
              float* srow = p.scr + (((long long)b * 2 + dir) * p.T + k) * Spad;
#pragma unroll
              for (int kk = 0; kk < K; ++kk) if (valid[kk]) srow[sglob[kk]] = a[kk];
              if (tid == 0) p.cs[((long long)b * 2 + dir) * p.T + k] = C;
            }
          } else {
            const float* orow = sm.othring + ((size_t)st * RB + r) * Spad;
            const double Co = p.cs[((long long)b * 2 + (1 - dir)) * p.T + (T - 1 - k)];
            float e[K];
#pragma unroll
            for (int kk = 0; kk < K; ++kk) e[kk] = valid[kk] ? (a[kk] + orow[sglob[kk]] - lp[kk]) : kNeg;
            if (k == n1) {
              // meeting point: log2 P = C + Co + log2 sum_s 2^(e_s)   (block reduction, once)
              float mx = kNeg;
#pragma unroll
              for (int kk = 0; kk < K; ++kk) mx = fmaxf(mx, e[kk]);
              mx = warp_max(mx);
              if (lane == 0) sm.wred[warp] = mx;
              __syncthreads();
              mx = sm.wred[0];
              for (int w = 1; w < nwarp; ++w) mx = fmaxf(mx, sm.wred[w]);
              __syncthreads();
              float sx = 0.f;
#pragma unroll
              for (int kk = 0; kk < K; ++kk) sx += valid[kk] ? ex2(e[kk] - mx) : 0.f;
              sx = warp_sum(sx);
              if (lane == 0) sm.wred[warp] = sx;
              __syncthreads();
              sx = 0.f;
              for (int w = 0; w < nwarp; ++w) sx += sm.wred[w];
              __syncthreads();
              feasible = mx > -1.0e29f;
              logP = C + Co + (double)mx + (double)log2f(sx);
              if (tid == 0 && (dir == 0 || !p.pair)) {
                p.nll[b] = feasible ? (float)(-logP * (double)kLn2) : INFINITY;
              }
            }
            if (feasible) {
              const float cst = (float)(C + Co - logP);
              float* ob = sm.obuf + (size_t)(k & 1) * Spad;
              float bsum = 0.f;
#pragma unroll
              for (int kk = 0; kk < K; ++kk) {
                const float o = valid[kk] ? ex2(fminf(e[kk] + cst, 0.f)) : 0.f;
                if (kk & 1) { if (valid[kk]) ob[sglob[kk]] = o; }
                else bsum += o;
              }
              bsum = warp_sum(bsum);
              if (lane == 0) sm.bsum[(k & 1) * 32 + warp] = bsum;
              pend_k = k;
            } else {
              float* grow = p.grad + (long long)t * p.sg_t + (long long)b * p.sg_b;
              for (int v = tid; v < p.V; v += nthr) grow[v] = __int_as_float(0x7fc00000);
            }
          }

          // renormalise at the end of every chunk: block max -> subtract, remember in C
          const bool last_in_chunk = (r == rows - 1);
          if (last_in_chunk) {
            float mx = kNeg;
#pragma unroll
            for (int kk = 0; kk < K; ++kk) mx = fmaxf(mx, valid[kk] ? a[kk] : kNeg);
            mx = warp_max(mx);
            if (lane == 0) sm.wred[warp] = mx;
          }
          __syncthreads();
          if (last_in_chunk) {
            float mx = sm.wred[0];
            for (int w = 1; w < nwarp; ++w) mx = fmaxf(mx, sm.wred[w]);
            if (mx > -1.0e29f) {
#pragma unroll
              for (int kk = 0; kk < K; ++kk) a[kk] = fmaxf(a[kk] - mx, kNeg);
              C += (double)mx;
              xshift = mx;                              // the hand-off slot still holds the unshifted value
              drift += mx / (float)rows;                // what the offsets failed to absorb over this chunk
            }
          }
        }
      }
      cp_async_wait<0>();
    }  // phase

    // ---- flush the last pending emission of this pass ------------------------------------------
    if (pend_k >= 0) {
      const int tp = dir ? (T - 1 - pend_k) : pend_k;
      const float* ob = sm.obuf + (size_t)(pend_k & 1) * Spad;
      float* grow = p.grad + (long long)tp * p.sg_t + (long long)b * p.sg_b;
#pragma unroll
      for (int kk = 1; kk < K; kk += 2) {
        if (headmask & (1u << kk)) {
          float sum = ob[sglob[kk]];
          int nx = sm.nxt[sglob[kk]];
          while (nx >= 0) { sum += ob[nx]; nx = sm.nxt[nx]; }
          atomicAdd(grow + cls[kk], -gs * sum);
        }
      }
      if (tid == 0) {
        float bs = 0.f;
        for (int w = 0; w < nwarp; ++w) bs += sm.bsum[(pend_k & 1) * 32 + w];
        atomicAdd(grow + p.blank, -gs * bs);
      }
      pend_k = -1;
    }

    // ---- loss-only: read log P off the last forward row ------------------------------------------
    if (!want_grad) {
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < K; ++kk) {
        const int jl = tid * K + kk;
        if (valid[kk] && jl >= S - 2) sm.obuf[jl - (S - 2) + 1] = a[kk];   // slots 1 (S-2) and 2 (S-1)
      }
      if (tid == 0 && S == 1) sm.obuf[1] = kNeg;
      __syncthreads();
      if (tid == 0) {
        float r;
        if (T == 0) r = (L == 0) ? 0.f : INFINITY;
        else {
          const float v = lse2_log2(sm.obuf[1], sm.obuf[2]);
          r = (v > -1.0e29f) ? (float)(-(C + (double)v) * (double)kLn2) : INFINITY;
        }
        p.nll[b] = r;
      }
    }
    __syncthreads();
  }  // pass

  if (want_grad && T == 0 && tid == 0 && (!p.pair || cluster_ctarank() == 0)) p.nll[b] = (L == 0) ? 0.f : INFINITY;

  // ---- batch reduction by the last CTA to finish (deterministic order) ----------------------
  if (p.loss != nullptr && (!p.pair || cluster_ctarank() == 0)) {
    __shared__ unsigned is_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) is_last = (atomicAdd(p.ticket, 1u) == (unsigned)p.B - 1u);
    __syncthreads();
    if (is_last && tid == 0) {
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
  int W, tab_gather, Spad, K, nthr, RB, NST, pair;
  size_t smem;
  size_t off_ticket, off_tab, off_scr, off_cs, total;
};

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s && *s) ? atoi(s) : dflt;
}

static int make_plan(int T, int B, int V, int Lmax, bool want_grad, CtcPlan& pl) {
  const int S = 2 * Lmax + 1;
  pl.tab_gather = (V > 2 * (Lmax + 1)) ? 1 : 0;
  pl.W = (int)align_up((size_t)(pl.tab_gather ? Lmax + 1 : V) + 1, 4);   // +1: per-frame offset slot
  pl.Spad = (int)align_up((size_t)S, 4);
  // K (even) states per thread: balance MUFU work over the 4 SM sub-partitions
  int bestK = -1; long bestCost = 0;
  for (int K = 2; K <= 12; K += 2) {
    const int thr = (S + K - 1) / K;
    const int nw = (thr + 31) / 32;
    if (nw * 32 > (K <= 4 ? 1024 : 512)) continue;
    const long cost = (long)((nw + 3) / 4) * K * 20 + (nw > 1 ? 60 : 0) + nw * 4;
    if (bestK < 0 || cost < bestCost) { bestK = K; bestCost = cost; }
  }
  const int forcedK = env_int("B2ASR_CTC_K", 0);
  if (forcedK >= 2 && forcedK <= 12 && forcedK % 2 == 0 && (S + forcedK - 1) / forcedK <= (forcedK <= 4 ? 1024 : 512)) bestK = forcedK;
  if (bestK < 0) { set_error("ctc: label length %d too long (max %d)", Lmax, (12 * 1024 - 1) / 2); return B2ASR_EUNSUPPORTED; }
  pl.K = bestK;
  pl.nthr = (int)align_up((size_t)((S + bestK - 1) / bestK), 32);
  pl.pair = want_grad ? env_int("B2ASR_CTC_PAIR", 1) : 0;
  pl.NST = 3;
  pl.RB = 8;
  const size_t cap = 200 * 1024;
  while (pl.RB > 1 && lattice_smem_bytes(pl.W, pl.Spad, Lmax, pl.RB, pl.NST, want_grad) > cap) pl.RB >>= 1;
  pl.smem = lattice_smem_bytes(pl.W, pl.Spad, Lmax, pl.RB, pl.NST, want_grad);
  if (pl.smem > 227 * 1024) { set_error("ctc: lattice needs %zu bytes of shared memory", pl.smem); return B2ASR_EUNSUPPORTED; }
  size_t off = 0;
  pl.off_ticket = off; off += 256;
  pl.off_tab = off; off += align_up((size_t)B * T * pl.W * sizeof(float), 256);
  pl.off_scr = off; if (want_grad) off += align_up((size_t)B * 2 * T * pl.Spad * sizeof(float), 256);
  pl.off_cs = off; if (want_grad) off += align_up((size_t)B * 2 * T * sizeof(double), 256);
  pl.total = off;
  return B2ASR_OK;
}

template <int K>
static int launch_lattice(const CtcArgs& a, const CtcPlan& pl, cudaStream_t st) {
  B2_CUDA_TRY(cudaFuncSetAttribute(ctc_lattice_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
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
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, ctc_lattice_kernel<K>, a));
  return B2ASR_OK;
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
  a.cs = (double*)(w + pl.off_cs);
  a.RB = pl.RB; a.NST = pl.NST; a.pair = pl.pair;

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
    case 6: rc = launch_lattice<6>(a, pl, st); break;
    case 8: rc = launch_lattice<8>(a, pl, st); break;
    case 10: rc = launch_lattice<10>(a, pl, st); break;
    case 12: rc = launch_lattice<12>(a, pl, st); break;
    default: set_error("ctc: internal: K=%d", pl.K); return B2ASR_EUNSUPPORTED;
  }
  return rc;
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
