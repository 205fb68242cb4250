// Kernel group (3): CTC prefix scorer + joint CTC/attention/RNN-LM beam step for sm_100a.
//
// Replaces, for all live hypotheses of one utterance and in ONE launch per decode step,
//   reference src/asr.py:210-256  (candidate top-k, score fusion, <sos> hack, LM fusion, top-beam,
//                                  child ranking and pruning),
//   reference src/ctc.py:65-101   (CTCPrefixScore.cheap_compute: float32 logaddexp recursion over the
//                                  encoder frames, -1e8 "log zero", the repeated-label rule :87-91),
//   reference src/postprocess.py:71-106 (Hypothesis.avgScore / addTopk bookkeeping).
// The recursion over T' frames is inherently sequential (each frame's logaddexp depends on the
// previous frame's result and re-association would change float32 roundings, hence decoded tokens);
// the kernel is latency-bound by design.  Parallelism: one CTA per hypothesis, the three dependent
// chains of every candidate (non-blank state, blank state, cumulative prefix score) on three
// adjacent lanes exchanging one value per frame by warp shuffle, ten candidates per warp.  Frame
// data are gathered once into shared memory by the whole CTA; both top-k selections are CTA radix
// selects followed by a warp-level rank sort; the last CTA to finish ranks the children and
// gathers the survivors' state, so the step needs no second launch and no host round trip.
#include "common.cuh"

namespace b2asr {

constexpr float kLogZero = -100000000.0f;     // ctc.py:11
constexpr float kNegCtc  = -1000000.0f;       // asr.py:225
constexpr float kNegSos  = -10000000.0f;      // asr.py:229
constexpr int   kEos     = 1;
constexpr int   kBeamNT  = 256;
constexpr int   kMaxK    = 32;

// numpy's float32 logaddexp (npy_logaddexpf): x==y -> x+ln2, else max + log1pf(expf(-|x-y|)).
// Default: the same two libm calls as numpy makes (CUDA's expf / log1pf differ from glibc's by an ulp in ~1.5 % of
// calls).  -DB2ASR_BEAM_SFU evaluates them on the special function unit instead -- e = 2^(-|x-y| log2 e) (MUFU.EX2),
// log1p(e) = ln2 * log2(u) + (e - (u-1))/u with u = 1 + e (MUFU.LG2; the second term restores what rounding 1 + e lost)
// -- absolute error per call <= ~2e-7 instead of ~6e-8, the recursion 26.7 instead of 39.3 us per step at M4 and all
// decode tests still token-identical; not the default because every rounding difference is a chance to flip a
// near-tie against the reference.
__device__ __forceinline__ float np_logaddexp(float x, float y) {
  const float tmp = __fsub_rn(x, y);
  const float m = tmp > 0.f ? x : y;
#ifndef B2ASR_BEAM_SFU
  const float r = __fadd_rn(m, log1pf(expf(-fabsf(tmp))));
#else
  const float e = ex2(-fabsf(tmp) * kLog2e);
  const float u = 1.f + e;
  const float r = m + fmaf(lg2(u), 0.693147180559945309417232121458176568f, __fdividef(e - (u - 1.f), u));
#endif
  return x == y ? __fadd_rn(x, 0.693147180559945309417232121458176568f) : r;
}

__device__ __forceinline__ unsigned f2key(float f) {   // order-preserving float -> uint
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int kTopkCollect = 256;             // candidates the fast path collects before ranking them exactly
struct TopkScratch {
  int cnt_iter[3];
  int cnt, minidx;
  float tv[kMaxK]; int ti[kMaxK];
  unsigned lmax[kBeamNT];                     // fast path: every thread's largest key
  unsigned thr;                               //   the k-th largest of those = a lower bound of the k-th largest key
  float cv[kTopkCollect]; int ci[kTopkCollect];
};

constexpr int kKeysPerThread = 24;            // register-resident keys: V <= 24*256 = 6144, else the shared-memory loop

// CTA-wide top-k (k <= 32) of vals[0..V): results sorted by (value desc, index asc) in outv/outi (shared).
// The k-th largest key is found by bisection over its 32 bits (one CTA-wide count per bit; the keys
// stay in registers), everything above it is collected, ties at the threshold are taken lowest index
// first, and the k survivors are ordered by a warp-level rank sort.
//
// Fast path (keys in registers, one CTA of kBeamNT threads): the k-th largest of the threads' LOCAL maxima is a lower
// bound of the k-th largest key (k different threads hold a key at least that large), found by one warp without any
// CTA-wide barrier; everything at or above it is collected (typically a few dozen entries) and ranked exactly by
// (value desc, index asc) -- the same total order, hence the same result, as the bisection below, which remains for
// large vocabularies and for inputs with more than kTopkCollect entries at or above the bound (long runs of ties).
__device__ void cta_topk(const float* vals, int V, int k, float* outv, int* outi, TopkScratch* S) {
  const int tid = threadIdx.x, NT = blockDim.x;
  if (k > V) k = V;
  const bool in_regs = V <= kKeysPerThread * NT;
  unsigned keys[kKeysPerThread];
#pragma unroll
  for (int j = 0; j < kKeysPerThread; ++j) {
    const int i = tid + j * NT;
    keys[j] = (in_regs && i < V) ? f2key(vals[i]) : 0u;       // key 0 never reaches a threshold (>= 1)
  }
  if (in_regs && NT == kBeamNT && k >= 1) {
    unsigned lm = 0u;
#pragma unroll
    for (int j = 0; j < kKeysPerThread; ++j) lm = max(lm, keys[j]);
    S->lmax[tid] = lm;
    if (tid == 0) S->cnt = 0;
    __syncthreads();
    if (tid < 32) {                                            // k-th largest of the kBeamNT local maxima, one warp
      unsigned mine[kBeamNT / 32];
#pragma unroll
      for (int i = 0; i < kBeamNT / 32; ++i) mine[i] = S->lmax[tid + 32 * i];
      unsigned prefix = 0u;
      for (int it = 0; it < 32; ++it) {
        const unsigned trial = prefix | (0x80000000u >> it);
        int c = 0;
#pragma unroll
        for (int i = 0; i < kBeamNT / 32; ++i) c += mine[i] >= trial;
        if (__reduce_add_sync(kFull, c) >= k) prefix = trial;
      }
      if (tid == 0) S->thr = prefix;
    }
    __syncthreads();
    const unsigned thr = S->thr;
    if (thr != 0u) {                                           // (0: fewer than k threads hold a key -- tiny V)
#pragma unroll
      for (int j = 0; j < kKeysPerThread; ++j) {
        if (keys[j] >= thr) {
          const int pos = atomicAdd(&S->cnt, 1);
          if (pos < kTopkCollect) { S->cv[pos] = vals[tid + j * NT]; S->ci[pos] = tid + j * NT; }
        }
      }
      __syncthreads();
      const int C = S->cnt;
      if (C <= kTopkCollect) {                                 // (C >= k by construction)
        if (tid < C) {
          const float v = S->cv[tid];
          const int ix = S->ci[tid];
          const unsigned kv = f2key(v);
          int rank = 0;
          for (int m = 0; m < C; ++m) {
            const unsigned km = f2key(S->cv[m]);
            rank += (km > kv) || (km == kv && S->ci[m] < ix);
          }
          if (rank < k) { outv[rank] = v; outi[rank] = ix; }
        }
        __syncthreads();
        return;
      }
    }
    __syncthreads();                                           // fall through to the bisection
  }
  unsigned prefix = 0u;
  if (tid == 0) { S->cnt_iter[0] = 0; S->cnt_iter[1] = 0; S->cnt_iter[2] = 0; }
  __syncthreads();
  for (int it = 0; it < 32; ++it) {
    const unsigned trial = prefix | (0x80000000u >> it);
    int c = 0;
    if (in_regs) {
#pragma unroll
      for (int j = 0; j < kKeysPerThread; ++j) c += keys[j] >= trial;
    } else {
      for (int i = tid; i < V; i += NT) c += f2key(vals[i]) >= trial;
    }
    c = __reduce_add_sync(kFull, c);
    int* slot = &S->cnt_iter[it % 3];
    if ((tid & 31) == 0 && c) atomicAdd(slot, c);
    __syncthreads();
    if (*slot >= k) prefix = trial;             // the k-th largest key has this bit set
    // re-arm the slot of iteration it+2: its last readers (iteration it-1) all passed the barrier above,
    // and nobody adds to it before the next barrier
    if (tid == 0) S->cnt_iter[(it + 2) % 3] = 0;
  }
  __syncthreads();
  const unsigned kth = prefix;
  if (tid == 0) S->cnt = 0;
  __syncthreads();
  for (int i = tid; i < V; i += NT) {
    if (f2key(vals[i]) > kth) { const int pos = atomicAdd(&S->cnt, 1); if (pos < kMaxK) { S->tv[pos] = vals[i]; S->ti[pos] = i; } }
  }
  __syncthreads();
  const int need = k - S->cnt;                 // entries equal to the threshold still to take (>= 1)
  __syncthreads();
  int last = -1;
  for (int r = 0; r < need; ++r) {             // usually a single round
    if (tid == 0) S->minidx = 0x7fffffff;
    __syncthreads();
    int best = 0x7fffffff;
    for (int i = tid; i < V; i += NT) if (i > last && f2key(vals[i]) == kth) { best = i; break; }
    if (best != 0x7fffffff) atomicMin(&S->minidx, best);
    __syncthreads();
    last = S->minidx;
    if (tid == 0 && last != 0x7fffffff) { const int pos = S->cnt++; if (pos < kMaxK) { S->tv[pos] = vals[last]; S->ti[pos] = last; } }
    __syncthreads();
  }
  // warp-level rank sort
  if (tid < 32) {
    const int n = min(S->cnt, k);
    const float v = tid < n ? S->tv[tid] : -INFINITY;
    const int ix = tid < n ? S->ti[tid] : 0x7fffffff;
    int rank = 0;
#pragma unroll
    for (int m = 0; m < 32; ++m) {
      const float vm = __shfl_sync(kFull, v, m);
      const int im = __shfl_sync(kFull, ix, m);
      if (m < n && (vm > v || (vm == v && im < ix))) ++rank;
    }
    if (tid < n) { outv[rank] = v; outi[rank] = ix; }
    for (int j = n + tid; j < k; j += 32) { outv[j] = -INFINITY; outi[j] = -1; }
  }
  __syncthreads();
}

// Shared-memory frame tables of one hypothesis, filled by the whole CTA.
// xk [32][xk_pitch(T)]: candidate-major; the pitch is odd, so the recursion's lanes (one candidate each, same frame) hit
// 32 different banks and the fill writes runs of consecutive frames
struct PrefixTables { float* xk; float* xb; float* phiD; float* phiS; };
__host__ __device__ inline int xk_pitch(int T) { return T | 1; }

__device__ void prefix_fill_tables(const PrefixTables& tb, const float* x, int T, int V, const float* r_prev,
                                   const int* cand_sm, int K, int start, const float* xt = nullptr, long long xt_ld = 0) {
  const int tid = threadIdx.x, NT = blockDim.x;
  const int first = start == 1 ? 0 : start;        // row 0 is read only for the empty prefix (start == 1)
  const int TP = xk_pitch(T);
  if (tb.xk == nullptr) {
    // long utterance: the recursion reads the candidates' rows of the transposed table directly
  } else if (xt) {
    // transposed posteriors: candidate k's column is the contiguous row xt[cand[k]][0:T] -- a warp reads 32
    // consecutive frames of one candidate (one 128-byte line) and scatters them down column k of xk
    const int lane = tid & 31, warp = tid >> 5, nw = NT >> 5;
    for (int k = warp; k < 32; k += nw) {
      const float* row = xt + (long long)cand_sm[k < K ? k : 0] * xt_ld;
      for (int t = lane; t < T; t += 32) tb.xk[k * TP + t] = (k < K && t >= first) ? __ldg(row + t) : 0.f;
    }
  } else {
#pragma unroll 8
    for (int i = first * 32 + tid; i < T * 32; i += NT) {
      const int t = i >> 5, k = i & 31;
      tb.xk[k * TP + t] = k < K ? __ldg(x + (long long)t * V + cand_sm[k]) : 0.f;
    }
  }
  for (int t = tid; t < T; t += NT) {
    tb.xb[t] = xt ? __ldg(xt + t) : __ldg(x + (long long)t * V);  // blank = 0 (ctc.py:12): row 0 of the transposed table
    if (t >= start) {
      const float pnb = r_prev[2 * (t - 1)], pb = r_prev[2 * (t - 1) + 1];
      tb.phiD[t] = np_logaddexp(pnb, pb);                        // ctc.py:87-93
      tb.phiS[t] = np_logaddexp(pnb, kLogZero);                  // candidate repeats the last label
    }
  }
}

// The sequential recursion (ctc.py:85-98).  Executed by warp 0, lane = candidate: a lane carries its candidate's three
// running values -- r[t,0] (prefix ends in the candidate), r[t,1] (ends in blank) and psi -- whose three logaddexp per
// frame are independent of each other, so their latencies overlap inside the thread and nothing crosses lanes
// (one lane per value with a shuffle per frame was 307 clk per frame: a pure dependency chain).  r rows are written
// with the caller's strides: element (t,k,j) at r_out[t*st_t + k*st_k + j].
__device__ void prefix_recursion(const PrefixTables& tb, int T, int K, int prefix_len, int last,
                                 const int* cand_sm, float* psi_out, float* r_out, long long st_t, long long st_k) {
  const int k = threadIdx.x & 31;
  if (k >= K) return;
  const int start = max(1, prefix_len);
  const bool same = cand_sm[k] == last;
  float* ro = r_out + k * st_k;
  // rows before start-1 (and the blank column of row start-1) are "log zero" (ctc.py:73)
  for (int t = 0; t < min(start, T); ++t) { ro[t * st_t] = kLogZero; ro[t * st_t + 1] = kLogZero; }
  float rn = kLogZero, rb = kLogZero;
  const float* xkk = tb.xk + k * xk_pitch(T);                   // this candidate's column of the posteriors
  if (prefix_len == 0) rn = xkk[0];                             // r[0,0,:] = x[0,cands]
  float psi = rn;                                               // psi = r[start-1,0,:]
  if (start - 1 < T) ro[(start - 1) * st_t] = rn;
  const float* phis = same ? tb.phiS : tb.phiD;
#pragma unroll 2
  for (int t = start; t < T; ++t) {
    const float phi = phis[t], xe = xkk[t], xbt = tb.xb[t];
    const float n0 = __fadd_rn(np_logaddexp(rn, phi), xe);      // ctc.py:94
    const float n1 = __fadd_rn(np_logaddexp(rb, rn), xbt);      // ctc.py:95 (reads r[t-1,0])
    psi = np_logaddexp(psi, __fadd_rn(phi, xe));                // ctc.py:96
    rn = n0; rb = n1;
    ro[t * st_t] = n0; ro[t * st_t + 1] = n1;
  }
  psi_out[k] = psi;
}

// The same recursion for utterances whose [32][T'] candidate table does not fit in shared memory: the lane reads its
// candidate's posteriors from a global row (`xcol` = row cand[k] of the transposed table), eight frames ahead of their
// use.  Same values, same arithmetic; a function of its own so that the common path above keeps its code.
__device__ __noinline__ void prefix_recursion_global(const PrefixTables& tb, int T, int K, int prefix_len, int last,
                                                     const int* cand_sm, float* psi_out, float* r_out, long long st_t,
                                                     long long st_k, const float* xt, long long xt_ld) {
  const int k = threadIdx.x & 31;
  if (k >= K) return;
  const float* xcol = xt + (long long)cand_sm[k] * xt_ld;
  const int start = max(1, prefix_len);
  const bool same = cand_sm[k] == last;
  float* ro = r_out + k * st_k;
  for (int t = 0; t < min(start, T); ++t) { ro[t * st_t] = kLogZero; ro[t * st_t + 1] = kLogZero; }
  float rn = kLogZero, rb = kLogZero;
  if (prefix_len == 0) rn = __ldg(xcol);
  float psi = rn;
  if (start - 1 < T) ro[(start - 1) * st_t] = rn;
  const float* phis = same ? tb.phiS : tb.phiD;
  constexpr int PF = 8;
  float q[PF];
#pragma unroll
  for (int u = 0; u < PF; ++u) q[u] = start + u < T ? __ldg(xcol + start + u) : 0.f;
  for (int t0 = start; t0 < T; t0 += PF) {
    float nq[PF];
#pragma unroll
    for (int u = 0; u < PF; ++u) nq[u] = t0 + PF + u < T ? __ldg(xcol + t0 + PF + u) : 0.f;
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const int t = t0 + u;
      if (t < T) {
        const float phi = phis[t], xe = q[u], xbt = tb.xb[t];
        const float n0 = __fadd_rn(np_logaddexp(rn, phi), xe);
        const float n1 = __fadd_rn(np_logaddexp(rb, rn), xbt);
        psi = np_logaddexp(psi, __fadd_rn(phi, xe));
        rn = n0; rb = n1;
        ro[t * st_t] = n0; ro[t * st_t + 1] = n1;
      }
    }
#pragma unroll
    for (int u = 0; u < PF; ++u) q[u] = nq[u];
  }
  psi_out[k] = psi;
}

__host__ __device__ inline size_t prefix_smem_floats(int T) { return (size_t)xk_pitch(T) * 32 + 3 * (size_t)((T + 3) / 4 * 4); }

// ------------------------------------------------------------------------------------------------
// Stand-alone scorer kernels (API parity with CTCPrefixScore)
// ------------------------------------------------------------------------------------------------
__global__ void prefix_init_kernel(const float* x, int T, int V, int blank, float* r) {
  // sequential fp32 running sum, one thread: bit-exact with ctc.py:21-27
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float acc = 0.f;
    for (int t = 0; t < T; ++t) {
      const float xv = x[(long long)t * V + blank];
      acc = t == 0 ? xv : __fadd_rn(acc, xv);
      r[2 * t] = kLogZero; r[2 * t + 1] = acc;
    }
  }
}

__global__ void __launch_bounds__(kBeamNT) prefix_score_kernel(const float* x, int T, int V, int prefix_len, int last,
                                                              const float* r_prev, const int* cands, int K,
                                                              float* psi, float* r_out) {
  extern __shared__ __align__(16) float smf[];
  __shared__ int cand_sm[kMaxK];
  const int Tp = (T + 3) / 4 * 4;
  const size_t xkf = (size_t)xk_pitch(T) * 32;
  PrefixTables tb{smf, smf + xkf, smf + xkf + Tp, smf + xkf + 2 * Tp};
  if (threadIdx.x < kMaxK) cand_sm[threadIdx.x] = threadIdx.x < K ? min(V - 1, max(0, cands[threadIdx.x])) : 0;
  __syncthreads();
  prefix_fill_tables(tb, x, T, V, r_prev, cand_sm, K, max(1, prefix_len));
  __syncthreads();
  if (threadIdx.x < 32) prefix_recursion(tb, T, K, prefix_len, last, cand_sm, psi, r_out, 2, (long long)T * 2);
}

// ------------------------------------------------------------------------------------------------
// Fused beam step
// ------------------------------------------------------------------------------------------------
// The best n children in stable descending order of their mean score (asr.py:251), by ONE warp.  Every row's children
// arrive sorted (cta_topk returns value desc, index asc, and the mean is a monotone function of the value), so this is
// an n-way merge of n sorted lists: lane r holds the head of row r, the warp takes the largest head n times
// (`redux.sync.max` on the order-preserving key; among equal heads the lowest row, i.e. the stable order) and only the
// winner advances.  Slots without a child (<eos>, dead rows) are skipped, as the reference never creates them.
__device__ __forceinline__ void rank_children(const float* avg, const int* slotk, int* slot_child, unsigned* ckey, int n) {
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  const int base = lane * (n + 1);
  // the row's children compacted to (key, position) pairs with a 0 terminator, so that the winner's advance is one
  // load at a known address, issued a round before its value is needed
  unsigned char* cpos = reinterpret_cast<unsigned char*>(ckey + n * (n + 1));
  if (lane < n) {
    int q = 0;
    for (int p = 0; p < n; ++p) {
      const int sk = slotk[lane * n + p];
      const unsigned k = f2key(avg[lane * n + p]);
      if (sk != -2) { ckey[base + q] = k; cpos[lane * n + q] = (unsigned char)p; ++q; }
    }
    ckey[base + q] = 0u;
  }
  __syncwarp();
  int q = 0;
  unsigned key = lane < n ? ckey[base] : 0u;                   // 0: row exhausted (every real key is > 0)
  unsigned nkey = key != 0u ? ckey[base + 1] : 0u;
  for (int j = 0; j < n; ++j) {
    const unsigned mx = __reduce_max_sync(kFull, key);
    if (mx == 0u) break;
    const int w = __ffs(__ballot_sync(kFull, key == mx)) - 1;
    if (lane == w) {
      slot_child[j] = lane * n + cpos[lane * n + q];
      ++q;
      key = nkey;
      nkey = key != 0u ? ckey[base + q + 1] : 0u;
    }
  }
}

// debug aid: clock64 stamps of CTA 0's phases (slots 0-7) and of the pruning CTA (8-13); see b2asr_debug_beam_trace
__device__ long long g_beam_trace[16];
#define BT(slot) do { if (threadIdx.x == 0 && ((slot) >= 8 || blockIdx.x == 0)) g_beam_trace[slot] = clock64(); } while (0)

__global__ void __launch_bounds__(kBeamNT) beam_step_kernel(b2asr_beam_io io) {
  extern __shared__ __align__(16) float smf[];
  __shared__ TopkScratch S;
  __shared__ int cand_sm[kMaxK];
  __shared__ float candv_sm[kMaxK];
  __shared__ float psi_sm[kMaxK];
  __shared__ float topv_sm[kMaxK];
  __shared__ int topi_sm[kMaxK];
  __shared__ unsigned is_last;

  const int tid = threadIdx.x, NT = blockDim.x;
  const int row = blockIdx.x, n = io.n, V = io.V, T = io.T, K = io.K;
  const bool ctc = io.x_ctc != nullptr;
  const int plen = *io.step;
  float* val = smf;                                            // [V] attention log-probs, then fused scores
  const int Vp = (V + 3) / 4 * 4;
  const bool live = io.valid_in[row] != 0;

  BT(0);
  if (live) {
    const float* att = io.att_logp + (long long)row * V;
    if (io.att_lse) {                                            // logits in: log_softmax = logit - lse (asr.py:210)
      const float lse = io.att_lse[row];
      for (int i = tid; i < V; i += NT) val[i] = __fsub_rn(att[i], lse);
    } else {
      for (int i = tid; i < V; i += NT) val[i] = att[i];
    }
    __syncthreads();
    if (ctc) {
      // ---- CTC candidates: top-K of the attention log-probs (asr.py:215-216) -----------------------
      BT(1);
      cta_topk(val, V, K, candv_sm, cand_sm, &S);
      BT(2);
      const int Tp = (T + 3) / 4 * 4;
      float* base = smf + Vp;
      // long utterances (V + 35 T' floats beyond the 200 KB the launch may ask for): the launch carries shared memory
      // for the three per-frame vectors only and the [32][T'] candidate table stays in the transposed global copy
      unsigned dyn_smem;
      asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_smem));
      const bool xk_global = ((size_t)Vp + prefix_smem_floats(T)) * sizeof(float) > dyn_smem;
      const size_t xkf = xk_global ? 0 : (size_t)xk_pitch(T) * 32;
      PrefixTables tb{xk_global ? nullptr : base, base + xkf, base + xkf + Tp, base + xkf + 2 * Tp};
      const float* r_prev = io.r_in + (long long)row * T * 2;
      prefix_fill_tables(tb, io.x_ctc, T, V, r_prev, cand_sm, K, max(1, plen), io.x_ctc_t, io.x_ctc_t_ld);
      __syncthreads();
      BT(3);
      // ---- prefix scores of prefix+[c] (ctc.py:65-101) ------------------------------------------------
      if (tid < 32) {
        float* rall = io.r_all + (long long)row * T * K * 2;    // [K][T][2]: a survivor's rows are contiguous
        if (!xk_global) prefix_recursion(tb, T, K, plen, (int)io.last_in[row], cand_sm, psi_sm, rall, 2, (long long)T * 2);
        else prefix_recursion_global(tb, T, K, plen, (int)io.last_in[row], cand_sm, psi_sm, rall, 2, (long long)T * 2,
                                     io.x_ctc_t, io.x_ctc_t_ld);
      }
      __syncthreads();
      BT(4);
      // ---- fusion (asr.py:222-229): separate fp32 multiply and add roundings, as torch does -----------
      const float neg = __fmul_rn(io.w_ctc, kNegCtc);
      for (int i = tid; i < V; i += NT) val[i] = __fadd_rn(__fmul_rn(io.w_att, val[i]), neg);
      __syncthreads();
      if (tid < K) {
        const float ctc_char = __fsub_rn(psi_sm[tid], io.ctc_prob_in[row]);      // asr.py:222
        val[cand_sm[tid]] = __fadd_rn(__fmul_rn(io.w_att, candv_sm[tid]), __fmul_rn(io.w_ctc, ctc_char));
        io.cand[row * K + tid] = cand_sm[tid];
        io.psi[row * K + tid] = psi_sm[tid];
      }
      __syncthreads();
      if (tid == 0) val[0] = kNegSos;                                           // asr.py:229
      __syncthreads();
    }
    if (io.lm_logp != nullptr && io.w_lm > 0.f) {                               // asr.py:232-235
      const float* lm = io.lm_logp + (long long)row * V;
      if (io.lm_lse) {
        const float lse = io.lm_lse[row];
        for (int i = tid; i < V; i += NT) val[i] = __fadd_rn(val[i], __fmul_rn(io.w_lm, __fsub_rn(lm[i], lse)));
      } else {
        for (int i = tid; i < V; i += NT) val[i] = __fadd_rn(val[i], __fmul_rn(io.w_lm, lm[i]));
      }
      __syncthreads();
    }
    // ---- top-`beam` of the fused scores (asr.py:238) ---------------------------------------------------
    BT(5);
    cta_topk(val, V, n, topv_sm, topi_sm, &S);
    BT(6);
    if (tid < n) { io.topv[row * n + tid] = topv_sm[tid]; io.topi[row * n + tid] = topi_sm[tid]; }
  }

  // ---- the last CTA to arrive prunes (asr.py:240-253, postprocess.py:71-106) ---------------------------
  __threadfence();
  __syncthreads();
  if (tid == 0) is_last = (atomicAdd(io.ticket, 1u) == (unsigned)n - 1u);
  __syncthreads();
  BT(7);
  if (!is_last) return;
  __threadfence();
  BT(8);

  float* avg = smf;                                            // [n*n] mean score of every child (or -inf)
  int* slotk = reinterpret_cast<int*>(smf + n * n);            // [n*n] candidate slot of the child's token
  const float cnt = (float)(plen + 1);
  const int nc = n * n;
  int* slot_child = slotk + nc;                                // [n] child index occupying each survivor slot
  int* cand_all = slot_child + n;                              // [n*K] every row's CTC candidates, staged once
  unsigned* ckey = reinterpret_cast<unsigned*>(cand_all + (ctc ? n * K : 0));   // [n*(n+1)] + [n*n] bytes: rank_children's lists
  for (int base = 0; base < nc; base += 2 * NT) {             // two children per thread; every load of a round is
    int tok[2], vin[2]; float tv[2], sum[2];                   // issued before the first one is consumed
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int c = min(base + u * NT + tid, nc - 1);
      const int r = c / n;
      vin[u] = io.valid_in[r];
      tok[u] = __ldcg(io.topi + c);
      tv[u] = __ldcg(io.topv + c);
      sum[u] = io.sum_in[r];
    }
    if (base == 0) {
      if (ctc) for (int i = tid; i < n * K; i += NT) cand_all[i] = __ldcg(io.cand + i);
      for (int j = tid; j < n; j += NT) slot_child[j] = -1;
      __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int c = base + u * NT + tid;
      if (c >= nc) break;
      const int r = c / n;
      float a = -INFINITY; int sk = -2;                         // -2: no child here; -1: child, token not a CTC candidate
      if (vin[u] && tok[u] >= 0 && tok[u] != kEos) {
        a = __fdiv_rn(__fadd_rn(sum[u], tv[u]), cnt);           // postprocess.py:71-74
        sk = -1;
        if (ctc) {                                              // a row's candidates are distinct: no early exit needed,
          const int* cr = cand_all + r * K;                     // so the K compares are independent
          for (int j = 0; j < K; ++j) sk = cr[j] == tok[u] ? j : sk;
        }
      }
      avg[c] = a; slotk[c] = sk;
    }
  }
  BT(9);
  __syncthreads();
  // the best n children in stable descending order (asr.py:251): n-way merge of the rows' sorted children
  rank_children(avg, slotk, slot_child, ckey, n);
  __syncthreads();
  BT(10);
  const long long hoff = (long long)min(plen, io.max_steps - 1) * n;
  for (int j = tid; j < n; j += NT) {
    const int c = slot_child[j];
    if (c >= 0) {
      const int r = c / n;
      const float tv = __ldcg(io.topv + c);
      const int tok = __ldcg(io.topi + c);
      io.valid_out[j] = 1; io.last_out[j] = tok; io.parent_out[j] = r;
      io.sum_out[j] = __fadd_rn(io.sum_in[r], tv);
      io.ctc_prob_out[j] = (ctc && slotk[c] >= 0) ? __ldcg(io.psi + r * K + slotk[c]) : 0.f;
      io.hist_parent[hoff + j] = r; io.hist_tok[hoff + j] = tok; io.hist_score[hoff + j] = tv;
    } else {
      io.valid_out[j] = 0; io.last_out[j] = 0; io.parent_out[j] = 0; io.sum_out[j] = 0.f; io.ctc_prob_out[j] = 0.f;
      io.hist_parent[hoff + j] = -1; io.hist_tok[hoff + j] = -1; io.hist_score[hoff + j] = 0.f;
    }
  }
  BT(11);
  if (ctc) {                                                   // survivors inherit r[cand slot] (postprocess.py:95-97)
    // r_all is candidate-major, so a survivor's T rows (r[t,0], r[t,1]) are one contiguous run of the parent's slot
    const float2* rall2 = reinterpret_cast<const float2*>(io.r_all);
    float2* rout2 = reinterpret_cast<float2*>(io.r_out);
    // one warp per survivor: its T rows are one contiguous run, every lane keeps all of its loads in flight
    const int warp = tid >> 5, lane = tid & 31, nw = NT >> 5;
    for (int j = warp; j < n; j += nw) {
      const int c = slot_child[j];
      const int sk = c >= 0 ? slotk[c] : -1;
      float2* dst = rout2 + (long long)j * T;
      if (sk < 0) {
        for (int t = lane; t < T; t += 32) dst[t] = make_float2(kLogZero, kLogZero);
        continue;
      }
      const float2* src = rall2 + ((long long)(c / n) * K + sk) * T;
      for (int t0 = 0; t0 < T; t0 += 32 * 16) {
        float2 v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) { const int t = t0 + u * 32 + lane; if (t < T) v[u] = __ldcg(src + t); }
#pragma unroll
        for (int u = 0; u < 16; ++u) { const int t = t0 + u * 32 + lane; if (t < T) dst[t] = v[u]; }
      }
    }
  }
  __syncthreads();
  BT(12);
  int* eos_at = slot_child;                                    // reuse: [n] child index holding <eos> per row, or -1
  for (int j = tid; j < n; j += NT) eos_at[j] = -1;
  __syncthreads();
  for (int c = tid; c < nc; c += NT)
    if (io.valid_in[c / n] && __ldcg(io.topi + c) == kEos) eos_at[c / n] = c;   // top-k indices are unique per row
  __syncthreads();
  if (tid == 0) {                                              // <eos> finalises its parent, in row order (asr.py:244-245)
    int f = *io.fin_count;
    for (int r = 0; r < n; ++r) {
      const int c = eos_at[r];
      if (c < 0) continue;
      const float tv = __ldcg(io.topv + c);
      if (f < io.max_steps * n) { io.fin_step[f] = plen; io.fin_row[f] = r; io.fin_score[f] = tv; io.fin_sum[f] = __fadd_rn(io.sum_in[r], tv); ++f; }
    }
    *io.fin_count = f;
    *io.step = plen + 1;
    *io.ticket = 0u;
  }
  BT(13);
}

}  // namespace b2asr

using namespace b2asr;

extern "C" size_t b2asr_beam_io_size(void) { return sizeof(b2asr_beam_io); }

extern "C" int b2asr_ctc_prefix_init(const float* x, int T, int V, int blank, float* r_out, void* stream) {
  B2_REQUIRE(x && r_out, B2ASR_EINVAL, "ctc_prefix_init: null pointer argument");
  B2_REQUIRE(T > 0 && V > 0 && blank >= 0 && blank < V, B2ASR_EINVAL, "ctc_prefix_init: bad sizes T=%d V=%d blank=%d", T, V, blank);
  prefix_init_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(x, T, V, blank, r_out);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

extern "C" int b2asr_ctc_prefix_score(const float* x, int T, int V, int prefix_len, int last,
                                      const float* r_prev, const int32_t* cands, int K,
                                      float* psi, float* r_out, void* stream) {
  B2_REQUIRE(x && r_prev && cands && psi && r_out, B2ASR_EINVAL, "ctc_prefix_score: null pointer argument");
  B2_REQUIRE(T > 0 && V > 0 && prefix_len >= 0, B2ASR_EINVAL, "ctc_prefix_score: bad sizes T=%d V=%d", T, V);
  B2_REQUIRE(K >= 1 && K <= kMaxK, B2ASR_EUNSUPPORTED, "ctc_prefix_score: K=%d candidates per call (1..%d)", K, kMaxK);
  const size_t smem = prefix_smem_floats(T) * sizeof(float);
  B2_REQUIRE(smem <= 200 * 1024, B2ASR_EUNSUPPORTED, "ctc_prefix_score: T=%d frames need %zu bytes of shared memory", T, smem);
  B2_CUDA_TRY(cudaFuncSetAttribute(prefix_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  prefix_score_kernel<<<1, kBeamNT, smem, (cudaStream_t)stream>>>(x, T, V, prefix_len, last, r_prev, cands, K, psi, r_out);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

extern "C" int b2asr_beam_step(const b2asr_beam_io* io, void* stream) {
  B2_REQUIRE(io != nullptr, B2ASR_EINVAL, "beam_step: null io");
  const b2asr_beam_io& a = *io;
  B2_REQUIRE(a.att_logp && a.valid_in && a.last_in && a.sum_in && a.valid_out && a.last_out && a.sum_out && a.parent_out &&
             a.topv && a.topi && a.step && a.ticket && a.hist_parent && a.hist_tok && a.hist_score && a.fin_count &&
             a.fin_step && a.fin_row && a.fin_score && a.fin_sum && a.ctc_prob_out, B2ASR_EINVAL, "beam_step: null pointer argument");
  B2_REQUIRE(a.n >= 1 && a.n <= kMaxK && a.V >= a.n && a.max_steps >= 1, B2ASR_EUNSUPPORTED, "beam_step: beam %d (1..%d), V=%d", a.n, kMaxK, a.V);
  size_t floats = (size_t)((a.V + 3) / 4 * 4);
  if (a.x_ctc) {
    B2_REQUIRE(a.K >= 1 && a.K <= kMaxK && a.K <= a.V && a.T > 0, B2ASR_EUNSUPPORTED, "beam_step: K=%d CTC candidates (1..%d), T=%d", a.K, kMaxK, a.T);
    B2_REQUIRE(a.r_in && a.r_out && a.ctc_prob_in && a.cand && a.psi && a.r_all, B2ASR_EINVAL, "beam_step: joint CTC needs r_in/r_out/ctc_prob_in/cand/psi/r_all");
    floats += prefix_smem_floats(a.T);
    if (floats * sizeof(float) > 200 * 1024 && a.x_ctc_t)          // long utterance: candidate table read from x_ctc_t
      floats = (size_t)((a.V + 3) / 4 * 4) + 3 * (size_t)((a.T + 3) / 4 * 4);
  }
  const size_t prune = (size_t)4 * a.n * a.n + 2 * a.n + (size_t)a.n * (a.x_ctc ? a.K : 0) + 2 + 2 * (size_t)a.n;
  if (floats < prune) floats = prune;
  const size_t smem = floats * sizeof(float);
  B2_REQUIRE(smem <= 200 * 1024, B2ASR_EUNSUPPORTED, "beam_step: V=%d, T=%d need %zu bytes of shared memory", a.V, a.T, smem);
  B2_CUDA_TRY(cudaFuncSetAttribute(beam_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  beam_step_kernel<<<a.n, kBeamNT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA_TRY(cudaGetLastError());
  return B2ASR_OK;
}

// debug aid: phase stamps of the last beam_step launch (synchronous copy)
extern "C" int b2asr_debug_beam_trace(long long* host_out, int n) {
  B2_REQUIRE(host_out && n > 0 && n <= 16, B2ASR_EINVAL, "beam trace: bad arguments");
  B2_CUDA_TRY(cudaMemcpyFromSymbol(host_out, b2asr::g_beam_trace, sizeof(long long) * n));
  return B2ASR_OK;
}
