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

struct TopkScratch {
  int cnt_iter[3];
  int cnt, minidx;
  float tv[kMaxK]; int ti[kMaxK];
};

constexpr int kKeysPerThread = 24;            // register-resident keys: V <= 24*256 = 6144, else the shared-memory loop

// CTA-wide top-k (k <= 32) of vals[0..V): results sorted by (value desc, index asc) in outv/outi (shared).
// The k-th largest key is found by bisection over its 32 bits (one CTA-wide count per bit; the keys
// stay in registers), everything above it is collected, ties at the threshold are taken lowest index
// first, and the k survivors are ordered by a warp-level rank sort.
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
  if (xt) {
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
// Stable descending rank of every child (asr.py:251): one warp per child; the lane's competitors (32*Q >= nc of
// them) stay in registers while the warp walks its children.  A slot without a child holds -inf: it never
// outranks anything and is never ranked itself.
template <int Q>
__device__ __forceinline__ void rank_children(const float* avg, const int* slotk, int* slot_child, int nc, int n) {
  const int lane = threadIdx.x & 31;
  float mine[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) { const int m = lane + 32 * q; mine[q] = m < nc ? avg[m] : -INFINITY; }
  for (int c = threadIdx.x >> 5; c < nc; c += blockDim.x >> 5) {
    if (slotk[c] == -2) continue;                              // warp-uniform
    const float a = avg[c];
    int rank = 0;
#pragma unroll
    for (int q = 0; q < Q; ++q) rank += (mine[q] > a) || (mine[q] == a && lane + 32 * q < c);
    rank = __reduce_add_sync(kFull, rank);
    if (lane == 0 && rank < n) slot_child[rank] = c;
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
      const size_t xkf = (size_t)xk_pitch(T) * 32;
      PrefixTables tb{base, base + xkf, base + xkf + Tp, base + xkf + 2 * Tp};
      const float* r_prev = io.r_in + (long long)row * T * 2;
      prefix_fill_tables(tb, io.x_ctc, T, V, r_prev, cand_sm, K, max(1, plen), io.x_ctc_t, io.x_ctc_t_ld);
      __syncthreads();
      BT(3);
      // ---- prefix scores of prefix+[c] (ctc.py:65-101) ------------------------------------------------
      if (tid < 32)
        prefix_recursion(tb, T, K, plen, (int)io.last_in[row], cand_sm, psi_sm,
                         io.r_all + (long long)row * T * K * 2, (long long)K * 2, 2);
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
  if (ctc) for (int i = tid; i < n * K; i += NT) cand_all[i] = __ldcg(io.cand + i);
  for (int j = tid; j < n; j += NT) slot_child[j] = -1;
  __syncthreads();
  for (int c = tid; c < nc; c += NT) {
    const int r = c / n;
    float a = -INFINITY; int sk = -2;                           // -2: no child here; -1: child, token not a CTC candidate
    if (io.valid_in[r]) {
      const int tok = __ldcg(io.topi + c);
      if (tok >= 0 && tok != kEos) {
        a = __fdiv_rn(__fadd_rn(io.sum_in[r], __ldcg(io.topv + c)), cnt);      // postprocess.py:71-74
        sk = -1;
        if (ctc) for (int j = 0; j < K; ++j) if (cand_all[r * K + j] == tok) { sk = j; break; }
      }
    }
    avg[c] = a; slotk[c] = sk;
  }
  BT(9);
  __syncthreads();
  // stable descending rank of every child (asr.py:251): one warp per child, lanes stride the competitors.
  // A slot without a child holds -inf: it never outranks anything and is never ranked itself.
  if (nc <= 512) rank_children<16>(avg, slotk, slot_child, nc, n); else rank_children<32>(avg, slotk, slot_child, nc, n);
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
    // one warp-strided pass per survivor: (r[t,0], r[t,1]) is one 8-byte load from the parent's candidate slot
    const float2* rall2 = reinterpret_cast<const float2*>(io.r_all);
    float2* rout2 = reinterpret_cast<float2*>(io.r_out);
    // survivors' source offsets once, then a flat (survivor, frame) sweep with 8 loads in flight per thread
    long long* src_of = reinterpret_cast<long long*>(reinterpret_cast<int*>(smf) + ((2 * nc + n + n * K + 1) & ~1));   // [n], 8-byte aligned
    for (int j = tid; j < n; j += NT) {
      const int c = slot_child[j];
      src_of[j] = (c >= 0 && slotk[c] >= 0) ? ((long long)(c / n) * T * K + slotk[c]) : -1;
    }
    __syncthreads();
    const int total = n * T;
    for (int i0 = 0; i0 < total; i0 += 8 * NT) {
      float2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * NT + tid;
        v[u] = make_float2(kLogZero, kLogZero);
        if (i < total) {
          const int j = i / T, t = i - j * T;
          const long long so = src_of[j];
          if (so >= 0) v[u] = __ldcg(rall2 + so + (long long)t * K);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * NT + tid;
        if (i < total) rout2[i] = v[u];
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
  }
  const size_t prune = (size_t)2 * a.n * a.n + a.n + (size_t)a.n * (a.x_ctc ? a.K : 0) + 2 + 2 * (size_t)a.n;
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
