// Kernel group (2b), backward: gradient of one Attention.forward step (reference src/asr.py:410-462;
// the reference gets it from autograd over ~12 ATen ops).  The forward's tanh activations are NOT
// stored (T'xA per step per utterance); they are recomputed here from the saved query, the previous
// alignment and the alignment itself.
//
//   given  d_ctx [B,D], d_att [B,T] (what the NEXT step's location conv sent back; may be null)
//   d_alpha = d_att + h . d_ctx                 de = scale * alpha * (d_alpha - <alpha, d_alpha>)
//   loc:  u = tanh(K + q + l), l = tanh(Wl f), f = conv(prev_att)
//         d_pre = de * wg * (1-u^2)             dK += d_pre          dq = sum_t d_pre
//         d_zl = d_pre * (1-l^2)                dWl += d_zl^T f      df = d_zl Wl
//         dWc += df (x) prev_att                d_prev_att = flipped conv of df with Wc
//   dot:  dK += de (x) q, dq = de^T K
//   d_zq = dq * (1-q^2), d_s = d_zq Wphi
// Parameter gradients go to one private slot per CTA (summed once per batch on the host side), the
// encoder-feature gradient dh = alpha (x) d_ctx is deferred to one batched GEMM per batch.
#include "attn.cuh"

namespace b2asr {

struct BwdExtra { size_t de, al, wls, dfo, dfw, dqp, dzq, tmp, total; };

__host__ __device__ inline BwdExtra attn_bwd_extra(const AttnArgs& p) {
  BwdExtra x; size_t o = 0;
  const int CP = (p.C + 3) / 4 * 4;
  auto take = [&](size_t n) { size_t at = o; o += (n + 3) / 4 * 4; return at; };
  x.de = take(p.chunk_max + 4);
  x.al = take(p.chunk_max + 4);
  x.wls = take((size_t)p.A * CP);                        // Wl as [A][CP] for the df pass
  x.dfo = take((size_t)p.chunk_max * CP);
  x.dfw = take((size_t)(p.chunk_max + 2 * p.Kw) * CP);
  x.dqp = take(p.A);
  x.dzq = take(p.A);
  { const size_t t1 = (size_t)(CP / 4) * p.chunk_max, t2 = 4 * (size_t)((p.dec + p.CS - 1) / p.CS); x.tmp = take(t1 > t2 ? t1 : t2); }
  x.total = o;
  return x;
}

// sum 16 per-lane values across the warp with 16 shuffles; on return lane L (L even) holds the
// total of value index ((L>>1)&15 bit-reversed per stage, see idx below) in v[0].
__device__ __forceinline__ float warp_reduce16(float (&v)[16], int lane, int& idx) {
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) { const float send = up ? v[i] : v[i + 8]; const float keep = up ? v[i + 8] : v[i]; v[i] = keep + __shfl_xor_sync(kFull, send, 16); }
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float send = up ? v[i] : v[i + 4]; const float keep = up ? v[i + 4] : v[i]; v[i] = keep + __shfl_xor_sync(kFull, send, 8); }
  }
  {
    const bool up = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) { const float send = up ? v[i] : v[i + 2]; const float keep = up ? v[i + 2] : v[i]; v[i] = keep + __shfl_xor_sync(kFull, send, 4); }
  }
  {
    const bool up = lane & 2;
    const float send = up ? v[0] : v[1]; const float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(kFull, send, 2);
  }
  v[0] += __shfl_xor_sync(kFull, v[0], 1);
  idx = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
  return v[0];
}

template <int CPT, int MAXNT, int MINB>
__global__ void __launch_bounds__(MAXNT, MINB) attn_step_bwd_kernel(AttnArgs p) {
  extern __shared__ __align__(128) unsigned char smraw[];
  const BwdExtra X = attn_bwd_extra(p);
  const AttnSmem L = attn_smem_layout(p, X.total);
  float* ring = reinterpret_cast<float*>(smraw + L.off_ring);
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smraw + L.off_bar);
  float* q_sm = reinterpret_cast<float*>(smraw + L.off_q);
  float* wcT = reinterpret_cast<float*>(smraw + L.off_wc);
  float* pa = reinterpret_cast<float*>(smraw + L.off_pa);
  float* fsm = reinterpret_cast<float*>(smraw + L.off_f);
  float* dctx = reinterpret_cast<float*>(smraw + L.off_ctxp);
  float* stat = reinterpret_cast<float*>(smraw + L.off_stat);
  float* red = reinterpret_cast<float*>(smraw + L.off_red);
  float* ex = reinterpret_cast<float*>(smraw + L.off_extra);
  float* de = ex + X.de; float* al = ex + X.al; float* wls = ex + X.wls; float* dfo = ex + X.dfo;
  float* dfw = ex + X.dfw; float* dqp = ex + X.dqp; float* dzq = ex + X.dzq; float* tmp = ex + X.tmp;

  const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = NT >> 5;
  const int CS = p.CS;
  const unsigned r = CS > 1 ? cluster_ctarank() : 0u;
  const int b = blockIdx.x / CS;
  const int A = p.A, D = p.D, T = p.T, C = p.C;
  const int CP = (C + 3) / 4 * 4, CQ = CP / 4;
  const int KW1 = 2 * p.Kw + 1;
  const int len = min(T, max(0, p.lens[b]));
  const int chunk = (len + CS - 1) / CS;
  const int t0 = min(len, (int)r * chunk), t1 = min(len, t0 + chunk), nf = t1 - t0;
  const float* Kb = p.keys + (long long)b * p.k_bstride + (long long)t0 * A;
  const float* Hb = p.h + (long long)b * p.h_bstride + (long long)t0 * D;
  const int TFK = p.TFK, TFH = p.TFH, NS = p.NS;
  const int nK = (nf + TFK - 1) / TFK, nH = (nf + TFH - 1) / TFH, ntiles = nK + nH;
  const unsigned stage_floats = p.stage_bytes / 4;
  const bool loc = p.mode == 1;
  float* slot = p.pgrad ? p.pgrad + (long long)blockIdx.x * p.pgrad_stride : nullptr;   // [A*C | C*KW1 | A | 1]

  B2_TRACE(p, 16);
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  auto issue = [&](int i) {          // thread 0 only; encoder-feature tiles first, then key tiles
    if (i >= ntiles) return;
    const int s = i % NS;
    float* dst = ring + (size_t)s * stage_floats;
    const float* src; unsigned bytes;
    if (i < nH) { const int f0 = i * TFH; const int n = min(TFH, nf - f0); src = Hb + (long long)f0 * D; bytes = (unsigned)n * D * 4u; }
    else { const int f0 = (i - nH) * TFK; const int n = min(TFK, nf - f0); src = Kb + (long long)f0 * A; bytes = (unsigned)n * A * 4u; }
    mbar_expect_tx(&full[s], bytes);
    bulk_g2s(dst, src, bytes, &full[s]);
  };
  if (tid == 0) for (int i = 0; i < NS; ++i) issue(i);
  B2_TRACE(p, 17);

  // ---- stage small operands: asynchronous 4-byte copies, all in flight at once, one wait ------------------
  // What the FORWARD pass left (query, alignments) and the weights first; the upstream gradients, which the previous
  // kernels of the backward walk produce, after pdl_wait() (programmatic dependent launch, common.cuh).
  for (int a = tid; a < A; a += NT) cp_async4(q_sm + a, p.q_saved + (long long)b * A + a);
  for (int j = tid; j < nf; j += NT) cp_async4(al + j, p.att_saved + (long long)b * T + t0 + j);
  if (loc) {
    const int npa = nf + 2 * p.Kw;
    const float unif = len > 0 ? 1.f / (float)len : 0.f;
    for (int j = tid; j < npa; j += NT) {
      const int t = t0 - p.Kw + j;
      if (t >= 0 && t < T && p.prev_att) cp_async4(pa + j, p.prev_att + (long long)b * T + t);
      else pa[j] = (t >= 0 && t < len) ? unif : 0.f;
    }
    for (int i = tid; i < C * KW1; i += NT) { const int c = i / KW1, k = i - c * KW1; cp_async4(wcT + k * CP + c, p.Wc + i); }
    for (int i = tid; i < KW1 * (CP - C); i += NT) { const int k = i / (CP - C), c = C + i - k * (CP - C); wcT[k * CP + c] = 0.f; }
    for (int i = tid; i < A * C; i += NT) { const int a = i / C, c = i - a * C; cp_async4(wls + a * CP + c, p.Wl + i); }
    for (int i = tid; i < A * (CP - C); i += NT) { const int a = i / (CP - C), c = C + i - a * (CP - C); wls[a * CP + c] = 0.f; }
  }
  pdl_wait();
  pdl_launch();
  for (int d = tid; d < D; d += NT) cp_async4(dctx + d, p.d_ctx + (long long)b * p.dctx_stride + d);
  for (int j = tid; j < nf; j += NT) {
    if (p.d_att) cp_async4(de + j, p.d_att + (long long)b * T + t0 + j); else de[j] = 0.f;
  }
  cp_async_wait_all();
  cluster_or_cta_sync(CS);            // #0: operands staged; every CTA of the cluster is running (DSMEM is legal)

  // ---- recompute the location features of my frames ------------------------------------------------
  if (loc) {
    const float4* w4 = reinterpret_cast<const float4*>(wcT);
    for (int item = tid; item < nf * CQ; item += NT) {
      const int cg = item / nf, fl = item - cg * nf;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
      for (int k = 0; k < KW1; ++k) {
        const float pv = pa[fl + k];
        const float4 w = w4[k * CQ + cg];
        acc.x = fmaf(pv, w.x, acc.x); acc.y = fmaf(pv, w.y, acc.y); acc.z = fmaf(pv, w.z, acc.z); acc.w = fmaf(pv, w.w, acc.w);
      }
      reinterpret_cast<float4*>(fsm)[fl * CQ + cg] = acc;
    }
  }
  __syncthreads();
  B2_TRACE(p, 18);

  // ---- d_alpha over my frames: external part + h . d_ctx (one warp per frame) -------------------
  const int ngroups = D / 4;
  for (int i = 0; i < nH; ++i) {
    const int s = i % NS;
    mbar_wait(&full[s], (unsigned)((i / NS) & 1));
    const float4* ht = reinterpret_cast<const float4*>(ring + (size_t)s * stage_floats);
    const int f0 = i * TFH, n = min(TFH, nf - f0);
    for (int j = warp; j < n; j += nwarp) {
      float acc = 0.f;
      for (int g = lane; g < ngroups; g += 32) {
        const float4 hv = ht[(size_t)j * ngroups + g];
        const float4 dv = reinterpret_cast<const float4*>(dctx)[g];
        acc = fmaf(hv.x, dv.x, acc); acc = fmaf(hv.y, dv.y, acc); acc = fmaf(hv.z, dv.z, acc); acc = fmaf(hv.w, dv.w, acc);
      }
      acc = warp_sum(acc);
      if (lane == 0) de[f0 + j] += acc;              // de[] was staged with the external d_att (or 0)
    }
    __syncthreads();
    if (tid == 0) { fence_proxy_async_smem(); issue(i + NS); }
  }

  B2_TRACE(p, 19);
  // ---- softmax backward: de = scale * alpha * (d_alpha - <alpha, d_alpha>) ------------------------
  float part = 0.f;
  for (int j = tid; j < nf; j += NT) part = fmaf(al[j], de[j], part);
  part = cta_reduce_sum(part, red);
  if (tid == 0) {
    if (CS > 1) { for (int rr = 0; rr < CS; ++rr) st_dsmem_f32(dsmem_addr(stat + r, rr), part); }
    else stat[0] = part;
  }
  cluster_or_cta_sync(CS);            // #1
  float dotsum = 0.f;
  for (int rr = 0; rr < CS; ++rr) dotsum += stat[rr];
  float dbg_part = 0.f;
  for (int j = tid; j < nf; j += NT) { const float v = p.scale * al[j] * (de[j] - dotsum); de[j] = v; dbg_part += v; }
  dbg_part = cta_reduce_sum(dbg_part, red);     // (also orders the de[] writes before the reads below)
  if (loc && slot && tid == 0) red_add_f32(slot + (long long)A * C + (long long)C * KW1 + A, dbg_part);

  B2_TRACE(p, 20);
  // ---- main pass over the key tiles: thread a owns attention dimension a -------------------------
  // Two frames per iteration (two independent recompute chains).  d_zl overwrites the key value it was computed
  // from in the ring stage; after the tile, df[t][c] = sum_a d_zl[t][a] Wl[a][c] is formed by one warp per
  // (frame, 4 channels) with lanes striding a -- 6 shuffles per 4 outputs instead of 16 per frame and warp.
  const bool active = tid < A;
  float wl[CPT], dwl[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) { wl[c] = (active && loc && c < C) ? wls[tid * CP + (c < CP ? c : 0)] : 0.f; dwl[c] = 0.f; }
  const float qa = active ? q_sm[tid] : 0.f;
  const float wga = (active && loc) ? __ldg(p.wg + tid) : 0.f;
  float dq_acc = 0.f, dwg_acc = 0.f;
  const int kcol = active ? tid : 0;
  float* dkb = p.d_keys + (long long)b * p.dk_bstride + (long long)t0 * A;
  for (int i = 0; i < nK; ++i) {
    const int ti = nH + i, s = ti % NS;
    mbar_wait(&full[s], (unsigned)((ti / NS) & 1));
    float* kt = ring + (size_t)s * stage_floats;
    const int f0 = i * TFK, n = min(TFK, nf - f0);
    constexpr int UNR = MINB > 1 ? 2 : 4;
    // Branch-free over the UNR frames so their chains interleave: a frame past the tile end recomputes the last
    // frame with d_e = 0 (every contribution vanishes) and its stores are predicated off; inactive threads
    // (a >= A) carry wga = wl = 0, so their d_pre and d_zl are 0 and only their (unused) accumulators see garbage.
    if (loc) {
      for (int j = 0; j < n; j += UNR) {
        // All shared-memory loads of the UNR frames come first: d_zl is written back INTO the ring stage, and a load
        // placed after such a store cannot be moved above it (may alias), which used to chain the UNR frames one
        // after the other instead of interleaving them (measured: key pass 48 us for 100 frames per CTA).
        float kv[UNR], dev[UNR], fv[UNR][CPT];
        int jjs[UNR];
        bool oks[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          oks[u] = j + u < n;
          jjs[u] = oks[u] ? j + u : n - 1;
          kv[u] = kt[jjs[u] * A + kcol];
          dev[u] = oks[u] ? de[f0 + jjs[u]] : 0.f;
          const float4* f4 = reinterpret_cast<const float4*>(fsm) + (size_t)(f0 + jjs[u]) * CQ;
#pragma unroll
          for (int c4 = 0; c4 < CPT / 4; ++c4) {
            float4 t4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 < CQ) t4 = f4[c4];
            fv[u][4 * c4 + 0] = t4.x; fv[u][4 * c4 + 1] = t4.y; fv[u][4 * c4 + 2] = t4.z; fv[u][4 * c4 + 3] = t4.w;
          }
        }
        float dpre[UNR], dzl[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          float zl0 = 0.f, zl1 = 0.f;                     // two half-length dependency chains
#pragma unroll
          for (int c = 0; c < CPT; c += 2) { zl0 = fmaf(wl[c], fv[u][c], zl0); zl1 = fmaf(wl[c + 1], fv[u][c + 1], zl1); }
          const float l = tanh_fast(zl0 + zl1);
          const float uu = tanh_fast(kv[u] + qa + l);
          dwg_acc = fmaf(dev[u], uu, dwg_acc);
          dpre[u] = dev[u] * wga * (1.f - uu * uu);
          dq_acc += dpre[u];
          dzl[u] = dpre[u] * (1.f - l * l);
#pragma unroll
          for (int c = 0; c < CPT; ++c) dwl[c] = fmaf(dzl[u], fv[u][c], dwl[c]);
        }
        // fire-and-forget reduction (RED.ADD.F32): one owner per element and launch, so no contention, and no L2
        // round trip on the loop's critical path as a load-add-store would put there.  Unconditional and branch-free
        // (ptxas turns a predicated RED into a branch, which would split the unrolled frames into basic blocks
        // again): masked lanes add 0.0 to a valid element and park their d_zl in a scratch slot; a frame past the
        // tile end recomputed the last frame with d_e = 0, so every contribution of it vanishes.
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const bool live = active && oks[u];
          red_add_f32(dkb + (long long)(f0 + jjs[u]) * A + kcol, live ? dpre[u] : 0.f);
          float* zdst = live ? kt + jjs[u] * A + tid : red + (tid & 31);
          *zdst = dzl[u];
        }
      }
    } else {
      float* dkp = dkb + (long long)f0 * A + kcol;
      for (int j = 0; j < n; ++j, dkp += A) {
        const float kv = kt[j * A + kcol];
        const float dev = de[f0 + j];
        dq_acc = fmaf(dev, kv, dq_acc);
        if (active) red_add_f32(dkp, dev * qa);
      }
    }
    __syncthreads();
    if (loc) {
      // df[j][c] = sum_a d_zl[j][a] Wl[a][c]: one warp per frame, lanes stride a, all CPT channels at once (CPT/4
      // independent float4 accumulators and butterflies -> instruction-level parallelism instead of a task loop)
      const float4* wl4 = reinterpret_cast<const float4*>(wls);
      for (int j = warp; j < n; j += nwarp) {
        float4 acc[CPT / 4];
#pragma unroll
        for (int q = 0; q < CPT / 4; ++q) acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int a = lane; a < A; a += 32) {
          const float dz = kt[j * A + a];
#pragma unroll
          for (int q = 0; q < CPT / 4; ++q) {
            if (q < CQ) {
              const float4 w = wl4[a * CQ + q];
              acc[q].x = fmaf(dz, w.x, acc[q].x); acc[q].y = fmaf(dz, w.y, acc[q].y);
              acc[q].z = fmaf(dz, w.z, acc[q].z); acc[q].w = fmaf(dz, w.w, acc[q].w);
            }
          }
        }
        const bool up16 = lane & 16, up8 = lane & 8;
        const int sub = ((lane >> 4) & 1) * 2 + ((lane >> 3) & 1);
#pragma unroll
        for (int q = 0; q < CPT / 4; ++q) {
          if (q < CQ) {
            const float s0 = up16 ? acc[q].x : acc[q].z, s1 = up16 ? acc[q].y : acc[q].w;
            const float k0 = up16 ? acc[q].z : acc[q].x, k1 = up16 ? acc[q].w : acc[q].y;
            float x0 = k0 + __shfl_xor_sync(kFull, s0, 16);
            const float x1 = k1 + __shfl_xor_sync(kFull, s1, 16);
            const float t0 = up8 ? x0 : x1, t1 = up8 ? x1 : x0;
            x0 = t1 + __shfl_xor_sync(kFull, t0, 8);
            x0 += __shfl_xor_sync(kFull, x0, 4);
            x0 += __shfl_xor_sync(kFull, x0, 2);
            x0 += __shfl_xor_sync(kFull, x0, 1);
            if ((lane & 7) == 0) dfo[(size_t)(f0 + j) * CP + q * 4 + sub] = x0;
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0) { fence_proxy_async_smem(); issue(ti + NS); }
  }

  B2_TRACE(p, 21);
  // ---- parameter-gradient slot: dWl rows and dwg (private to this CTA; plain read-modify-write) ---
  if (loc && slot && active) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) if (c < C) red_add_f32(slot + (long long)tid * C + c, dwl[c]);
    red_add_f32(slot + (long long)A * C + (long long)C * KW1 + tid, dwg_acc);
  }
  if (active) dqp[tid] = dq_acc;
  cluster_or_cta_sync(CS);            // #2: dq partials and df of every CTA are complete
  B2_TRACE(p, 22);

  // ---- dq: slice-reduce over the cluster, through tanh', broadcast for the d_s product --------
  {
    const int aper = (A + CS - 1) / CS;
    const int a0 = (int)r * aper, a1 = min(A, a0 + aper);
    for (int a = a0 + tid; a < a1; a += NT) {
      float v = 0.f;
      if (CS > 1) { for (int rr = 0; rr < CS; ++rr) v += ld_dsmem_f32(dsmem_addr(dqp + a, rr)); }
      else v = dqp[a];
      if (p.Wphi) { const float qv = q_sm[a]; v *= (1.f - qv * qv); }
      if (p.d_zq) p.d_zq[(long long)b * A + a] = v;
      if (CS > 1) { for (int rr = 0; rr < CS; ++rr) st_dsmem_f32(dsmem_addr(dzq + a, rr), v); }
      else dzq[a] = v;
    }
  }

  B2_TRACE(p, 23);
  // ---- location branch: gather df halo from the neighbours, then dWc and d_prev_att -------------
  if (loc) {
    const int nw = nf + 2 * p.Kw;
#pragma unroll 4
    for (int item = tid; item < nw * CQ; item += NT) {       // one 128-bit (4-channel) load per item
      const int w = item / CQ, cg = item - w * CQ;
      const int tg = t0 - p.Kw + w;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (tg >= 0 && tg < len && chunk > 0) {
        const int rr = tg / chunk, li = tg - rr * chunk;
        const float* src = dfo + (size_t)li * CP + cg * 4;
        v = (CS > 1) ? ld_dsmem_f32x4(dsmem_addr(src, (unsigned)rr)) : *reinterpret_cast<const float4*>(src);
      }
      reinterpret_cast<float4*>(dfw)[item] = v;
    }
    __syncthreads();
    if (slot) {                     // dWc[c][k] += sum_j df[j][c] * prev_att[t0+j+k-Kw]; thread = (tap, 4 channels)
      float* sWc = slot + (long long)A * C;
      const float4* d4 = reinterpret_cast<const float4*>(dfo);
      for (int item = tid; item < KW1 * CQ; item += NT) {
        const int cg = item / KW1, k = item - cg * KW1;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int j = 0; j < nf; ++j) {
          const float pv = pa[j + k];
          const float4 d = d4[(size_t)j * CQ + cg];
          acc.x = fmaf(d.x, pv, acc.x); acc.y = fmaf(d.y, pv, acc.y); acc.z = fmaf(d.z, pv, acc.z); acc.w = fmaf(d.w, pv, acc.w);
        }
        const int c0 = cg * 4;
        if (c0 + 0 < C) red_add_f32(sWc + (c0 + 0) * KW1 + k, acc.x);
        if (c0 + 1 < C) red_add_f32(sWc + (c0 + 1) * KW1 + k, acc.y);
        if (c0 + 2 < C) red_add_f32(sWc + (c0 + 2) * KW1 + k, acc.z);
        if (c0 + 3 < C) red_add_f32(sWc + (c0 + 3) * KW1 + k, acc.w);
      }
    }
    if (p.d_prev_att) {             // d_prev_att[t0+i] = sum_k sum_c Wc[c][k] * df[t0+i-k+Kw][c]
      const float4* w4 = reinterpret_cast<const float4*>(wcT);
      const float4* d4 = reinterpret_cast<const float4*>(dfw);
      for (int item = tid; item < nf * CQ; item += NT) {
        const int cg = item / nf, i = item - cg * nf;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);       // four independent chains
#pragma unroll 4
        for (int k = 0; k < KW1; ++k) {
          const float4 w = w4[k * CQ + cg];
          const float4 d = d4[(size_t)(i - k + 2 * p.Kw) * CQ + cg];
          acc.x = fmaf(w.x, d.x, acc.x); acc.y = fmaf(w.y, d.y, acc.y); acc.z = fmaf(w.z, d.z, acc.z); acc.w = fmaf(w.w, d.w, acc.w);
        }
        tmp[cg * p.chunk_max + i] = (acc.x + acc.y) + (acc.z + acc.w);
      }
      __syncthreads();
      for (int i = tid; i < nf; i += NT) {
        float acc = 0.f;
        for (int cg = 0; cg < CQ; ++cg) acc += tmp[cg * p.chunk_max + i];
        p.d_prev_att[(long long)b * T + t0 + i] = acc;
      }
      for (int t = len + (int)r * NT + tid; t < T; t += CS * NT) p.d_prev_att[(long long)b * T + t] = 0.f;
    }
  }
  B2_TRACE(p, 24);
  cluster_or_cta_sync(CS);            // #3: d_zq broadcast complete (and all df halo reads done)
  B2_TRACE(p, 25);

  // ---- d_s = d_zq Wphi (slice of the decoder dimension per CTA) -----------------------------------
  if (p.d_s) {
    const int jper = (p.dec + CS - 1) / CS;
    const int j0 = (int)r * jper, j1 = min(p.dec, j0 + jper);
    if (p.Wphi) {
      // thread = (output j, quarter of the attention dimension): independent coalesced loads, batched by unrolling
      const int nj = j1 - j0;
      const int aq = (A + 3) / 4;
      float* part = tmp;                                  // [4][nj], sized in attn_bwd_extra
      for (int item = tid; item < nj * 4; item += NT) {
        const int qd = item / nj, j = j0 + item - qd * nj;
        const int a0 = qd * aq, a1 = min(A, a0 + aq);
        float acc = 0.f;
#pragma unroll 8
        for (int a = a0; a < a1; ++a) acc = fmaf(dzq[a], __ldg(p.Wphi + (long long)a * p.dec + j), acc);
        part[item] = acc;
      }
      __syncthreads();
      for (int j = tid; j < nj; j += NT) p.d_s[(long long)b * p.dec + j0 + j] = part[j] + part[nj + j] + part[2 * nj + j] + part[3 * nj + j];
    } else {
      for (int j = j0 + tid; j < j1; j += NT) p.d_s[(long long)b * p.dec + j] = dzq[j];
    }
  }
  B2_TRACE(p, 26);
  if (CS > 1) cluster_sync_all();     // #4
  B2_TRACE(p, 27);
}

int attn_configure(AttnArgs& a, int cluster_hint, size_t extra_floats, size_t* smem_out, size_t cap_bytes = 0);
int attn_launch(void (*kern)(AttnArgs), const AttnArgs& a, size_t smem, cudaStream_t st);

}  // namespace b2asr

using namespace b2asr;

extern "C" int64_t b2asr_attn_pgrad_stride(int A, int C, int Kw) {
  return (int64_t)A * C + (int64_t)C * (2 * Kw + 1) + A + 1;
}

extern "C" int b2asr_attn_slots(int B, int T, int D, int A, int dec, int mode, int C, int Kw, int cluster) {
  AttnArgs a = {};
  a.mode = mode; a.B = B; a.T = T; a.D = D; a.A = A; a.dec = dec; a.C = C; a.Kw = Kw;
  size_t smem = 0;
  if (attn_configure(a, cluster, 0, &smem) != B2ASR_OK) return B2ASR_EUNSUPPORTED;
  // the backward may settle on fewer ring stages, and on twice the clusters when two CTAs per SM fit (see
  // b2asr_attn_step_bwd): a few unused zero slots cost nothing
  const int cs2 = a.CS * 2;
  const bool may_widen = cluster <= 0 && a.NT <= 320 && cs2 <= kAttMaxCluster && B * cs2 > 148 && B * cs2 <= 2 * 148 &&
                         (T + cs2 * 2 - 1) / (cs2 * 2) >= 16;
  return B * (may_widen ? cs2 : a.CS);
}

extern "C" int b2asr_attn_step_bwd(int mode, const float* Wphi, const float* keys, int64_t k_bstride,
                                   const float* h, int64_t h_bstride, const int32_t* lens, const float* prev_att,
                                   const float* Wc, int C, int Kw, const float* Wl, const float* wg, float scale,
                                   const float* q_saved, const float* att_saved, const float* d_ctx, const float* d_att,
                                   float* d_keys, int64_t dk_bstride, float* d_prev_att, float* d_zq, float* d_s,
                                   float* pgrad, int64_t pgrad_stride,
                                   int B, int T, int D, int A, int dec, int cluster, void* stream) {
  const long long tl_stride = g_attn_row_stride;       // consumed before any early return (see attn_fwd.cu)
  g_attn_row_stride = 0;
  B2_REQUIRE(keys && h && lens && q_saved && att_saved && d_ctx && d_keys, B2ASR_EINVAL, "attn_step_bwd: null pointer argument");
  B2_REQUIRE(mode == 0 || mode == 1, B2ASR_EINVAL, "attn_step_bwd: mode %d (0 dot, 1 loc)", mode);
  B2_REQUIRE(B > 0 && T > 0 && D > 0 && A > 0 && dec > 0, B2ASR_EINVAL, "attn_step_bwd: bad sizes");
  B2_REQUIRE(mode == 0 || (Wc && Wl && wg), B2ASR_EINVAL, "attn_step_bwd: loc mode needs Wc, Wl, wg");
  B2_REQUIRE(mode == 0 || pgrad == nullptr || pgrad_stride >= b2asr_attn_pgrad_stride(A, C, Kw), B2ASR_EINVAL,
             "attn_step_bwd: pgrad_stride too small");
  B2_REQUIRE(Wphi || A == dec, B2ASR_EINVAL, "attn_step_bwd: proj=False needs A == dec");
  B2_REQUIRE((((uintptr_t)keys | (uintptr_t)h) & 15) == 0 && (k_bstride % 4 == 0) && (h_bstride % 4 == 0), B2ASR_EINVAL,
             "attn_step_bwd: keys/h must be 16-byte aligned");
  AttnArgs a = {};
  a.mode = mode; a.Wphi = Wphi; a.keys = keys; a.k_bstride = k_bstride; a.h = h; a.h_bstride = h_bstride;
  a.lens = lens; a.prev_att = prev_att; a.Wc = Wc; a.C = C; a.Kw = Kw; a.Wl = Wl; a.wg = wg; a.scale = scale;
  a.q_saved = q_saved; a.att_saved = att_saved; a.d_ctx = d_ctx; a.d_att = d_att;
  a.d_keys = d_keys; a.dk_bstride = dk_bstride; a.d_prev_att = d_prev_att; a.d_zq = d_zq; a.d_s = d_s;
  a.pgrad = pgrad; a.pgrad_stride = pgrad_stride;
  a.B = B; a.T = T; a.D = D; a.A = A; a.dec = dec;
  a.dctx_stride = tl_stride ? tl_stride : D;
  a.trace = attn_trace_buffer();
  // first pass fixes CS/NT/tiles, second pass sizes the ring with the backward's extra arrays
  size_t smem = 0;
  const AttnArgs a0 = a;
  int rc = attn_configure(a, cluster, 0, &smem);
  if (rc != B2ASR_OK) return rc;
  size_t cap = 227 * 1024;
  { const char* e = getenv("B2ASR_ATT_CAP_KB"); if (e && *e) cap = (size_t)atoi(e) * 1024; }
  // second pass: size the ring with the backward's extra arrays under a shared-memory cap
  auto size_ring = [&](AttnArgs& q, size_t limit) {
    if (q.TFK > 32) q.TFK = 32;
    q.NS = 4;
    AttnSmem R = attn_smem_layout(q, attn_bwd_extra(q).total);
    while (R.total > limit && q.NS > 2) { q.NS--; R = attn_smem_layout(q, attn_bwd_extra(q).total); }
    while (R.total > limit && q.TFK > 2 && q.TFH > 1) {     // long slices: shrink the ring tiles
      q.stage_bytes = (unsigned)align_up((size_t)q.stage_bytes / 2, 128);
      const unsigned rk = (unsigned)A * 4u, rh = (unsigned)D * 4u;
      q.TFK = (int)(q.stage_bytes / rk); if (q.TFK > 32) q.TFK = 32; if (q.TFK < 1) q.TFK = 1;
      q.TFH = (int)(q.stage_bytes / rh); if (q.TFH < 1) q.TFH = 1;
      const unsigned need = (rk * q.TFK > rh * q.TFH) ? rk * q.TFK : rh * q.TFH;
      if (q.stage_bytes < need) q.stage_bytes = (unsigned)align_up((size_t)need, 128);
      R = attn_smem_layout(q, attn_bwd_extra(q).total);
    }
    return R;
  };
  // Like the forward kernel (attn_fwd.cu): when the automatic cluster size stopped only at one CTA per SM, twice the
  // clusters with the ring capped at 113 KB (two CTAs per SM, the <= 102-register build) is faster inside the decoder
  // loop (M2: 16.0 -> 15.5 ms per batch), although the kernel timed alone is slower that way (73 -> 79 us).
  static const bool env_cfg = (getenv("B2ASR_ATT_CS") && *getenv("B2ASR_ATT_CS")) || (getenv("B2ASR_ATT_CAP_KB") && *getenv("B2ASR_ATT_CAP_KB"));
  bool wide = false;
  AttnSmem L = {};
  if (cluster <= 0 && !env_cfg && a.NT <= 320) {
    const int cs2 = a.CS * 2;
    if (cs2 <= kAttMaxCluster && a.B * cs2 > 148 && a.B * cs2 <= 2 * 148 && (a.T + cs2 * 2 - 1) / (cs2 * 2) >= 16) {
      AttnArgs a2 = a0;
      size_t smem2 = 0;
      if (attn_configure(a2, cs2, 0, &smem2, (size_t)113 * 1024) == B2ASR_OK) {
        const AttnSmem L2 = size_ring(a2, (size_t)113 * 1024);
        if (L2.total <= (size_t)113 * 1024) { a = a2; L = L2; wide = true; }
      }
    }
  }
  const int cs = a.CS;
  if (!wide) L = size_ring(a, cap);
  B2_REQUIRE(L.total <= 227 * 1024, B2ASR_EUNSUPPORTED, "attn_step_bwd: needs %zu bytes of shared memory (T=%d, cluster %d)", L.total, T, cs);
  const bool two = a.NT <= 320 && L.total <= 113 * 1024;      // two CTAs per SM fit: compile for <= 102 registers
  void (*kern)(AttnArgs);
  const bool small = a.NT <= 320;                              // one CTA per SM but <= 320 threads: up to 200 registers
  if (a.C <= 12) kern = two ? attn_step_bwd_kernel<12, 320, 2> : (small ? attn_step_bwd_kernel<12, 320, 1> : attn_step_bwd_kernel<12, 512, 1>);
  else kern = two ? attn_step_bwd_kernel<16, 320, 2> : (small ? attn_step_bwd_kernel<16, 320, 1> : attn_step_bwd_kernel<16, 512, 1>);
  return attn_launch(kern, a, L.total, (cudaStream_t)stream);
}
