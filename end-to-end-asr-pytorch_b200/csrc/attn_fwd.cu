// Kernel group (2b): one fused launch per decoder step of the reference's Attention.forward
// (src/asr.py:410-462): query projection tanh(phi s) (asr.py:422), location conv over the previous
// alignment + projection + tanh (asr.py:452), tanh energy against the cached keys (asr.py:453),
// padding mask + softmax(scale*e) (asr.py:454-455) and the context reduction (asr.py:457);
// `dot` mode (asr.py:426-432) is the same kernel without the location branch.
//
// Work split: a cluster of CS CTAs per utterance, each owning a contiguous slice of the valid
// frames.  Keys and encoder features of the slice stream through a shared-memory ring filled by
// 1-D bulk (TMA) copies that complete on mbarriers; the query, the softmax statistics and the
// partial contexts cross the cluster through distributed shared memory.
#include "attn.cuh"

namespace b2asr {

// CPT: location channels held in registers (12 or 16); MAXNT/MINB: launch bounds (two CTAs per SM when NT <= 320)
template <int CPT, int MAXNT, int MINB>
__global__ void __launch_bounds__(MAXNT, MINB) attn_step_fwd_kernel(AttnArgs p) {
  extern __shared__ __align__(128) unsigned char smraw[];
  const AttnSmem L = attn_smem_layout(p, 0);
  float* ring = reinterpret_cast<float*>(smraw + L.off_ring);
  unsigned long long* full = reinterpret_cast<unsigned long long*>(smraw + L.off_bar);
  float* q_sm = reinterpret_cast<float*>(smraw + L.off_q);
  float* s_sm = reinterpret_cast<float*>(smraw + L.off_s);
  float* wcT = reinterpret_cast<float*>(smraw + L.off_wc);
  float* pa = reinterpret_cast<float*>(smraw + L.off_pa);
  float* fsm = reinterpret_cast<float*>(smraw + L.off_f);
  float* e = reinterpret_cast<float*>(smraw + L.off_e);
  float* epart = reinterpret_cast<float*>(smraw + L.off_epart);
  float* ctxp = reinterpret_cast<float*>(smraw + L.off_ctxp);
  float* stat = reinterpret_cast<float*>(smraw + L.off_stat);
  float* red = reinterpret_cast<float*>(smraw + L.off_red);

  const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = NT >> 5;
  const int CS = p.CS;
  const unsigned r = CS > 1 ? cluster_ctarank() : 0u;
  const int b = blockIdx.x / CS;
  const int A = p.A, D = p.D, T = p.T, C = p.C;
  const int CP = (C + 3) / 4 * 4, CQ = CP / 4;
  const int len = min(T, max(0, p.lens[b]));
  const int chunk = (len + CS - 1) / CS;
  const int t0 = min(len, (int)r * chunk), t1 = min(len, t0 + chunk), nf = t1 - t0;
  const float* Kb = p.keys + (long long)b * p.k_bstride + (long long)t0 * A;
  const float* Hb = p.h + (long long)b * p.h_bstride + (long long)t0 * D;
  const int TFK = p.TFK, TFH = p.TFH, NS = p.NS;
  const int tfmax = TFK > TFH ? TFK : TFH;
  const int nK = (nf + TFK - 1) / TFK, nH = (nf + TFH - 1) / TFH, ntiles = nK + nH;
  const unsigned stage_floats = p.stage_bytes / 4;
  // NOTE: row_map is written by the kernel right before this one in the decode loop (the beam step); reading it here,
  // before pdl_wait(), is safe only because that kernel never triggers its dependents early (common.cuh).
  const long long bsrc = p.row_map ? p.row_map[b] : (long long)b;

  B2_TRACE(p, 0);
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  auto issue = [&](int i) {          // thread 0 only: key tiles first, then encoder-feature tiles
    if (i >= ntiles) return;
    const int s = i % NS;
    float* dst = ring + (size_t)s * stage_floats;
    const float* src; unsigned bytes;
    if (i < nK) { const int f0 = i * TFK; const int n = min(TFK, nf - f0); src = Kb + (long long)f0 * A; bytes = (unsigned)n * A * 4u; }
    else { const int f0 = (i - nK) * TFH; const int n = min(TFH, nf - f0); src = Hb + (long long)f0 * D; bytes = (unsigned)n * D * 4u; }
    mbar_expect_tx(&full[s], bytes);
    bulk_g2s(dst, src, bytes, &full[s]);
  };
  if (tid == 0) for (int i = 0; i < NS; ++i) issue(i);
  B2_TRACE(p, 1);

  // ---- stage the small operands: asynchronous 4-byte copies, all in flight at once, one wait ---------------
  // Weights first: they (like the key / feature tiles requested above) do not depend on the previous kernel of the
  // decoder loop, so with a programmatic dependent launch all of this overlaps the previous kernel's tail.
  const int KW1 = 2 * p.Kw + 1;
  if (p.mode == 1) {
    for (int i = tid; i < C * KW1; i += NT) {            // transposed [tap][channel] so one float4 feeds 4 channels
      const int c = i / KW1, k = i - c * KW1;
      cp_async4(wcT + k * CP + c, p.Wc + i);
    }
    for (int i = tid; i < KW1 * (CP - C); i += NT) { const int k = i / (CP - C), c = C + i - k * (CP - C); wcT[k * CP + c] = 0.f; }
  }
  // The previous alignment was written by the PREVIOUS attention kernel, which had completed before the decoder
  // cell between the two (the only kernel that may still be running) could start: safe before the wait.
  if (p.mode == 1) {
    const int npa = nf + 2 * p.Kw;
    const float unif = len > 0 ? 1.f / (float)len : 0.f;
    for (int j = tid; j < npa; j += NT) {
      const int t = t0 - p.Kw + j;
      if (t >= 0 && t < T && p.prev_att) cp_async4(pa + j, p.prev_att + bsrc * T + t);
      else pa[j] = (t >= 0 && t < len) ? unif : 0.f;
    }
    cp_async_wait_all();
    __syncthreads();
    // ---- location features f[t][c] = sum_k Wc[c][k] * prev_att[t+k-Kw] for my frames ---------------
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
  pdl_wait();                           // the decoder state comes from the cell kernel before this one
  pdl_launch();
  for (int i = tid; i < p.dec; i += NT) cp_async4(s_sm + i, p.s + bsrc * p.s_stride + i);
  cp_async_wait_all();
  cluster_or_cta_sync(CS);            // #0: operands staged; every CTA of the cluster is running (DSMEM is legal)
  B2_TRACE(p, 2);

  // ---- query: q = tanh(phi s), each CTA computes a slice and scatters it through DSMEM ----------
  if (p.Wphi) {
    const int aper = (A + CS - 1) / CS;
    const int a0 = (int)r * aper, a1 = min(A, a0 + aper);
    const bool v32 = (p.dec % 32 == 0) && (((uintptr_t)p.Wphi & 15) == 0);
    if (v32) {
      // eight lanes per row, an eighth of the row each, all of its 128-bit loads in flight at once
      const int q8 = p.dec / 32;                          // float4 per lane
      for (int item0 = 0; item0 < (a1 - a0) * 8; item0 += NT) {
        const int item = item0 + tid;
        const int a = a0 + (item >> 3), qd = item & 7;
        float acc = 0.f;
        if (a < a1) {
          const float4* w = reinterpret_cast<const float4*>(p.Wphi + (long long)a * p.dec) + qd * q8;
          const float4* sv4 = reinterpret_cast<const float4*>(s_sm) + qd * q8;
          for (int i0 = 0; i0 < q8; i0 += 5) {             // five 128-bit loads in flight per lane
            float4 wv[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) wv[i] = (i0 + i < q8) ? __ldg(w + i0 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 5; ++i) {
              if (i0 + i < q8) {
                const float4 sv = sv4[i0 + i];
                acc = fmaf(wv[i].x, sv.x, acc); acc = fmaf(wv[i].y, sv.y, acc); acc = fmaf(wv[i].z, sv.z, acc); acc = fmaf(wv[i].w, sv.w, acc);
              }
            }
          }
        }
        acc += __shfl_xor_sync(kFull, acc, 1);
        acc += __shfl_xor_sync(kFull, acc, 2);
        acc += __shfl_xor_sync(kFull, acc, 4);
        if (a < a1 && qd == 0) {
          const float qv = tanh_acc(acc);
          if (CS > 1) { for (int rr = 0; rr < CS; ++rr) st_dsmem_f32(dsmem_addr(q_sm + a, rr), qv); }
          else q_sm[a] = qv;
          if (p.q_out) p.q_out[(long long)b * A + a] = qv;
        }
      }
    } else {
      for (int a = a0 + warp; a < a1; a += nwarp) {
        const float* w = p.Wphi + (long long)a * p.dec;
        float acc = 0.f;
        for (int i = lane; i < p.dec; i += 32) acc = fmaf(__ldg(w + i), s_sm[i], acc);
        acc = warp_sum(acc);
        if (lane == 0) {
          const float qv = tanh_acc(acc);
          if (CS > 1) { for (int rr = 0; rr < CS; ++rr) st_dsmem_f32(dsmem_addr(q_sm + a, rr), qv); }
          else q_sm[a] = qv;
          if (p.q_out) p.q_out[(long long)b * A + a] = qv;
        }
      }
    }
  } else {
    for (int a = tid; a < A; a += NT) q_sm[a] = s_sm[a];
  }

  B2_TRACE(p, 3);
  cluster_or_cta_sync(CS);            // #1: q and f visible
  B2_TRACE(p, 4);

  const bool active = tid < A;
  float wl[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) wl[c] = (active && p.mode == 1 && c < C) ? __ldg(p.Wl + (long long)tid * C + c) : 0.f;
  const float qa = active ? q_sm[tid] : 0.f;
  const float wga = (active && p.mode == 1) ? __ldg(p.wg + tid) : 0.f;
  const float bgv = p.mode == 1 ? __ldg(p.bg) : 0.f;
  const int kcol = active ? tid : 0;
  // ---- energies: thread a owns attention dimension a, frames stream through the ring --------
  // Four frames per iteration: four independent tanh chains per thread, and the four per-frame sums over the
  // warp's 32 dimensions cost 6 shuffles together (transposing butterfly) instead of 5 each.
  for (int i = 0; i < nK; ++i) {
    const int s = i % NS;
    mbar_wait(&full[s], (unsigned)((i / NS) & 1));
    const float* kt = ring + (size_t)s * stage_floats;
    const int f0 = i * TFK, n = min(TFK, nf - f0);
    for (int j = 0; j < n; j += 4) {
      float part[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int jj = min(j + u, n - 1);              // tail: recompute the last frame, the result is dropped
        const float kv = kt[jj * A + kcol];
        if (p.mode == 1) {
          const float4* f4 = reinterpret_cast<const float4*>(fsm) + (size_t)(f0 + jj) * CQ;
          float zl0 = 0.f, zl1 = 0.f;                    // two half-length dependency chains
#pragma unroll
          for (int c4 = 0; c4 < CPT / 4; ++c4) {
            if (c4 < CQ) {
              const float4 fv = f4[c4];
              zl0 = fmaf(wl[4 * c4 + 0], fv.x, zl0); zl1 = fmaf(wl[4 * c4 + 1], fv.y, zl1);
              zl0 = fmaf(wl[4 * c4 + 2], fv.z, zl0); zl1 = fmaf(wl[4 * c4 + 3], fv.w, zl1);
            }
          }
          part[u] = wga * tanh_fast(kv + qa + tanh_fast(zl0 + zl1));
        } else {
          part[u] = active ? kv * qa : 0.f;
        }
      }
      {
        const bool up = lane & 16;
        const float s0 = up ? part[0] : part[2], s1 = up ? part[1] : part[3];
        const float k0 = up ? part[2] : part[0], k1 = up ? part[3] : part[1];
        part[0] = k0 + __shfl_xor_sync(kFull, s0, 16);
        part[1] = k1 + __shfl_xor_sync(kFull, s1, 16);
      }
      {
        const bool up = lane & 8;
        const float s0 = up ? part[0] : part[1], k0 = up ? part[1] : part[0];
        part[0] = k0 + __shfl_xor_sync(kFull, s0, 8);
      }
      part[0] += __shfl_xor_sync(kFull, part[0], 4);
      part[0] += __shfl_xor_sync(kFull, part[0], 2);
      part[0] += __shfl_xor_sync(kFull, part[0], 1);
      if ((lane & 7) == 0) {
        const int u = ((lane >> 4) & 1) * 2 + ((lane >> 3) & 1);
        if (j + u < n) epart[warp * tfmax + j + u] = part[0];
      }
    }
    __syncthreads();
    for (int j = tid; j < n; j += NT) {
      float acc = 0.f;
      for (int w = 0; w < nwarp; ++w) acc += epart[w * tfmax + j];
      e[f0 + j] = (acc + bgv) * p.scale;
    }
    __syncthreads();                  // epart reusable; ring stage s fully consumed
    if (tid == 0) { fence_proxy_async_smem(); issue(i + NS); }
  }

  B2_TRACE(p, 5);
  // ---- masked softmax over the utterance: local statistics, then one exchange -----------------
  float m = -INFINITY;
  for (int j = tid; j < nf; j += NT) m = fmaxf(m, e[j]);
  m = cta_reduce_max(m, red);
  const float mm = (m == -INFINITY) ? 0.f : m;
  float ssum = 0.f;
  for (int j = tid; j < nf; j += NT) ssum += ex2((e[j] - mm) * kLog2e);
  ssum = cta_reduce_sum(ssum, red);
  if (tid == 0) {
    if (CS > 1) {
      for (int rr = 0; rr < CS; ++rr) { st_dsmem_f32(dsmem_addr(stat + 2 * r, rr), m); st_dsmem_f32(dsmem_addr(stat + 2 * r + 1, rr), ssum); }
    } else { stat[0] = m; stat[1] = ssum; }
  }
  cluster_or_cta_sync(CS);            // #2
  float M = -INFINITY;
  for (int rr = 0; rr < CS; ++rr) M = fmaxf(M, stat[2 * rr]);
  float Z = 0.f;
  for (int rr = 0; rr < CS; ++rr) if (stat[2 * rr] != -INFINITY) Z += stat[2 * rr + 1] * ex2((stat[2 * rr] - M) * kLog2e);
  const float inv = Z > 0.f ? 1.f / Z : 0.f;
  for (int j = tid; j < nf; j += NT) {
    const float al = ex2((e[j] - M) * kLog2e) * inv;
    e[j] = al;
    p.att[(long long)b * T + t0 + j] = al;
  }
  for (int t = len + (int)r * NT + tid; t < T; t += CS * NT) p.att[(long long)b * T + t] = 0.f;
  __syncthreads();

  B2_TRACE(p, 6);
  // ---- context: ctx = sum_t alpha_t h_t over my frames, then a DSMEM slice reduction ----------
  const int ngroups = D / 4, items = ngroups * p.FS;
  float4 acc[kAttMaxIPT];
#pragma unroll
  for (int it = 0; it < kAttMaxIPT; ++it) acc[it] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = 0; i < nH; ++i) {
    const int ti = nK + i, s = ti % NS;
    mbar_wait(&full[s], (unsigned)((ti / NS) & 1));
    const float4* ht = reinterpret_cast<const float4*>(ring + (size_t)s * stage_floats);
    const int f0 = i * TFH, n = min(TFH, nf - f0);
#pragma unroll
    for (int it = 0; it < kAttMaxIPT; ++it) {
      const int item = tid + it * NT;
      if (it < p.IPT && item < items) {
        const int g = item % ngroups, fs = item / ngroups;
        for (int j = fs; j < n; j += p.FS) {
          const float al = e[f0 + j];
          const float4 hv = ht[(size_t)j * ngroups + g];
          acc[it].x = fmaf(al, hv.x, acc[it].x); acc[it].y = fmaf(al, hv.y, acc[it].y);
          acc[it].z = fmaf(al, hv.z, acc[it].z); acc[it].w = fmaf(al, hv.w, acc[it].w);
        }
      }
    }
    __syncthreads();
    if (tid == 0) { fence_proxy_async_smem(); issue(ti + NS); }
  }
#pragma unroll
  for (int it = 0; it < kAttMaxIPT; ++it) {
    const int item = tid + it * NT;
    if (it < p.IPT && item < items) reinterpret_cast<float4*>(ctxp)[item] = acc[it];   // [fs][D]
  }
  __syncthreads();
  for (int d = tid; d < D; d += NT) {
    float v = 0.f;
    for (int fs = 0; fs < p.FS; ++fs) v += ctxp[(size_t)fs * D + d];
    ctxp[d] = v;
  }
  B2_TRACE(p, 7);
  cluster_or_cta_sync(CS);            // #3: partial contexts visible cluster-wide
  B2_TRACE(p, 8);
  {
    const int dper = (D + CS - 1) / CS;
    const int d0 = (int)r * dper, d1 = min(D, d0 + dper);
    for (int d = d0 + tid; d < d1; d += NT) {
      float v = 0.f;
      if (CS > 1) { for (int rr = 0; rr < CS; ++rr) v += ld_dsmem_f32(dsmem_addr(ctxp + d, rr)); }
      else v = ctxp[d];
      p.ctx[(long long)b * p.ctx_stride + d] = v;
    }
  }
  if (CS > 1) cluster_sync_all();     // #4: nobody leaves while its shared memory is still being read
  B2_TRACE(p, 9);
}


// ---------------------------------------------------------------------------------------------
int attn_configure(AttnArgs& a, int cluster_hint, size_t extra_floats, size_t* smem_out, size_t cap_bytes = 0) {
  B2_REQUIRE(a.A % 4 == 0 && a.D % 4 == 0, B2ASR_EUNSUPPORTED, "attn: A=%d and D=%d must be multiples of 4", a.A, a.D);
  B2_REQUIRE(a.A <= 512, B2ASR_EUNSUPPORTED, "attn: attention dim %d > 512 unsupported", a.A);
  B2_REQUIRE(a.mode == 0 || (a.C >= 1 && a.C <= kAttMaxC && a.Kw >= 0 && a.Kw <= 256), B2ASR_EUNSUPPORTED,
             "attn: location conv C=%d Kw=%d unsupported", a.C, a.Kw);
  if (a.mode == 0) { a.C = 4; a.Kw = 0; }
  int CS = cluster_hint;
  { const char* e = getenv("B2ASR_ATT_CS"); if (CS <= 0 && e && *e) CS = atoi(e); }
  if (CS <= 0) {
    CS = 1;
    while (CS < kAttMaxCluster && a.B * CS * 2 <= 148 && (a.T + CS * 2 - 1) / (CS * 2) >= 16) CS *= 2;
  }
  B2_REQUIRE(CS == 1 || CS == 2 || CS == 4 || CS == 8, B2ASR_EINVAL, "attn: cluster size %d not in {1,2,4,8}", CS);
  a.CS = CS;
  a.NT = (int)align_up((size_t)(a.A > 128 ? a.A : 128), 32);
  a.chunk_max = (a.T + CS - 1) / CS;
  a.stage_bytes = 40960u;
  { const char* e = getenv("B2ASR_ATT_STAGE"); if (e && *e) a.stage_bytes = (unsigned)atoi(e); }
  const unsigned rowK = (unsigned)a.A * 4u, rowH = (unsigned)a.D * 4u;
  const unsigned rowmax = rowK > rowH ? rowK : rowH;
  if (rowmax * 4u > a.stage_bytes) a.stage_bytes = (unsigned)align_up((size_t)rowmax * 4u, 128);
  a.TFK = (int)(a.stage_bytes / rowK); if (a.TFK > 64) a.TFK = 64;
  a.TFH = (int)(a.stage_bytes / rowH); if (a.TFH > 64) a.TFH = 64;
  a.FS = a.NT / (a.D / 4); if (a.FS < 1) a.FS = 1; if (a.FS > 4) a.FS = 4;
  a.IPT = ((a.D / 4) * a.FS + a.NT - 1) / a.NT;
  B2_REQUIRE(a.IPT <= kAttMaxIPT, B2ASR_EUNSUPPORTED, "attn: encoder dim %d too wide", a.D);
  a.NS = 4;
  { const char* e = getenv("B2ASR_ATT_NS"); if (e && *e) a.NS = atoi(e); }
  AttnSmem L = attn_smem_layout(a, extra_floats);
  size_t cap = 227 * 1024;
  { const char* e = getenv("B2ASR_ATT_CAP_KB"); if (e && *e) cap = (size_t)atoi(e) * 1024; }
  if (cap_bytes) cap = cap_bytes;
  while (L.total > cap && a.NS > 2) { a.NS--; L = attn_smem_layout(a, extra_floats); }
  B2_REQUIRE(L.total <= 227 * 1024, B2ASR_EUNSUPPORTED, "attn: needs %zu bytes of shared memory (T=%d too long for cluster %d)", L.total, a.T, CS);
  *smem_out = L.total;
  return B2ASR_OK;
}

__device__ long long g_attn_trace[32];
long long* attn_trace_buffer() {
  static long long* buf = nullptr;
  static int init = 0;
  if (!init) {
    init = 1;
    const char* e = getenv("B2ASR_ATT_TRACE");
    if (e && *e && *e != '0') { void* q = nullptr; if (cudaGetSymbolAddress(&q, g_attn_trace) == cudaSuccess) buf = (long long*)q; }
  }
  return buf;
}

// B2ASR_PDL=0 launches the decoder loop's kernels without programmatic dependent launch (A/B runs)
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("B2ASR_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

int attn_launch(void (*kern)(AttnArgs), const AttnArgs& a, size_t smem, cudaStream_t st) {
  B2_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(a.B * a.CS));
  cfg.blockDim = dim3((unsigned)a.NT);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)a.CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // see pdl_wait() in common.cuh
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 2 : 1;
  B2_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a));
  return B2ASR_OK;
}

thread_local long long g_attn_row_stride = 0;

}  // namespace b2asr

using namespace b2asr;

extern "C" int b2asr_attn_step_fwd(int mode, const float* s, int64_t s_stride, const float* Wphi,
                                   const float* keys, int64_t k_bstride, const float* h, int64_t h_bstride,
                                   const int32_t* lens, const float* prev_att,
                                   const float* Wc, int C, int Kw, const float* Wl, const float* wg, const float* bg,
                                   float scale, float* att, float* ctx, float* q_out,
                                   int B, int T, int D, int A, int dec, int cluster, void* stream) {
  return b2asr_attn_step_fwd_rows(mode, s, s_stride, Wphi, keys, k_bstride, h, h_bstride, lens, prev_att, nullptr, Wc, C,
                                  Kw, Wl, wg, bg, scale, att, ctx, (int64_t)0, q_out, B, T, D, A, dec, cluster, stream);
}

extern "C" int b2asr_attn_step_fwd_rows(int mode, const float* s, int64_t s_stride, const float* Wphi,
                                        const float* keys, int64_t k_bstride, const float* h, int64_t h_bstride,
                                        const int32_t* lens, const float* prev_att, const int64_t* row_map,
                                        const float* Wc, int C, int Kw, const float* Wl, const float* wg, const float* bg,
                                        float scale, float* att, float* ctx, int64_t ctx_stride, float* q_out,
                                        int B, int T, int D, int A, int dec, int cluster, void* stream) {
  const long long tl_stride = g_attn_row_stride;       // consumed before any early return: a failed
  g_attn_row_stride = 0;                               // check must not leave the stride behind for the next call
  B2_REQUIRE(s && keys && h && lens && att && ctx, B2ASR_EINVAL, "attn_step_fwd: null pointer argument");
  B2_REQUIRE(mode == 0 || mode == 1, B2ASR_EINVAL, "attn_step_fwd: mode %d (0 dot, 1 loc)", mode);
  B2_REQUIRE(B > 0 && T > 0 && D > 0 && A > 0 && dec > 0, B2ASR_EINVAL, "attn_step_fwd: bad sizes");
  B2_REQUIRE(mode == 0 || (Wc && Wl && wg && bg), B2ASR_EINVAL, "attn_step_fwd: loc mode needs Wc, Wl, wg, bg");
  B2_REQUIRE(Wphi || A == dec, B2ASR_EINVAL, "attn_step_fwd: proj=False needs A == dec");
  B2_REQUIRE((((uintptr_t)keys | (uintptr_t)h) & 15) == 0 && (k_bstride % 4 == 0) && (h_bstride % 4 == 0), B2ASR_EINVAL,
             "attn_step_fwd: keys/h must be 16-byte aligned");
  AttnArgs a = {};
  a.mode = mode; a.s = s; a.s_stride = s_stride; a.Wphi = Wphi; a.keys = keys; a.k_bstride = k_bstride;
  a.h = h; a.h_bstride = h_bstride; a.lens = lens; a.prev_att = prev_att; a.Wc = Wc; a.C = C; a.Kw = Kw;
  a.Wl = Wl; a.wg = wg; a.bg = bg; a.scale = scale; a.att = att; a.ctx = ctx; a.q_out = q_out;
  a.B = B; a.T = T; a.D = D; a.A = A; a.dec = dec;
  a.row_map = (const long long*)row_map;
  a.ctx_stride = ctx_stride > 0 ? ctx_stride : (tl_stride ? tl_stride : D);
  B2_REQUIRE(a.ctx_stride >= D, B2ASR_EINVAL, "attn_step_fwd: ctx_stride %lld smaller than D=%d", (long long)a.ctx_stride, D);
  a.trace = attn_trace_buffer();
  size_t smem = 0;
  static const int fwd_cs = [] { const char* e = getenv("B2ASR_ATT_FWD_CS"); return e && *e ? atoi(e) : 0; }();
  static const int fwd_cap = [] { const char* e = getenv("B2ASR_ATT_FWD_CAP_KB"); return e && *e ? atoi(e) : 0; }();
  const AttnArgs a0 = a;
  int rc = attn_configure(a, cluster > 0 ? cluster : fwd_cs, 0, &smem, (size_t)fwd_cap * 1024);
  if (rc != B2ASR_OK) return rc;
  // The automatic cluster size stops doubling when the grid would exceed one CTA per SM.  The forward kernel is
  // compiled for two CTAs per SM, and twice the clusters at half the frames each -- ring capped at 113 KB so that two
  // CTAs share an SM -- is faster when the slices stay long enough (M2: 36.4 -> 34.3 us alone, 43.3 -> 39.3 us per
  // step inside the decoder loop).  The backward kernel keeps one CTA per SM (A/B in DESIGN.md section 5).
  if (cluster <= 0 && fwd_cs == 0 && a.NT <= 320) {
    const int cs2 = a.CS * 2;
    if (cs2 <= kAttMaxCluster && a.B * cs2 > 148 && a.B * cs2 <= 2 * 148 && (a.T + cs2 * 2 - 1) / (cs2 * 2) >= 16) {
      AttnArgs a2 = a0;
      size_t smem2 = 0;
      if (attn_configure(a2, cs2, 0, &smem2, (size_t)113 * 1024) == B2ASR_OK && smem2 <= (size_t)113 * 1024) {
        a = a2;
        smem = smem2;
      }
    }
  }
  void (*kern)(AttnArgs);
  const bool small = a.NT <= 320;
  if (a.C <= 12) kern = small ? attn_step_fwd_kernel<12, 320, 2> : attn_step_fwd_kernel<12, 512, 1>;
  else kern = small ? attn_step_fwd_kernel<16, 320, 2> : attn_step_fwd_kernel<16, 512, 1>;
  return attn_launch(kern, a, smem, (cudaStream_t)stream);
}

// debug aid: copy the phase stamps of the last traced launch (B2ASR_ATT_TRACE=1) to the host
extern "C" int b2asr_debug_attn_trace(long long* host_out, int n) {
  long long* buf = attn_trace_buffer();
  B2_REQUIRE(buf && host_out && n > 0 && n <= 32, B2ASR_EINVAL, "attn trace: not enabled (set B2ASR_ATT_TRACE=1) or bad arguments");
  B2_CUDA_TRY(cudaMemcpy(host_out, buf, sizeof(long long) * n, cudaMemcpyDeviceToHost));
  return B2ASR_OK;
}
