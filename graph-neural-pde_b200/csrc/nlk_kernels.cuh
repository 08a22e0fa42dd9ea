// nlk_kernels.cuh - third generation of the fused GRAND-nl row kernels (ODEFuncTransformerAtt.forward,
// src/function_transformer_attention.py:44-59, scaled_dot scores, attention_norm_idx 0, and the row pass of its VJP).
//
// Why: ncu of the second generation (profiles/r02h_ncu_full_nl2_c4.csv) shows the single-gather kernels bound by the
// shared-memory data pipe (l1tex 82 - 86 %, 3.55 G wavefronts per adjoint launch on the ogbn-products shape; issue
// 52 %, DRAM 31 %), and about half of those wavefronts - and ~45 % of the 8.8 G warp instructions - are PER-ROW work:
// forming p_ih = Wk_h^T q_ih and dq_ih = Wk_h u_ih delivers all of Wk (32 KB) to the registers of the warp twice per
// row, 16 k MACs per row against 43 k for its 26 edges.  The single gather saves 256 B of DRAM traffic per edge
// (k[col] is never fetched) and pays ~1250 LSU cycles + ~1100 instructions per row for it; at 44 GB/s of HBM per SM
// the 256 B x 26 edges cost ~290 cycles.  So for d > A these kernels gather BOTH rows per edge,
//
//     x[col] (4 d bytes)  and  k[col] = Wk x[col] + bk (4 A bytes; projected together with q by ONE tensor-core GEMM),
//
// and everything per row disappears:
//   score      s_eh = <q_ih, k_jh> / sqrt(dk): dk MACs per (edge, head) on a lane of its own - no cross-lane
//              reduction, no p, no Wk in shared memory
//   dq         dq_ih = sum_e ds_eh k_jh / sqrt(dk) = (K2_h / (H l_h) - dot_h K1_h / l_h) / sqrt(dk) with the online sums
//              K1_h = sum_e e^{s-m} k_jh, K2_h = sum_e e^{s-m} dabar_e k_jh: A floats each (one float4 on A/4 lanes)
//              instead of the d-wide T2_h accumulators and the Wk u product
//   (the bias term <q_ih, bk_h> is constant over the edges of a row and cancels in the row softmax, as before;
//   sum_e ds_eh = 0, so the bias does not reach dq either)
//
// Staging: as in nl2_kernels.cuh - every warp owns a ring of kNlkStages stages of EB edges; one `cp.async.bulk`
// (UBLKCP, TMA unit) per row copies x[col] and k[col] into it, completion on an mbarrier; the warp issues a batch
// ahead of its arithmetic, across row boundaries inside a chunk of rows.
//   phase 0 (adjoint)  lanes span the columns: the x chunks of the batch go to registers ONCE (they are reused in
//                      phase 3), dabar_e = alpha <g_i, x_j> by a transposing butterfly over the warp
//   phase 1            lane = (head h, edge e): the score from q_ih (shared memory, broadcast) and k_jh
//   phase 2            online softmax per (head, edge slot); exp(s - m) [, exp(s - m) dabar] -> a record per edge
//   phase 3            lanes span the columns: T1_h += w_eh x_j (2 FFMA2 per head);  adjoint: lanes < A/4 also
//                      K1 += w k_j, K2 += w dabar k_j for their float4 of k
// Per-row work: stage q (A floats), epilogue, dq (one float4 store on A/4 lanes), ds sweep (unchanged).
// Hub rows (> long_threshold edges) are worked CTA-wide inside the same persistent launch.  No float atomics; every sum
// has a fixed order: bit-reproducible run to run.
//
// Roofline: HBM.  Algorithmic bytes per launch (perfect reuse of gathered rows; q, k are inputs here):
//   forward  4E + 4N + 8NA (q, k) + 12Nd (x, x0, out)
//   adjoint  4E + 4N + 12NA (q, k, dq) + 8Nd (x, g) [+ 8Nd when f is emitted] + 20E (ds [E,H] + abar [E] out)
// Gather model (no reuse): + 4 (E - N) (d + A).
#pragma once
#include "nl2_kernels.cuh"

namespace gb {

#ifndef GB_NLK_STAGES
#define GB_NLK_STAGES 2
#endif
constexpr int kNlkStages = GB_NLK_STAGES;
constexpr int kNlkXRow = 16 * 32 + 16;             // bytes between staged x rows (d <= 128; +16: bank spread)

// shared-memory geometry (bytes).  CTA: [W warp regions] [statistics of the CTA-wide rows [W][3][H]]
// [accumulators of the CTA-wide rows [H + 2][32] float4].  Warp region: [ring: NS x (EB x rows | EB k rows)]
// [q] [weight records [EB][2H or H]] [mbarriers]
struct NlkGeom {
  int ks;                    // bytes between staged k rows: 4 A + 16
  int stage_bytes;           // EB (kNlkXRow + ks)
  int off_q, off_w, off_bar;
  int warp_bytes;
  int pst_off, pacc_off;
  int total;
};
template <int H, bool ADJ, int W>
__host__ __device__ inline NlkGeom nlk_geom(int adim) {
  constexpr int EB = (32 / H) < 8 ? (32 / H) : 8;
  constexpr int WRECF = ((ADJ ? 2 * H : H) + 3) / 4 * 4;
  NlkGeom g;
  g.ks = 4 * adim + 16;
  g.stage_bytes = EB * (kNlkXRow + g.ks);
  g.off_q = kNlkStages * g.stage_bytes;
  g.off_w = g.off_q + ((4 * adim + 15) / 16) * 16;
  g.off_bar = g.off_w + EB * WRECF * 4;
  g.warp_bytes = ((g.off_bar + 8 * kNlkStages + 127) / 128) * 128;
  g.pst_off = W * g.warp_bytes;
  g.pacc_off = g.pst_off + ((W * 3 * H * 4 + 15) / 16) * 16;
  g.total = g.pacc_off + (H + 2) * 32 * 16;
  return g;
}

// Register / shared-memory state of one warp.
template <int H, bool ADJ>
struct NlkWarp {
  static constexpr int EB = (32 / H) < 8 ? (32 / H) : 8;    // edges per batch (H * EB <= 32)
  static constexpr int NS = kNlkStages;
  static constexpr int WREC = ADJ ? 2 * H : H;              // floats per edge in the weight record
  static constexpr int WRECF = (WREC + 3) / 4 * 4;          // adjoint: (w_h, w_h dabar) interleaved; forward: w_h
  static constexpr int SHD = 5 - ilog2_c(EB);               // dabar of edge e ends on lanes [e << SHD, (e + 1) << SHD)
  static_assert(H == 1 || H == 2 || H == 4 || H == 8, "unsupported head count");

  int h1, e1;                // phase-1/2 identity of the lane (lanes < H * EB): head, edge slot
  bool act1;
  int kl, hk;                // phase-3 k identity: lanes < A/4 own float4 `kl` of k, which belongs to head hk
  bool actk;
  float m, l, dsum;          // online softmax state of (head h1, edge slot e1)
  float4 acc1[H];            // T1_h, this lane's 4 columns
  float4 k1, k2;             // K1 / K2, this lane's float4 of k (adjoint)
  float4 gi;                 // alpha * g_i, this lane's 4 columns (adjoint)
  // shared memory of the warp
  unsigned char* ring;
  float* q_s;
  float* w_s;                // [EB][WRECF]
  uint32_t ring_a, bar_a;    // shared-space addresses of the ring and of the NS mbarriers
  int ks, stage_bytes;
  int cstage, cphase;        // stage / parity of the next batch to consume
  int pstage;                // stage of the next batch to issue

  __device__ __forceinline__ void init(unsigned char* base, const NlkGeom& g, const NlArgs& a, int lane) {
    h1 = lane / EB;
    e1 = lane % EB;
    act1 = lane < H * EB;
    actk = 4 * lane < a.adim;
    kl = actk ? lane : 0;
    hk = (4 * kl) / a.dk;
    ring = base;
    q_s = reinterpret_cast<float*>(base + g.off_q);
    w_s = reinterpret_cast<float*>(base + g.off_w);
    ring_a = smem_addr(base);
    bar_a = smem_addr(base + g.off_bar);
    ks = g.ks;
    stage_bytes = g.stage_bytes;
    cstage = 0;
    cphase = 0;
    pstage = 0;
    if (lane == 0)
      for (int s = 0; s < NS; ++s) mbar_init(bar_a + 8 * s, 1);
  }

  __device__ __forceinline__ void reset() {
    m = -CUDART_INF_F;
    l = 0.f;
    dsum = 0.f;
#pragma unroll
    for (int h = 0; h < H; ++h) acc1[h] = f4_zero();
    k1 = f4_zero();
    k2 = f4_zero();
  }

  // ---- producer ------------------------------------------------------------------------------------------------
  // issue the gather of one batch of `cnt` edges: lanes [0, cnt) copy x[col], lanes [EB, EB + cnt) copy k[col];
  // `cols` = the column id of edge lane % EB (valid on those lanes)
  __device__ __forceinline__ void issue(const NlArgs& a, int cols, int cnt, int lane) {
    const uint32_t bar = bar_a + 8 * pstage;
    const uint32_t xb = 16u * a.dv, kb = 4u * a.adim;
    if (lane == 0) mbar_expect_tx(bar, (xb + kb) * cnt);
    __syncwarp();
    const uint32_t st = ring_a + (uint32_t)(pstage * stage_bytes);
    if (lane < cnt)
      bulk_row_g2s(st + (uint32_t)(lane * kNlkXRow), a.sp.src + (size_t)cols * a.sp.lds, xb, bar);
    else if (lane >= EB && lane - EB < cnt)
      bulk_row_g2s(st + (uint32_t)(EB * kNlkXRow + (lane - EB) * ks), a.wk + (size_t)cols * a.ldwk, kb, bar);
    pstage = pstage + 1 == NS ? 0 : pstage + 1;
  }

  // ---- per-row setup: q -> shared, alpha g_i -> registers ---------------------------------------------------------
  __device__ __forceinline__ void setup_row(const NlArgs& a, int row, int lane, float alpha) {
    for (int j = lane; j < a.adim; j += 32) q_s[j] = __ldg(a.q + (size_t)row * a.ldq + j);
    if constexpr (ADJ) {
      gi = lane < a.dv ? ldg4(a.g + (size_t)row * a.ldg + 4 * lane) : f4_zero();
      f4_scale(gi, alpha);
    }
    __syncwarp();
  }

  // ---- consumer: one batch of `cnt` edges starting at CSR slot `base`, staged in ring stage `cstage` ------------------
  __device__ __forceinline__ void batch(const NlArgs& a, int base, int cnt, int lane, float scale) {
    mbar_wait(bar_a + 8 * cstage, (uint32_t)cphase);
    const unsigned char* stage = ring + cstage * stage_bytes;
    if (++cstage == NS) { cstage = 0; cphase ^= 1; }
    const unsigned char* xcol = stage + 16 * (lane < a.dv ? lane : 0);   // lanes past the row width: column 0, unused
    const unsigned char* krows = stage + EB * kNlkXRow;
    // phase 0 (adjoint): x chunks -> registers, dabar_e = <alpha g_i, x_j>
    float4 xr[ADJ ? EB : 1];
    float dab = 0.f;
    if constexpr (ADJ) {
      float v[EB];
#pragma unroll
      for (int e = 0; e < EB; ++e) {
        xr[e] = *reinterpret_cast<const float4*>(xcol + e * kNlkXRow);
        v[e] = dot4p(gi, xr[e]);                           // gi = 0 on the lanes past the row width
      }
      ReduceT<EB, 16>::run(v, lane);                       // v[0] of lane l = total of edge l >> SHD
      dab = __shfl_sync(kFullMask, v[0], e1 << SHD);
    }
    // phase 1: lane = (head h1, edge e1): s = <q_ih, k_jh> / sqrt(dk)
    float s = 0.f;
    if (act1) {
      const float4* kr = reinterpret_cast<const float4*>(krows + e1 * ks) + h1 * (a.dk >> 2);
      const float4* qr = reinterpret_cast<const float4*>(q_s) + h1 * (a.dk >> 2);
      float2 s0 = make_float2(0.f, 0.f), s1 = s0;
#pragma unroll 4
      for (int t = 0; t < (a.dk >> 2); ++t) {
        const float4 kv = kr[t], qv = qr[t];
        s0 = fma2(make_float2(qv.x, qv.y), make_float2(kv.x, kv.y), s0);
        s1 = fma2(make_float2(qv.z, qv.w), make_float2(kv.z, kv.w), s1);
      }
      s = ((s0.x + s0.y) + (s1.x + s1.y)) * scale;
    }
    // phase 2: online softmax of head h1 over the batch
    const bool valid = act1 && e1 < cnt;                   // slots past the batch hold stale rows: mask them
    s = valid ? s : -CUDART_INF_F;
    if constexpr (ADJ) {
      dab = valid ? dab : 0.f;
      if (valid) {                                         // park the raw score / dabar for the ds sweep
        a.ds[(size_t)(base + e1) * H + h1] = s;
        if (h1 == 0) a.dabar[base + e1] = dab;
      }
    }
    float mb = s;
#pragma unroll
    for (int o = 1; o < EB; o <<= 1) mb = fmaxf(mb, __shfl_xor_sync(kFullMask, mb, o));
    const float mnew = fmaxf(m, mb);
    const float corr = mnew == -CUDART_INF_F ? 1.f : expf(m - mnew);   // first batch: exp(-inf) = 0 on zero sums
    const float pe = valid ? expf(s - mnew) : 0.f;
    m = mnew;
    l = fmaf(l, corr, pe);
    if constexpr (ADJ) dsum = fmaf(dsum, corr, pe * dab);
    if (act1) {
      if constexpr (ADJ) *reinterpret_cast<float2*>(w_s + e1 * WRECF + 2 * h1) = make_float2(pe, pe * dab);
      else w_s[e1 * WRECF + h1] = pe;
    }
    if (__any_sync(kFullMask, corr != 1.f)) {              // some head's running maximum moved: rescale
#pragma unroll
      for (int h = 0; h < H; ++h) f4_scale(acc1[h], __shfl_sync(kFullMask, corr, h * EB));
      if constexpr (ADJ) {
        const float ck = __shfl_sync(kFullMask, corr, hk * EB);
        f4_scale(k1, ck);
        f4_scale(k2, ck);
      }
    }
    __syncwarp();
    // phase 3: lanes span the columns
#pragma unroll
    for (int e = 0; e < EB; ++e) {
      if (e < cnt) {
        float4 xv;
        if constexpr (ADJ) xv = xr[e];
        else xv = *reinterpret_cast<const float4*>(xcol + e * kNlkXRow);
        float w[WRECF];
#pragma unroll
        for (int j = 0; j < WRECF; j += 4) {
          const float4 t = *reinterpret_cast<const float4*>(w_s + e * WRECF + j);
          w[j] = t.x; w[j + 1] = t.y; w[j + 2] = t.z; w[j + 3] = t.w;
        }
#pragma unroll
        for (int h = 0; h < H; ++h) fma4p(acc1[h], w[ADJ ? 2 * h : h], xv);
        if constexpr (ADJ) {
          const float4 kv = *reinterpret_cast<const float4*>(krows + e * ks + 16 * kl);
          const float2 wk = *reinterpret_cast<const float2*>(w_s + e * WRECF + 2 * hk);
          fma4p(k1, wk.x, kv);
          fma4p(k2, wk.y, kv);
        }
      }
    }
    __syncwarp();                                          // the stage and the weight records are free again
  }

  // combine the lanes of the warp: every lane ends with the per-head statistics.  Fixed order -> deterministic.
  __device__ __forceinline__ void warp_merge(NlStats<H>& st) {
#pragma unroll
    for (int o = 1; o < EB; o <<= 1) {                     // over the edge slots (they share the running maximum)
      l += __shfl_xor_sync(kFullMask, l, o);
      if constexpr (ADJ) dsum += __shfl_xor_sync(kFullMask, dsum, o);
    }
#pragma unroll
    for (int h = 0; h < H; ++h) {
      st.mh[h] = __shfl_sync(kFullMask, m, h * EB);
      st.lh[h] = __shfl_sync(kFullMask, l, h * EB);
      st.dh[h] = ADJ ? __shfl_sync(kFullMask, dsum, h * EB) : 0.f;
    }
  }

  // f (diffusion epilogue) and dq of the row from the merged state; executed by one warp
  __device__ __forceinline__ void emit(const NlArgs& a, int row, int lane, const NlStats<H>& st, float alpha,
                                       float beta, float scale) {
    float inv[H];
#pragma unroll
    for (int h = 0; h < H; ++h) inv[h] = st.lh[h] > 0.f ? 1.0f / st.lh[h] : 0.f;
    const float inv_h = 1.0f / (float)H;
    if (!ADJ || a.write_f) {
      float4 ax = f4_zero();
#pragma unroll
      for (int h = 0; h < H; ++h) fma4(ax, inv[h], acc1[h]);
      f4_scale(ax, inv_h);
      if (lane < a.dv) {
        float s1 = 0.f, s2 = 0.f;
        row_epilogue<4, false, !ADJ>(a.sp, row, 4 * lane, Vec<4>{ax}, alpha, beta, s1, s2);
      }
    }
    if constexpr (ADJ) {
      float invk = 0.f, dotk = 0.f;
#pragma unroll
      for (int h = 0; h < H; ++h)
        if (hk == h) { invk = inv[h]; dotk = st.dh[h] * inv[h] * inv_h; }
      const float c2 = invk * inv_h * scale, c1 = invk * dotk * scale;
      if (actk) {
        const float4 dq4 = make_float4(k2.x * c2 - k1.x * c1, k2.y * c2 - k1.y * c1, k2.z * c2 - k1.z * c1,
                                       k2.w * c2 - k1.w * c1);
        *reinterpret_cast<float4*>(a.dq + (size_t)row * a.lddq + 4 * kl) = dq4;
      }
    }
  }
};

// A hub row (more edges than a.long_threshold) is worked by the whole CTA: warp w takes the batches w, w + W, ...
// through its own ring, then the partial online-softmax states are merged in warp order: statistics through the
// CTA-level [W][3][H] slots, accumulators summed warp after warp into the CTA-level [H + 2][32] float4 slots
// (T1_h, K1, K2).  Called by every thread of the CTA.
template <int H, bool ADJ, int W>
__device__ __forceinline__ void nlk_long_row(const NlArgs& a, NlkWarp<H, ADJ>& r, int row, float* pst, float4* pacc,
                                             float alpha, float beta, float scale) {
  using Wp = NlkWarp<H, ADJ>;
  constexpr int EB = Wp::EB, NS = Wp::NS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int start = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
  r.setup_row(a, row, lane, alpha);
  r.reset();
  {
    const int stride = W * EB;
    int poff = start + warp * EB;                          // producer cursor
    auto load_cols = [&]() {
      return (poff < end && lane < 2 * EB && lane % EB < min(EB, end - poff)) ? ld_stream_i32(a.colidx + poff + lane % EB) : 0;
    };
    int pcols = load_cols();
    auto produce = [&]() {
      if (poff < end) {
        r.issue(a, pcols, min(EB, end - poff), lane);
        poff += stride;
        pcols = load_cols();
      }
    };
#pragma unroll 1
    for (int s = 0; s < NS - 1; ++s) produce();
#pragma unroll 1
    for (int base = start + warp * EB; base < end; base += stride) {
      produce();
      r.batch(a, base, min(EB, end - base), lane, scale);
    }
  }
  NlStats<H> st;
  r.warp_merge(st);
  if (lane == 0) {
#pragma unroll
    for (int h = 0; h < H; ++h) {
      pst[(warp * 3 + 0) * H + h] = st.mh[h];
      pst[(warp * 3 + 1) * H + h] = st.lh[h];
      pst[(warp * 3 + 2) * H + h] = st.dh[h];
    }
  }
  __syncthreads();
  float scw[H];                                             // this warp's rescale to the merged maximum
#pragma unroll
  for (int h = 0; h < H; ++h) {
    float mx = -CUDART_INF_F;
    for (int w = 0; w < W; ++w) mx = fmaxf(mx, pst[(w * 3 + 0) * H + h]);
    float ls = 0.f, dsum = 0.f;
    for (int w = 0; w < W; ++w) {
      const float mw = pst[(w * 3 + 0) * H + h];
      const float sc = mw == -CUDART_INF_F ? 0.f : expf(mw - mx);
      ls = fmaf(pst[(w * 3 + 1) * H + h], sc, ls);
      dsum = fmaf(pst[(w * 3 + 2) * H + h], sc, dsum);
    }
    scw[h] = st.mh[h] == -CUDART_INF_F ? 0.f : expf(st.mh[h] - mx);
    st.mh[h] = mx;
    st.lh[h] = ls;
    st.dh[h] = dsum;
  }
  float sck = 0.f;
#pragma unroll
  for (int h = 0; h < H; ++h)
    if (r.hk == h) sck = scw[h];
  for (int w = 0; w < W; ++w) {                             // warp 0 stores, warps 1.. add in turn (fixed order)
    if (warp == w) {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float4 t1 = w == 0 ? f4_zero() : pacc[h * 32 + lane];
        fma4(t1, scw[h], r.acc1[h]);
        pacc[h * 32 + lane] = t1;
      }
      if constexpr (ADJ) {
        float4 t = w == 0 ? f4_zero() : pacc[H * 32 + lane];
        fma4(t, sck, r.k1);
        pacc[H * 32 + lane] = t;
        t = w == 0 ? f4_zero() : pacc[(H + 1) * 32 + lane];
        fma4(t, sck, r.k2);
        pacc[(H + 1) * 32 + lane] = t;
      }
    }
    __syncthreads();
  }
  if (warp == 0) {
#pragma unroll
    for (int h = 0; h < H; ++h) r.acc1[h] = pacc[h * 32 + lane];
    if constexpr (ADJ) {
      r.k1 = pacc[H * 32 + lane];
      r.k2 = pacc[(H + 1) * 32 + lane];
    }
    r.emit(a, row, lane, st, alpha, beta, scale);
  }
  if constexpr (ADJ) nl_sweep_ds<H>(a, start, end, lane, st, warp * 32, W * 32);
  __syncthreads();                                          // pst / pacc / q slots are reused by the next row
}

// persistent kernel: one CTA of W warps per SM, warps pull chunks of rows from an atomic counter
template <int H, bool ADJ, int W>
__global__ void __launch_bounds__(W * 32, 1) nlk_rows_kernel(const NlArgs a, const int chunk_rows) {
  extern __shared__ __align__(128) unsigned char nlk_smem[];
  using Wp = NlkWarp<H, ADJ>;
  constexpr int EB = Wp::EB, NS = Wp::NS;
  const NlkGeom geo = nlk_geom<H, ADJ, W>(a.adim);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Wp r;
  r.init(nlk_smem + (size_t)warp * geo.warp_bytes, geo, a, lane);
  float* pst = reinterpret_cast<float*>(nlk_smem + geo.pst_off);
  float4* pacc = reinterpret_cast<float4*>(nlk_smem + geo.pacc_off);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const float alpha = load_alpha(a.sp);
  const float beta = a.sp.beta_train != nullptr ? __ldg(a.sp.beta_train) : 0.f;
  const float scale = 1.0f / sqrtf((float)a.dk);
  const int nrows = a.row_hi - a.row_lo;
  // hub rows first, one CTA each: they finish long before the scheduler loop below runs dry
  for (int li = blockIdx.x; li < a.n_long; li += gridDim.x) {
    const int row = __ldg(a.long_list + li);
    if (row < a.row_lo || row >= a.row_hi) continue;
    if (__ldg(a.rowptr + row + 1) - __ldg(a.rowptr + row) <= a.long_threshold) continue;   // a warp takes it below
    nlk_long_row<H, ADJ, W>(a, r, row, pst, pacc, alpha, beta, scale);
  }
  for (;;) {
    int r0 = 0;
    if (lane == 0) r0 = atomicAdd(a.counter, chunk_rows);
    r0 = __shfl_sync(kFullMask, r0, 0);
    if (r0 >= nrows) break;
    const int nch = min(chunk_rows, nrows - r0);
    // the chunk's row pointers in one load (lanes 0 .. nch), handed out by shuffles
    const int rp = lane <= nch ? __ldg(a.rowptr + a.row_lo + r0 + lane) : 0;
    // producer cursor: runs NS - 1 batches ahead of the arithmetic, across the rows of the chunk
    int pr = -1, poff = 0, pend = 0, pcols = 0;
    auto next_row = [&]() {                                  // move to the next row a warp works on (or past the chunk)
      while (poff >= pend) {
        if (++pr >= nch) break;
        poff = __shfl_sync(kFullMask, rp, pr);
        pend = __shfl_sync(kFullMask, rp, pr + 1);
        if (pend - poff > a.long_threshold) poff = pend;    // done CTA-wide above
      }
      pcols = (pr < nch && lane < 2 * EB && lane % EB < min(EB, pend - poff)) ? ld_stream_i32(a.colidx + poff + lane % EB) : 0;
    };
    auto produce = [&]() {
      if (pr < nch) {
        r.issue(a, pcols, min(EB, pend - poff), lane);
        poff += EB;
        next_row();
      }
    };
    next_row();
#pragma unroll 1
    for (int s = 0; s < NS - 1; ++s) produce();
#pragma unroll 1
    for (int rr = 0; rr < nch; ++rr) {
      const int row = a.row_lo + r0 + rr;
      const int start = __shfl_sync(kFullMask, rp, rr), end = __shfl_sync(kFullMask, rp, rr + 1);
      if (end - start > a.long_threshold) continue;
      r.setup_row(a, row, lane, alpha);
      r.reset();
#pragma unroll 1
      for (int base = start; base < end; base += EB) {
        produce();
        r.batch(a, base, min(EB, end - base), lane, scale);
      }
      NlStats<H> st;
      r.warp_merge(st);
      r.emit(a, row, lane, st, alpha, beta, scale);
      if constexpr (ADJ) {
        __syncwarp();
        nl_sweep_ds<H>(a, start, end, lane, st, 0, 32);
      }
      __syncwarp();                                          // the q slot is reused by the next row
    }
  }
}

// GB_ERR_UNSUPPORTED: the shape does not fit the shared memory of W warps
template <int H, bool ADJ, int W>
int nlk_launch(NlArgs a, int n_long, cudaStream_t stream) {
  static bool configured = false;
  const NlkGeom geo = nlk_geom<H, ADJ, W>(a.adim);
  if ((size_t)geo.total > kNlSmemLimit) return GB_ERR_UNSUPPORTED;
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(nlk_rows_kernel<H, ADJ, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNlSmemLimit));
    configured = true;
  }
  const int nrows = a.row_hi - a.row_lo;
  if (nrows <= 0) return GB_OK;
  a.n_long = n_long;
  GB_CUDA(cudaMemsetAsync(a.counter, 0, sizeof(int32_t), stream));
  // rows per scheduler grab: up to kNl2ChunkRows, fewer on small graphs so that every warp of the machine gets work
  const int warps = W * sm_count();
  int chunk = nrows / (4 * warps);
  chunk = chunk < 1 ? 1 : (chunk > kNl2ChunkRows ? kNl2ChunkRows : chunk);
  const int want = (nrows + W * chunk - 1) / (W * chunk);
  const int blocks = min_int(want, sm_count());
  {
    LaunchScope ls(ADJ ? K_NL_ADJ_ROWS : K_NL_FWD, stream);
    nlk_rows_kernel<H, ADJ, W><<<blocks, W * 32, geo.total, stream>>>(a, chunk);
  }
  return check_launch("nlk_rows_kernel");
}

template <bool ADJ>
int nlk_dispatch(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
#define GB_NLK_CASE(HV)                                                           \
  if (heads == HV) {                                                              \
    const int rc = nlk_launch<HV, ADJ, 16>(a, n_long, stream);                    \
    return rc != GB_ERR_UNSUPPORTED ? rc : nlk_launch<HV, ADJ, 8>(a, n_long, stream); \
  }
  GB_NLK_CASE(4) GB_NLK_CASE(8) GB_NLK_CASE(2) GB_NLK_CASE(1)
#undef GB_NLK_CASE
  set_error("nlk kernels: unsupported heads=%d", heads);
  return GB_ERR_UNSUPPORTED;
}

}  // namespace gb
