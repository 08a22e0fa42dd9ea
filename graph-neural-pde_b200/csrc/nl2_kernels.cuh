// nl2_kernels.cuh - second generation of the single-gather GRAND-nl row kernels (forward evaluation and the row pass
// of the adjoint; ODEFuncTransformerAtt.forward, src/function_transformer_attention.py:44-59, and its VJP).
// Same mathematics, arguments and results as nl_kernels.cuh (read its header first); what changed is how a warp
// spends its issue slots.  The first generation is bound by instruction issue and dependency latency (ncu,
// profiles/r01x: 84 warp instructions per edge in the adjoint loop, 58 % issue utilisation, 31 % DRAM throughput):
// the gathered row lives in registers with the lanes spanning its columns, so every score needs a 32-lane
// transposing butterfly (124 instructions per 8 edges), every softmax weight a shuffle, every gather its own address
// arithmetic, and the loads of a warp are in flight only while it is not computing.  Here the gathered rows are
// STAGED IN SHARED MEMORY BY THE TMA UNIT:
//
//   gather    one `cp.async.bulk` (UBLKCP: global -> shared, completion on an mbarrier) per edge copies the whole
//             x[col] row; a warp issues a batch of 8 rows with ONE instruction (one row per lane) into its own ring
//             of kNl2Stages stages and runs a batch ahead of its arithmetic - across row boundaries inside a chunk
//             of rows - so the gather is off every critical path and costs no registers and no scoreboard slots
//             (scratch/gather_bench.cu: 7.4 - 7.6 TB/s of random 512-byte rows with 12 - 16 warps per SM and 2 stages)
//   phase 1   lane = (edge e = lane % 8, row quarter cq = lane / 8): the lane reads every fourth 16-byte chunk of
//             row e ONCE (conflict-free: the 8 lanes of a quarter warp read 8 rows whose stride is 16 XL + 16 bytes)
//             and the matching chunks of p[h] (= Wk_h^T q_ih / sqrt(dk), formed once per row, broadcast reads) for
//             all heads, [adjoint: and of alpha g_i]; the quarter sums are combined over the 4 lanes of an edge
//             with a 2-stage transposing exchange that leaves head cq on lane (e, cq)
//   phase 2   online softmax bookkeeping on one (edge, head) pair per lane; exp(s - m) [and exp(s - m) dabar] go to a
//             small record per edge in shared memory; raw score / dabar are parked for the ds sweep (coalesced)
//   phase 3   lanes span the columns: per edge one LDS.128 of the row chunk, the weight record as broadcast LDS.128,
//             2 FFMA2 per (head, accumulator) with the weight as scalar-broadcast operand
//
// Shared-memory traffic is ~1.5 KB per edge (each staged row is read once in phase 1 and once in phase 3).
// Per-row work (p, f epilogue, dq, ds sweep) is unchanged.  Hub rows (> long_threshold edges) are worked CTA-wide
// inside the same persistent launch as before.  No float atomics; the order of every sum is fixed by the row's edge
// order, so results are bit-reproducible run to run (they differ from the first generation's in the last bits:
// different association of the score sums).
//
// Roofline: HBM.  Algorithmic bytes per launch: see nl_kernels.cuh (identical).
#pragma once
#include "nl_kernels.cuh"

namespace gb {

#ifndef GB_NL2_STAGES
#define GB_NL2_STAGES 2
#endif
#ifndef GB_NL2_CHUNK_ROWS
#define GB_NL2_CHUNK_ROWS 16
#endif
constexpr int kNl2Stages = GB_NL2_STAGES;          // ring stages per warp (kNl2Stages - 1 batches in flight)
constexpr int kNl2ChunkRows = GB_NL2_CHUNK_ROWS;   // rows per scheduler grab (<= 31: one lane per row pointer)
constexpr int kNl2Edges = 8;                       // edges per batch
static_assert(kNl2Stages >= 2 && kNl2Stages <= 8, "ring depth");
static_assert(kNl2ChunkRows >= 1 && kNl2ChunkRows <= 31, "a chunk's row pointers are loaded by one warp instruction");

// ---- mbarrier / bulk-copy primitives (PTX ISA 8.x; SASS: SYNCS.*, UBLKCP) ---------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n NL2_WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra NL2_DONE_%=;\n"
      " bra NL2_WAIT_%=;\n NL2_DONE_%=:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// one whole row, global -> shared, through the TMA unit; completes `bytes` on the mbarrier
__device__ __forceinline__ void bulk_row_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// shared-memory geometry (bytes); the same numbers on the host (launch) and on the device.
// CTA: [Wk staged [A][ldw]] [W warp regions] [CTA-level statistics of the CTA-wide rows [W][3][H]]
// warp region: [ring: NS stages x 8 rows, stride RS] [p / u: H rows, stride RS] [alpha g_i] [q] [weight records] [mbarriers]
// RS = 16 XL + 16 bytes: the +16 puts the 8 rows of a batch (and the heads of p) on different banks.
struct Nl2Geom {
  int off_p, off_g, off_q, off_w, off_bar;   // offsets inside a warp's region (the ring comes first)
  int warp_bytes;
  int wk_bytes;
  int pst_off;
  int total;
};
template <int XL, int H, bool ADJ, int W>
__host__ __device__ inline Nl2Geom nl2_geom(int adim, int dv) {
  constexpr int WRECF = ((ADJ ? 2 * H : H) + 3) / 4 * 4;
  constexpr int RS = 16 * XL + 16;
  Nl2Geom g;
  g.off_p = kNl2Stages * kNl2Edges * RS;
  g.off_g = g.off_p + H * RS;
  g.off_q = g.off_g + (ADJ ? 16 * XL : 0);
  g.off_w = g.off_q + ((4 * adim + 15) / 16) * 16;
  g.off_bar = g.off_w + kNl2Edges * WRECF * 4;
  g.warp_bytes = ((g.off_bar + 8 * kNl2Stages + 127) / 128) * 128;
  g.wk_bytes = ((adim * nl_ldw(dv) * 4 + 127) / 128) * 128;
  g.pst_off = g.wk_bytes + W * g.warp_bytes;
  g.total = g.pst_off + W * 3 * H * 4;
  return g;
}

// 2-stage transposing exchange over the 4 lanes (e, cq = 0..3) of an edge (xor 16, then xor 8): every lane holds NV
// quarter sums; afterwards v[0 .. max(NV/4, 1)) are edge totals - see Nl2Warp::HPL / head0 / owner for which heads.
template <int NV>
__device__ __forceinline__ void quad_reduce(float* v, int lane) {
  if constexpr (NV >= 2) {
    constexpr int HN = NV / 2;
    const bool up = (lane & 16) != 0;
#pragma unroll
    for (int i = 0; i < HN; ++i) {
      const float send = up ? v[i] : v[i + HN], keep = up ? v[i + HN] : v[i];
      v[i] = keep + __shfl_xor_sync(kFullMask, send, 16);
    }
    if constexpr (HN >= 2) {
      constexpr int QN = HN / 2;
      const bool up8 = (lane & 8) != 0;
#pragma unroll
      for (int i = 0; i < QN; ++i) {
        const float send = up8 ? v[i] : v[i + QN], keep = up8 ? v[i + QN] : v[i];
        v[i] = keep + __shfl_xor_sync(kFullMask, send, 8);
      }
    } else {
      v[0] += __shfl_xor_sync(kFullMask, v[0], 8);
    }
  } else {
    v[0] += __shfl_xor_sync(kFullMask, v[0], 16);
    v[0] += __shfl_xor_sync(kFullMask, v[0], 8);
  }
}

// Register / shared-memory state of one warp.
template <int XL, int H, bool ADJ>
struct Nl2Warp {
  static constexpr int EB = kNl2Edges;                      // edges per batch
  static constexpr int EPI = 32 / XL;                       // edge groups in the accumulate phase
  static constexpr int NS = kNl2Stages;
  static constexpr int HA = ADJ ? H : 1;
  static constexpr int WREC = ADJ ? 2 * H : H;              // floats per edge in the weight record
  static constexpr int WRECF = (WREC + 3) / 4 * 4;
  static constexpr int RS = 16 * XL + 16;                   // bytes between staged rows, and between p[h] / u[h]
  static constexpr int HPL = H >= 4 ? H / 4 : 1;            // heads per lane after quad_reduce
  static constexpr int SHQ = H >= 4 ? 0 : (H == 2 ? 1 : 2);  // lanes (e, cq) with equal cq >> SHQ hold the same heads
  static_assert(H == 1 || H == 2 || H == 4 || H == 8, "unsupported head count");

  int e1, cq;                // phase-1/2 identity of the lane: edge slot, row quarter
  int head0;                 // first of the HPL heads this lane keeps the softmax state of
  bool owner;                // this lane writes the records of its heads (H < 4: the heads are replicated)
  int xl, es;                // phase-3 identity of the lane: column chunk, edge group
  float m[HPL], l[HPL], dsum[HPL];   // online softmax state of (this lane's heads, this lane's edge slot)
  float4 acc1[H];            // T1_h, this lane's 4 columns
  float4 acc2[HA];           // T2_h
  // shared memory of the warp
  unsigned char* ring;
  float* p_s;                // [H] rows of stride RS; reused as u_s by emit()
  float* g_s;                // alpha * g_i (adjoint)
  float* q_s;
  float* w_s;                // [EB][WRECF]
  uint32_t ring_a, bar_a;    // shared-space addresses of the ring and of the NS mbarriers
  int cstage, cphase;        // stage / parity of the next batch to consume
  int pstage;                // stage of the next batch to issue

  __device__ __forceinline__ void init(unsigned char* base, const Nl2Geom& g, int lane) {
    e1 = lane & 7;
    cq = lane >> 3;
    head0 = (cq >> SHQ) * HPL;
    owner = (cq & ((1 << SHQ) - 1)) == 0;
    xl = lane % XL;
    es = lane / XL;
    ring = base;
    p_s = reinterpret_cast<float*>(base + g.off_p);
    g_s = reinterpret_cast<float*>(base + g.off_g);
    q_s = reinterpret_cast<float*>(base + g.off_q);
    w_s = reinterpret_cast<float*>(base + g.off_w);
    ring_a = smem_addr(base);
    bar_a = smem_addr(base + g.off_bar);
    cstage = 0;
    cphase = 0;
    pstage = 0;
    if (lane == 0)
      for (int s = 0; s < NS; ++s) mbar_init(bar_a + 8 * s, 1);
  }

  __device__ __forceinline__ void reset() {
#pragma unroll
    for (int k = 0; k < HPL; ++k) { m[k] = -CUDART_INF_F; l[k] = 0.f; dsum[k] = 0.f; }
#pragma unroll
    for (int h = 0; h < H; ++h) acc1[h] = f4_zero();
#pragma unroll
    for (int h = 0; h < HA; ++h) acc2[h] = f4_zero();
  }

  // ---- producer ------------------------------------------------------------------------------------------------
  // issue the gather of one batch: `cnt` rows, column ids in lanes [0, cnt) of `cols`
  __device__ __forceinline__ void issue(const NlArgs& a, int cols, int cnt, int lane) {
    const uint32_t bar = bar_a + 8 * pstage;
    const uint32_t row_bytes = 16u * a.dv;
    if (lane == 0) mbar_expect_tx(bar, row_bytes * cnt);
    __syncwarp();
    if (lane < cnt)
      bulk_row_g2s(ring_a + (uint32_t)((pstage * EB + lane) * RS), a.sp.src + (size_t)cols * a.sp.lds, row_bytes, bar);
    pstage = pstage + 1 == NS ? 0 : pstage + 1;
  }

  // ---- per-row setup ---------------------------------------------------------------------------------------------
  // q -> shared, p[h] = Wk_h^T q_ih / sqrt(dk) -> shared, alpha g_i -> shared
  __device__ __forceinline__ void setup_row(const NlArgs& a, int row, int lane, const float* wk_s, int ldw, float alpha) {
    for (int j = lane; j < a.adim; j += 32) q_s[j] = __ldg(a.q + (size_t)row * a.ldq + j);
    float4 gi = f4_zero();
    if constexpr (ADJ)
      if (xl < a.dv) gi = ldg4(a.g + (size_t)row * a.ldg + 4 * xl);
    __syncwarp();
    const float scale = 1.0f / sqrtf((float)a.dk);
    const bool act = xl < a.dv;
    const float* wcol = wk_s + 4 * (act ? xl : 0);
#pragma unroll
    for (int h = 0; h < H; ++h) {
      float4 acc = f4_zero();
#pragma unroll 2
      for (int t = 0; t < a.dk; t += 4) {
        const int j = h * a.dk + t;
        const float4 qv = *reinterpret_cast<const float4*>(q_s + j);
        fma4p(acc, qv.x, *reinterpret_cast<const float4*>(wcol + (size_t)(j + 0) * ldw));
        fma4p(acc, qv.y, *reinterpret_cast<const float4*>(wcol + (size_t)(j + 1) * ldw));
        fma4p(acc, qv.z, *reinterpret_cast<const float4*>(wcol + (size_t)(j + 2) * ldw));
        fma4p(acc, qv.w, *reinterpret_cast<const float4*>(wcol + (size_t)(j + 3) * ldw));
      }
      f4_scale(acc, act ? scale : 0.f);
      if (es == 0) *reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(p_s) + h * RS + 16 * xl) = acc;
    }
    if constexpr (ADJ) {
      f4_scale(gi, alpha);
      if (es == 0) *reinterpret_cast<float4*>(g_s + 4 * xl) = gi;
    }
    __syncwarp();
  }

  // ---- consumer: one batch of `cnt` edges starting at CSR slot `base`, staged in ring stage `cstage` ------------------
  __device__ __forceinline__ void batch(const NlArgs& a, int base, int cnt, int lane) {
    mbar_wait(bar_a + 8 * cstage, (uint32_t)cphase);
    const unsigned char* stage = ring + cstage * (EB * RS);
    if (++cstage == NS) { cstage = 0; cphase ^= 1; }
    // phase 1: lane = (edge e1, quarter cq) reads chunks cq, cq + 4, ... of row e1 and of every p[h]
    float sc[H];
    float dab = 0.f;
    {
      const unsigned char* xrow = stage + e1 * RS;
      const unsigned char* prow = reinterpret_cast<const unsigned char*>(p_s);
      float2 lo[H], hi[H];
#pragma unroll
      for (int h = 0; h < H; ++h) lo[h] = hi[h] = make_float2(0.f, 0.f);
      float2 dlo = make_float2(0.f, 0.f), dhi = dlo;
#pragma unroll 4
      for (int c = cq; c < a.dv; c += 4) {
        const float4 xv = *reinterpret_cast<const float4*>(xrow + 16 * c);
#pragma unroll
        for (int h = 0; h < H; ++h) {
          const float4 pv = *reinterpret_cast<const float4*>(prow + h * RS + 16 * c);
          lo[h] = fma2(make_float2(pv.x, pv.y), make_float2(xv.x, xv.y), lo[h]);
          hi[h] = fma2(make_float2(pv.z, pv.w), make_float2(xv.z, xv.w), hi[h]);
        }
        if constexpr (ADJ) {
          const float4 gv = *reinterpret_cast<const float4*>(reinterpret_cast<const unsigned char*>(g_s) + 16 * c);
          dlo = fma2(make_float2(gv.x, gv.y), make_float2(xv.x, xv.y), dlo);
          dhi = fma2(make_float2(gv.z, gv.w), make_float2(xv.z, xv.w), dhi);
        }
      }
#pragma unroll
      for (int h = 0; h < H; ++h) sc[h] = (lo[h].x + lo[h].y) + (hi[h].x + hi[h].y);
      quad_reduce<H>(sc, lane);                            // sc[0 .. HPL) = scores of heads head0 .. of edge e1
      if constexpr (ADJ) {
        dab = (dlo.x + dlo.y) + (dhi.x + dhi.y);
        dab += __shfl_xor_sync(kFullMask, dab, 16);
        dab += __shfl_xor_sync(kFullMask, dab, 8);
      }
    }
    // phase 2: online softmax of this lane's heads over the batch
    const bool valid = e1 < cnt;                           // slots past the batch hold a stale row: mask them
    if constexpr (ADJ) {
      dab = valid ? dab : 0.f;
      if (valid && cq == 0) a.dabar[base + e1] = dab;      // park dabar / the raw scores for the ds sweep
    }
    bool moved = false;
    float corr[HPL];
#pragma unroll
    for (int k = 0; k < HPL; ++k) {
      const float s = valid ? sc[k] : -CUDART_INF_F;
      if constexpr (ADJ)
        if (valid && owner) a.ds[(size_t)(base + e1) * H + head0 + k] = s;
      float mb = s;
      mb = fmaxf(mb, __shfl_xor_sync(kFullMask, mb, 1));
      mb = fmaxf(mb, __shfl_xor_sync(kFullMask, mb, 2));
      mb = fmaxf(mb, __shfl_xor_sync(kFullMask, mb, 4));
      const float mnew = fmaxf(m[k], mb);                  // finite: edge 0 of the batch is valid
      corr[k] = expf(m[k] - mnew);                         // first batch: exp(-inf) = 0 on zero accumulators
      const float pe = valid ? expf(s - mnew) : 0.f;
      m[k] = mnew;
      l[k] = fmaf(l[k], corr[k], pe);
      if constexpr (ADJ) dsum[k] = fmaf(dsum[k], corr[k], pe * dab);
      if (owner) {
        w_s[e1 * WRECF + head0 + k] = pe;
        if constexpr (ADJ) w_s[e1 * WRECF + H + head0 + k] = pe * dab;
      }
      moved = moved || corr[k] != 1.f;
    }
    if (__any_sync(kFullMask, moved)) {                    // some head's running maximum moved: rescale
#pragma unroll
      for (int h = 0; h < H; ++h) {
        // head h lives on the lanes with cq >> SHQ == h / HPL; lane (e = 0, that cq) answers
        const float c = __shfl_sync(kFullMask, corr[h % HPL], ((h / HPL) << SHQ) << 3);
        f4_scale(acc1[h], c);
        if constexpr (ADJ) f4_scale(acc2[h], c);
      }
    }
    __syncwarp();
    // phase 3: lanes span the columns (lanes past the row width read column 0 and carry unused sums)
    {
      const unsigned char* xcol = stage + 16 * (xl < a.dv ? xl : 0);
      auto edge = [&](int e) {
        const float4 xv = *reinterpret_cast<const float4*>(xcol + e * RS);
        float w[WRECF];
#pragma unroll
        for (int j = 0; j < WRECF; j += 4) {
          const float4 t = *reinterpret_cast<const float4*>(w_s + e * WRECF + j);
          w[j] = t.x; w[j + 1] = t.y; w[j + 2] = t.z; w[j + 3] = t.w;
        }
#pragma unroll
        for (int h = 0; h < H; ++h) {
          fma4p(acc1[h], w[h], xv);
          if constexpr (ADJ) fma4p(acc2[h], w[H + h], xv);
        }
      };
      if (cnt == EB) {                                     // full batch: no per-edge tests
#pragma unroll
        for (int i = 0; i < EB / EPI; ++i) edge(i * EPI + es);
      } else {
#pragma unroll 1
        for (int e = es; e < cnt; e += EPI) edge(e);
      }
    }
    __syncwarp();                                          // the stage and the weight records are free again
  }

  // combine the lanes of the warp: every lane ends with the per-head statistics and (for its 4 columns) the
  // accumulators of the whole warp.  Fixed order -> deterministic.
  __device__ __forceinline__ void warp_merge(NlStats<H>& st) {
#pragma unroll
    for (int k = 0; k < HPL; ++k) {
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {                    // over the edge slots (they share the running maximum)
        l[k] += __shfl_xor_sync(kFullMask, l[k], o);
        if constexpr (ADJ) dsum[k] += __shfl_xor_sync(kFullMask, dsum[k], o);
      }
    }
#pragma unroll
    for (int h = 0; h < H; ++h) {
      const int src = ((h / HPL) << SHQ) << 3;
      st.mh[h] = __shfl_sync(kFullMask, m[h % HPL], src);
      st.lh[h] = __shfl_sync(kFullMask, l[h % HPL], src);
      st.dh[h] = ADJ ? __shfl_sync(kFullMask, dsum[h % HPL], src) : 0.f;
    }
    if constexpr (EPI > 1) {                               // the edge groups share the running maximum: plain sums
#pragma unroll
      for (int h = 0; h < H; ++h) {
#pragma unroll
        for (int o = XL; o < 32; o <<= 1) {
          f4_add(acc1[h], f4_shfl_xor(acc1[h], o));
          if constexpr (ADJ) f4_add(acc2[h], f4_shfl_xor(acc2[h], o));
        }
      }
    }
  }

  // f (diffusion epilogue) and dq of the row from the merged state; executed by one warp.  u goes through the p slots.
  __device__ __forceinline__ void emit(const NlArgs& a, int row, int lane, const NlStats<H>& st, float alpha,
                                       float beta, const float* wk_s, int ldw) {
    float inv[H];
#pragma unroll
    for (int h = 0; h < H; ++h) inv[h] = st.lh[h] > 0.f ? 1.0f / st.lh[h] : 0.f;
    const float inv_h = 1.0f / (float)H;
    if (!ADJ || a.write_f) {
      float4 ax = f4_zero();
#pragma unroll
      for (int h = 0; h < H; ++h) fma4(ax, inv[h], acc1[h]);
      f4_scale(ax, inv_h);
      if (es == 0 && xl < a.dv) {
        float s1 = 0.f, s2 = 0.f;
        row_epilogue<4, false>(a.sp, row, 4 * xl, Vec<4>{ax}, alpha, beta, s1, s2);
      }
    }
    if constexpr (ADJ) {
      constexpr int ups = RS / 4;                          // floats between u[h] and u[h + 1]
      float* u_s = p_s;
      __syncwarp();                                        // every lane is done with p
#pragma unroll
      for (int h = 0; h < H; ++h) {
        const float dot = st.dh[h] * inv[h] * inv_h;
        const float c2 = inv[h] * inv_h, c1 = inv[h] * dot;
        float4 uh;
        uh.x = acc2[h].x * c2 - acc1[h].x * c1;
        uh.y = acc2[h].y * c2 - acc1[h].y * c1;
        uh.z = acc2[h].z * c2 - acc1[h].z * c1;
        uh.w = acc2[h].w * c2 - acc1[h].w * c1;
        if (es == 0) *reinterpret_cast<float4*>(u_s + h * ups + 4 * xl) = uh;
      }
      __syncwarp();
      const float scale = 1.0f / sqrtf((float)a.dk);
      for (int j = lane; j < a.adim; j += 32) {              // dq[i, j] = <Wk[j, :], u_{head(j)}> / sqrt(dk)
        const float* wr = wk_s + (size_t)j * ldw;
        const float* ur = u_s + (j / a.dk) * ups;
        float2 t0 = make_float2(0.f, 0.f), t1 = make_float2(0.f, 0.f);
#pragma unroll 4
        for (int c = 0; c < a.dv; ++c) {
          const float4 wv = *reinterpret_cast<const float4*>(wr + 4 * c), uv = *reinterpret_cast<const float4*>(ur + 4 * c);
          t0 = fma2(make_float2(wv.x, wv.y), make_float2(uv.x, uv.y), t0);
          t1 = fma2(make_float2(wv.z, wv.w), make_float2(uv.z, uv.w), t1);
        }
        a.dq[(size_t)row * a.lddq + j] = ((t0.x + t1.x) + (t0.y + t1.y)) * scale;
      }
    }
  }
};

__device__ __forceinline__ void nl2_stage_wk(const NlArgs& a, float* wk_s, int ldw) {
  const int total = a.adim * a.dv;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int j = i / a.dv, c = i - j * a.dv;
    *reinterpret_cast<float4*>(wk_s + (size_t)j * ldw + 4 * c) = ldg4(a.wk + (size_t)j * a.ldwk + 4 * c);
  }
}

// A hub row (more edges than a.long_threshold) is worked by the whole CTA: warp w takes the batches w, w + W, ...
// through its own ring, then the partial online-softmax states are merged in warp order.  Statistics go through the
// CTA-level [W][3][H] slots, accumulators are summed warp after warp into the p slots of warps 0 (T1) and 1 (T2).
// Called by every thread of the CTA.
template <int XL, int H, bool ADJ, int W>
__device__ __forceinline__ void nl2_long_row(const NlArgs& a, Nl2Warp<XL, H, ADJ>& r, int row, const float* wk_s, int ldw,
                                             float* pst, float* pacc1, float* pacc2, float alpha, float beta) {
  using Wp = Nl2Warp<XL, H, ADJ>;
  constexpr int EB = Wp::EB, NS = Wp::NS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int start = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
  r.setup_row(a, row, lane, wk_s, ldw, alpha);             // every warp forms p in its own slots
  r.reset();
  {
    const int stride = W * EB;
    int poff = start + warp * EB;                          // producer cursor
    int pcols = (poff < end && lane < min(EB, end - poff)) ? ld_stream_i32(a.colidx + poff + lane) : 0;
    auto produce = [&]() {
      if (poff < end) {
        r.issue(a, pcols, min(EB, end - poff), lane);
        poff += stride;
        pcols = (poff < end && lane < min(EB, end - poff)) ? ld_stream_i32(a.colidx + poff + lane) : 0;
      }
    };
#pragma unroll 1
    for (int s = 0; s < NS - 1; ++s) produce();
#pragma unroll 1
    for (int base = start + warp * EB; base < end; base += stride) {
      produce();
      r.batch(a, base, min(EB, end - base), lane);
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
  __syncthreads();                                          // also: every warp is done with its p
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
  constexpr int ups = Wp::RS / 4;
  for (int w = 0; w < W; ++w) {                             // warp 0 stores, warps 1.. add in turn (fixed order)
    if (warp == w && r.es == 0) {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float4* d1 = reinterpret_cast<float4*>(pacc1 + h * ups + 4 * r.xl);
        float4 t1 = w == 0 ? f4_zero() : *d1;
        fma4(t1, scw[h], r.acc1[h]);
        *d1 = t1;
        if constexpr (ADJ) {
          float4* d2 = reinterpret_cast<float4*>(pacc2 + h * ups + 4 * r.xl);
          float4 t2 = w == 0 ? f4_zero() : *d2;
          fma4(t2, scw[h], r.acc2[h]);
          *d2 = t2;
        }
      }
    }
    __syncthreads();
  }
  if (warp == 0) {
#pragma unroll
    for (int h = 0; h < H; ++h) {
      r.acc1[h] = *reinterpret_cast<const float4*>(pacc1 + h * ups + 4 * r.xl);
      if constexpr (ADJ) r.acc2[h] = *reinterpret_cast<const float4*>(pacc2 + h * ups + 4 * r.xl);
    }
    r.emit(a, row, lane, st, alpha, beta, wk_s, ldw);      // u goes through warp 0's p slots (= pacc1, already read)
  }
  if constexpr (ADJ) nl_sweep_ds<H>(a, start, end, lane, st, warp * 32, W * 32);
  __syncthreads();                                          // the p slots go back to their warps
}

// persistent kernel: one CTA of W warps per SM, warps pull chunks of rows from an atomic counter
template <int XL, int H, bool ADJ, int W>
__global__ void __launch_bounds__(W * 32, 1) nl2_rows_kernel(const NlArgs a, const int chunk_rows) {
  extern __shared__ __align__(128) unsigned char nl2_smem[];
  using Wp = Nl2Warp<XL, H, ADJ>;
  constexpr int EB = Wp::EB, NS = Wp::NS;
  static_assert(W >= 2, "CTA-wide rows park their accumulators in the p slots of warps 0 and 1");
  const Nl2Geom geo = nl2_geom<XL, H, ADJ, W>(a.adim, a.dv);
  const int ldw = nl_ldw(a.dv);
  float* wk_s = reinterpret_cast<float*>(nl2_smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Wp r;
  r.init(nl2_smem + geo.wk_bytes + (size_t)warp * geo.warp_bytes, geo, lane);
  float* pst = reinterpret_cast<float*>(nl2_smem + geo.pst_off);
  float* pacc1 = reinterpret_cast<float*>(nl2_smem + geo.wk_bytes + geo.off_p);
  float* pacc2 = reinterpret_cast<float*>(nl2_smem + geo.wk_bytes + geo.warp_bytes + geo.off_p);
  nl2_stage_wk(a, wk_s, ldw);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const float alpha = load_alpha(a.sp);
  const float beta = a.sp.beta_train != nullptr ? __ldg(a.sp.beta_train) : 0.f;
  const int nrows = a.row_hi - a.row_lo;
  // hub rows first, one CTA each: they finish long before the scheduler loop below runs dry
  for (int li = blockIdx.x; li < a.n_long; li += gridDim.x) {
    const int row = __ldg(a.long_list + li);
    if (row < a.row_lo || row >= a.row_hi) continue;
    if (__ldg(a.rowptr + row + 1) - __ldg(a.rowptr + row) <= a.long_threshold) continue;   // a warp takes it below
    nl2_long_row<XL, H, ADJ, W>(a, r, row, wk_s, ldw, pst, pacc1, pacc2, alpha, beta);
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
      pcols = (pr < nch && lane < min(EB, pend - poff)) ? ld_stream_i32(a.colidx + poff + lane) : 0;
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
      r.setup_row(a, row, lane, wk_s, ldw, alpha);
      r.reset();
#pragma unroll 1
      for (int base = start; base < end; base += EB) {
        produce();
        r.batch(a, base, min(EB, end - base), lane);
      }
      NlStats<H> st;
      r.warp_merge(st);
      r.emit(a, row, lane, st, alpha, beta, wk_s, ldw);
      if constexpr (ADJ) {
        __syncwarp();
        nl_sweep_ds<H>(a, start, end, lane, st, 0, 32);
      }
      __syncwarp();                                          // q / p / u slots are reused by the next row
    }
  }
}

// GB_ERR_UNSUPPORTED: the shape does not fit the shared memory of W warps
template <int XL, int H, bool ADJ, int W>
int nl2_launch(NlArgs a, int n_long, cudaStream_t stream) {
  static bool configured = false;
  const Nl2Geom geo = nl2_geom<XL, H, ADJ, W>(a.adim, a.dv);
  if ((size_t)geo.total > kNlSmemLimit) return GB_ERR_UNSUPPORTED;
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(nl2_rows_kernel<XL, H, ADJ, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNlSmemLimit));
    configured = true;
  }
  const int nrows = a.row_hi - a.row_lo;
  if (nrows <= 0) return GB_OK;
  a.n_long = n_long;
  GB_CUDA(cudaMemsetAsync(a.counter, 0, sizeof(int32_t), stream));
  // rows per scheduler grab: up to kNl2ChunkRows, fewer on small graphs so that every warp of the machine gets work
  const int warps = W * persistent_sms();
  int chunk = nrows / (4 * warps);
  chunk = chunk < 1 ? 1 : (chunk > kNl2ChunkRows ? kNl2ChunkRows : chunk);
  const int want = (nrows + W * chunk - 1) / (W * chunk);
  const int blocks = min_int(want, persistent_sms());
  {
    LaunchScope ls(ADJ ? K_NL_ADJ_ROWS : K_NL_FWD, stream);
    nl2_rows_kernel<XL, H, ADJ, W><<<blocks, W * 32, geo.total, stream>>>(a, chunk);
  }
  return check_launch("nl2_rows_kernel");
}

// GB_ERR_UNSUPPORTED = shape not covered here (the caller uses the first-generation kernels)
template <bool ADJ>
int nl2_dispatch(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
#define GB_NL2_CASE(XLV, HV)                                                           \
  if (xl == XLV && heads == HV) {                                                      \
    const int rc = nl2_launch<XLV, HV, ADJ, 16>(a, n_long, stream);                    \
    return rc != GB_ERR_UNSUPPORTED ? rc : nl2_launch<XLV, HV, ADJ, 8>(a, n_long, stream); \
  }
  const int xl = a.dv <= 16 ? 16 : 32;
  GB_NL2_CASE(32, 4) GB_NL2_CASE(32, 8) GB_NL2_CASE(32, 2) GB_NL2_CASE(32, 1)
  GB_NL2_CASE(16, 4) GB_NL2_CASE(16, 8) GB_NL2_CASE(16, 2) GB_NL2_CASE(16, 1)
#undef GB_NL2_CASE
  return GB_ERR_UNSUPPORTED;
}

}  // namespace gb
