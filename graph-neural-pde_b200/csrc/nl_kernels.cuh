// nl_kernels.cuh - single-gather GRAND-nl kernels (ODEFuncTransformerAtt.forward,
// src/function_transformer_attention.py:44-59, scaled_dot scores, attention_norm_idx 0) and their VJP.
//
// Why these exist: on graphs without locality the evaluation is bound by the per-edge row gathers
// (ncu, profiles/r01c: 49 GB of DRAM reads per forward launch on the ogbn-products shape = E' x (A + d) x 4 B).
// They gather ONE row per edge - x[col] - instead of k[col] and x[col]:
//
//   score[e,h] = <q[i,h,:], k[j,h,:]> / sqrt(dk),   k[j] = Wk x[j] + bk
//              = <p[i,h,:], x[j,:]> + const(i,h),   p[i,h,:] = Wk_h^T q[i,h,:] / sqrt(dk)        (d-vector per head)
//   and the softmax over the edges of row i is invariant to const(i,h), so the K projection is never formed.
//
//   forward   one pass over the row's edges with an ONLINE softmax per head (running max m_h, running sum l_h,
//             running sum T1_h = sum_e exp(s_eh - m_h) x_j):   (A x)_i = 1/H sum_h T1_h / l_h
//   adjoint   the same pass also carries dabar_e = alpha <g_i, x_j> and T2_h = sum_e exp(s_eh - m_h) dabar_e x_j,
//             D_h = sum_e exp(s_eh - m_h) dabar_e, which give without a second gather
//                 dot_h = D_h / (H l_h)                     (= sum_e att_eh dabar_e / H)
//                 u_h   = T2_h / (H l_h) - dot_h T1_h / l_h (= sum_e ds_eh x_j)
//                 dq[i,h,:] = Wk_h u_h / sqrt(dk)
//             raw scores / dabar are parked per edge and turned into ds[e,h] = att (dabar/H - dot_h) and
//             abar[e] in a coalesced second sweep over the row's own slots (no gather).
//
// Work decomposition: persistent CTAs of 8 warps (GB_NL_WARPS_FWD / _ADJ), Wk staged once per CTA in shared memory, rows handed out
// by a dynamic (integer atomic) scheduler in chunks; a warp owns a row; XL lanes span the d columns
// (128-bit chunks), 32/XL edge groups run side by side, each lane keeps U gathers in flight.  Per-head dot
// products are reduced with a transposing butterfly (31 shuffles for 32 values).  Rows longer than the plan's
// long threshold get one CTA each (warps stride the batches; partial online-softmax states merged in a
// fixed order).  No float atomics: results are bit-reproducible run to run.
//
// Roofline: HBM.  Algorithmic bytes per launch (perfect reuse of gathered rows):
//   forward  4E + 4N + 4NA (q) + 12Nd (x, x0, out)
//   adjoint  4E + 4N + 8NA (q, dq) + 8Nd (x, g) [+ 8Nd when f is emitted] + 20E (ds [E,H] + abar [E] out;
//            the parked scores are re-read from L2)
// Gather model (no reuse): + 4 (E - N) d.
#pragma once
#include <math_constants.h>
#include "spmm_common.cuh"

namespace gb {

// GB_NL_* are the compile-time tuning knobs of these kernels; csrc/Makefile's `variant` target builds a library per
// setting (scripts/sweep_variants.py runs them side by side on the GPU).  The defaults are the measured optimum.
#ifndef GB_NL_WARPS
#define GB_NL_WARPS 8
#endif
#ifndef GB_NL_WARPS_FWD
#define GB_NL_WARPS_FWD GB_NL_WARPS
#endif
#ifndef GB_NL_WARPS_ADJ
#define GB_NL_WARPS_ADJ GB_NL_WARPS
#endif
#ifndef GB_NL_CHUNK_ROWS
#define GB_NL_CHUNK_ROWS 4
#endif
template <bool ADJ> constexpr int kNlWarpsOf = ADJ ? GB_NL_WARPS_ADJ : GB_NL_WARPS_FWD;   // warps per CTA
constexpr int kNlChunkRows = GB_NL_CHUNK_ROWS;   // rows per scheduler grab (<= 31: one lane per row pointer)
static_assert(kNlChunkRows >= 1 && kNlChunkRows <= 31, "a chunk's row pointers are loaded by one warp instruction");
static_assert(kNlWarpsOf<false> >= 4 && kNlWarpsOf<false> <= 32 && kNlWarpsOf<true> >= 4 && kNlWarpsOf<true> <= 32,
              "CTA-wide hub rows park [W][3][H] statistics in the q slots of warps 1.. ((W-1) A floats, dk >= 4)");
// Tuning knobs (measured on the ogbn-products shape, profiles/r01zi, r01zj): the forward kernel keeps 4 gathers in
// flight per lane group and fits 80 registers -> 3 CTAs (24 warps) per SM, 3 % faster than 8 gathers at 2 CTAs;
// the adjoint kernel carries twice the accumulators: 8 gathers at 2 CTAs (125 registers); 4 gathers at 3 CTAs
// spills and is 12 % slower.
#ifndef GB_NL_UFWD
#define GB_NL_UFWD 4
#endif
#ifndef GB_NL_CTAS_FWD
#define GB_NL_CTAS_FWD 3
#endif
#ifndef GB_NL_UADJ
#define GB_NL_UADJ 8
#endif
#ifndef GB_NL_CTAS_ADJ
#define GB_NL_CTAS_ADJ 2
#endif
constexpr int kNlUFwd = GB_NL_UFWD, kNlUAdj = GB_NL_UADJ;              // gathers in flight per lane group
constexpr int kNlCtasFwd = GB_NL_CTAS_FWD, kNlCtasAdj = GB_NL_CTAS_ADJ;  // resident CTAs per SM compiled for
constexpr int kNlLongRow = 2048;         // rows longer than this get a whole CTA (the online softmax needs no
                                         // per-edge shared memory, so a warp can take rows far beyond the plan's
                                         // 256-edge list threshold; the scheduler balances them)
constexpr unsigned kFullMask = 0xffffffffu;
constexpr size_t kNlSmemLimit = 227 * 1024;   // opt-in dynamic shared memory per CTA on sm_100

struct NlArgs {
  const int32_t* rowptr;
  const int32_t* colidx;
  int32_t row_lo, row_hi;    // rows evaluated by this launch (1-D row partition; [0, n) on one GPU)
  int32_t dv;                // d / 4
  int32_t adim, dk;          // attention_dim = heads * dk
  const float* q;            // [*, A] indexed by the global row id
  int32_t ldq;
  const float* wk;           // [A, d] (torch Linear.weight layout)
  int32_t ldwk;
  SpmmArgs sp;               // x (src/lds), diffusion epilogue, optional Runge-Kutta stage
  int32_t write_f;           // adjoint: also emit f (the adjoint solver needs it)
  // adjoint only
  const float* g;            // dL/df, indexed by the global row id
  int32_t ldg;
  float* ds;                 // [E,H]  out: dL/d raw score (parked raw scores in between)
  float* dabar;              // [E]    scratch: alpha <g_i, x_j>
  float* abar;               // [E]    out: mean_h att
  float* dq;                 // [*, A] out
  int32_t lddq;
  int32_t* counter;          // scheduler counter, zeroed before the launch
  const int32_t* long_list;  // rows with more edges than the plan's list threshold (256)
  int32_t n_long;
  int32_t long_threshold;    // rows with more edges than this are worked by a whole CTA
};

constexpr int ilog2_c(int v) { return v <= 1 ? 0 : 1 + ilog2_c(v / 2); }

__device__ __forceinline__ float4 f4_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// Packed fp32 (sm_100: fma/mul.rn.f32x2 -> FFMA2 / FMUL2).  The row kernels are bound by instruction issue,
// not by the fma pipe: one FFMA2 takes one issue slot for two FMAs, and its scalar-broadcast operand form
// (R.F32) makes "scalar x pair" free.  Results are the same IEEE fma per component as the scalar code.
__device__ __forceinline__ unsigned long long pk2(float2 v) { return *reinterpret_cast<unsigned long long*>(&v); }
__device__ __forceinline__ float2 unpk2(unsigned long long r) { return *reinterpret_cast<float2*>(&r); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk2(a)), "l"(pk2(b)), "l"(pk2(c)));
  return unpk2(r);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a)), "l"(pk2(b)));
  return unpk2(r);
}
__device__ __forceinline__ float2 bc2(float s) { return make_float2(s, s); }
// acc += s * v, two FFMA2
__device__ __forceinline__ void fma4p(float4& acc, float s, const float4& v) {
  const float2 lo = fma2(bc2(s), make_float2(v.x, v.y), make_float2(acc.x, acc.y));
  const float2 hi = fma2(bc2(s), make_float2(v.z, v.w), make_float2(acc.z, acc.w));
  acc = make_float4(lo.x, lo.y, hi.x, hi.y);
}
// <a, b> over 4 columns: FMUL2 + FFMA2 + FADD
__device__ __forceinline__ float dot4p(const float4& a, const float4& b) {
  const float2 t = fma2(make_float2(a.z, a.w), make_float2(b.z, b.w), mul2(make_float2(a.x, a.y), make_float2(b.x, b.y)));
  return t.x + t.y;
}
__device__ __forceinline__ void f4_scale(float4& a, float s) {
  const float2 lo = mul2(bc2(s), make_float2(a.x, a.y)), hi = mul2(bc2(s), make_float2(a.z, a.w));
  a = make_float4(lo.x, lo.y, hi.x, hi.y);
}
__device__ __forceinline__ float4 f4_shfl_xor(const float4& a, int o) {
  return make_float4(__shfl_xor_sync(kFullMask, a.x, o), __shfl_xor_sync(kFullMask, a.y, o),
                     __shfl_xor_sync(kFullMask, a.z, o), __shfl_xor_sync(kFullMask, a.w, o));
}
__device__ __forceinline__ void f4_add(float4& a, const float4& b) {
  a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
}

// Transposing butterfly: every lane of an XL-lane group holds NV partial sums (NV <= XL, powers of two);
// on return v[0] of lane xl is the group total of value index (xl >> (log2 XL - log2 NV)).
// O = first xor offset (XL/2).  Costs NV-1 + (log2 XL - log2 NV) shuffles instead of NV log2 XL.
template <int NV, int O>
struct ReduceT {
  static __device__ __forceinline__ void run(float* v, int xl) {
    if constexpr (O >= 1) {
      if constexpr (NV > 1) {
        constexpr int HN = NV / 2;
        const bool up = (xl & O) != 0;
#pragma unroll
        for (int i = 0; i < HN; ++i) {
          const float send = up ? v[i] : v[i + HN];
          const float keep = up ? v[i + HN] : v[i];
          v[i] = keep + __shfl_xor_sync(kFullMask, send, O);
        }
        ReduceT<HN, O / 2>::run(v, xl);
      } else {
        v[0] += __shfl_xor_sync(kFullMask, v[0], O);
        ReduceT<1, O / 2>::run(v, xl);
      }
    }
  }
};

template <int H>
struct NlStats {
  float mh[H], lh[H], dh[H];     // per head: running max, sum of exp, sum of exp * dabar
};

// Register state of one warp working on (part of) one row.
template <int XL, int H, bool ADJ>
struct NlRow {
  static constexpr int EPI = 32 / XL;                       // edge groups per warp
  static constexpr int UMAX = ADJ ? kNlUAdj : kNlUFwd;
  static constexpr int U = (XL / H) < UMAX ? (XL / H) : UMAX;   // edges in flight per group
  static constexpr int NV = U * H;                          // scores per group batch (<= XL)
  static constexpr int SH = ilog2_c(XL) - ilog2_c(NV);      // score index of lane xl = xl >> SH
  static constexpr int SHU = ilog2_c(XL) - ilog2_c(U);
  static constexpr int BATCH = EPI * U;                     // edges per warp batch
  static constexpr int HA = ADJ ? H : 1;
  static_assert(H <= XL && NV <= XL && 32 % H == 0, "unsupported head count");

  float m, l, dsum;          // online softmax state of (this lane's head, this lane's edge slots)
  float4 acc1[H];            // T1_h, this lane's 4 columns
  float4 acc2[HA];           // T2_h
  static constexpr bool PAIRS = (H % 2 == 0);               // heads paired in FFMA2 score products
  float4 p[PAIRS ? 1 : H];   // Wk_h^T q_ih / sqrt(dk), this lane's 4 columns (odd head counts)
  float2 pp[PAIRS ? 2 * H : 1];   // the same as head pairs: pp[(h/2)*4 + c] = (p[h][c], p[h+1][c])
  float4 gi;                 // g_i chunk (adjoint)
  int xl, es, gbase, myh, myu;

  __device__ __forceinline__ void init_lane(int lane) {
    xl = lane % XL;
    es = lane / XL;
    gbase = lane - xl;
    const int idx = xl >> SH;
    myh = idx % H;
    myu = idx / H;
  }

  __device__ __forceinline__ void reset() {
    m = -CUDART_INF_F;
    l = 0.f;
    dsum = 0.f;
#pragma unroll
    for (int h = 0; h < H; ++h) acc1[h] = f4_zero();
#pragma unroll
    for (int h = 0; h < HA; ++h) acc2[h] = f4_zero();
  }

  // p[h] = Wk_h^T q_ih / sqrt(dk); wk_s = staged Wk [A][ldw], q_s = this row's q [A] (shared memory)
  __device__ __forceinline__ void load_p(const NlArgs& a, const float* wk_s, int ldw, const float* q_s) {
    const float scale = 1.0f / sqrtf((float)a.dk);
    const bool act = xl < a.dv;
    const float* wcol = wk_s + 4 * (act ? xl : 0);
    float4 ph[H];
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
      ph[h] = acc;
    }
    const float2 sc2 = bc2(act ? scale : 0.f);
    if constexpr (PAIRS) {
      // the scaling multiply doubles as the transposition to head pairs: its FMUL2 results ARE the register
      // pairs the score products read (a plain repack is re-done by ptxas with two MOVs at every use)
#pragma unroll
      for (int h = 0; h < H; h += 2) {
        pp[2 * h + 0] = mul2(make_float2(ph[h].x, ph[h + 1].x), sc2);
        pp[2 * h + 1] = mul2(make_float2(ph[h].y, ph[h + 1].y), sc2);
        pp[2 * h + 2] = mul2(make_float2(ph[h].z, ph[h + 1].z), sc2);
        pp[2 * h + 3] = mul2(make_float2(ph[h].w, ph[h + 1].w), sc2);
      }
    } else {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        f4_scale(ph[h], sc2.x);
        p[h] = ph[h];
      }
    }
  }

  // column index of edge base + lane (lanes < BATCH), -1 past the end of the row
  __device__ __forceinline__ int load_cols(const NlArgs& a, int base, int end, int lane) const {
    return (lane < BATCH && base + lane < end) ? ld_stream_i32(a.colidx + base + lane) : -1;
  }

  // pull the rows of a coming batch towards the SM: lane -> (edge lane / LPE, 1/LPE of the row), one
  // prefetch instruction per batch.  `cnext` = load_cols() of that batch.
  __device__ __forceinline__ void prefetch_rows(const NlArgs& a, int cnext, int lane) const {
    constexpr int LPE = 32 / BATCH;
    const int c = __shfl_sync(kFullMask, cnext, lane / LPE);
    const int row_bytes = 16 * a.dv;
    const int off = (lane % LPE) * ((row_bytes + LPE - 1) / LPE);
    if (c >= 0 && off < row_bytes) {
      const char* ptr = reinterpret_cast<const char*>(a.sp.src + (size_t)c * a.sp.lds) + off;
      asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr));
    }
  }

  // one warp batch: edges [base, base + BATCH) of a row ending at `end`; `cmine` = load_cols(base), issued by
  // the caller one batch ahead so that the index load is off the critical path of the gather
  __device__ __forceinline__ void batch(const NlArgs& a, int base, int end, int lane, float alpha, int cmine) {
    const bool act = xl < a.dv;
    float4 xr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = __shfl_sync(kFullMask, cmine, es * U + u);
      xr[u] = (c >= 0 && act) ? ldg4(a.sp.src + (size_t)c * a.sp.lds + 4 * xl) : f4_zero();
    }
    float v[NV];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if constexpr (PAIRS) {                                  // two heads per FFMA2, x broadcast
#pragma unroll
        for (int h = 0; h < H; h += 2) {
          float2 s2 = mul2(pp[2 * h + 0], bc2(xr[u].x));
          s2 = fma2(pp[2 * h + 1], bc2(xr[u].y), s2);
          s2 = fma2(pp[2 * h + 2], bc2(xr[u].z), s2);
          s2 = fma2(pp[2 * h + 3], bc2(xr[u].w), s2);
          v[u * H + h] = s2.x;
          v[u * H + h + 1] = s2.y;
        }
      } else {
#pragma unroll
        for (int h = 0; h < H; ++h) v[u * H + h] = dot4p(p[h], xr[u]);
      }
    }
    ReduceT<NV, XL / 2>::run(v, xl);
    const int pos = base + es * U + myu;
    const bool valid = pos < end;
    const float s = valid ? v[0] : -CUDART_INF_F;
    float dab = 0.f;
    if constexpr (ADJ) {
      float w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) w[u] = dot4p(gi, xr[u]);
      ReduceT<U, XL / 2>::run(w, xl);
      dab = alpha * __shfl_sync(kFullMask, w[0], gbase + (myu << SHU));
      if (valid && (xl & ((1 << SH) - 1)) == 0) {           // park the raw score / dabar of this edge
        a.ds[(size_t)pos * H + myh] = s;
        if (myh == 0) a.dabar[pos] = dab;
      }
    }
    float mb = s;
#pragma unroll
    for (int o = H << SH; o < XL; o <<= 1) mb = fmaxf(mb, __shfl_xor_sync(kFullMask, mb, o));
    const float mnew = fmaxf(m, mb);
    const float corr = (mnew == -CUDART_INF_F) ? 1.f : expf(m - mnew);
    const float pe = valid ? expf(s - mnew) : 0.f;
    m = mnew;
    l = fmaf(l, corr, pe);
    if constexpr (ADJ) dsum = fmaf(dsum, corr, pe * dab);
    if (__any_sync(kFullMask, corr != 1.f)) {               // some head's running max moved: rescale
#pragma unroll
      for (int h = 0; h < H; ++h) {
        const float c = __shfl_sync(kFullMask, corr, gbase + (h << SH));
        f4_scale(acc1[h], c);
        if constexpr (ADJ) f4_scale(acc2[h], c);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const float du = ADJ ? __shfl_sync(kFullMask, dab, gbase + ((u * H) << SH)) : 0.f;
#pragma unroll
      for (int h = 0; h < H; ++h) {
        const float pu = __shfl_sync(kFullMask, pe, gbase + ((u * H + h) << SH));
        fma4p(acc1[h], pu, xr[u]);
        if constexpr (ADJ) fma4p(acc2[h], pu * du, xr[u]);
      }
    }
  }

  // combine the lanes / edge groups of the warp: afterwards every lane holds the per-head statistics and
  // (for its 4 columns) the merged accumulators.  Fixed order -> deterministic.
  __device__ __forceinline__ void warp_merge(NlStats<H>& st) {
#pragma unroll
    for (int o = H << SH; o < XL; o <<= 1) {
      l += __shfl_xor_sync(kFullMask, l, o);
      if constexpr (ADJ) dsum += __shfl_xor_sync(kFullMask, dsum, o);
    }
#pragma unroll
    for (int h = 0; h < H; ++h) {
      st.mh[h] = __shfl_sync(kFullMask, m, gbase + (h << SH));
      st.lh[h] = __shfl_sync(kFullMask, l, gbase + (h << SH));
      st.dh[h] = ADJ ? __shfl_sync(kFullMask, dsum, gbase + (h << SH)) : 0.f;
    }
    if (EPI > 1) {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float mx = st.mh[h];
#pragma unroll
        for (int o = XL; o < 32; o <<= 1) mx = fmaxf(mx, __shfl_xor_sync(kFullMask, mx, o));
        const float sc = (st.mh[h] == -CUDART_INF_F) ? 0.f : expf(st.mh[h] - mx);
        st.mh[h] = mx;
        st.lh[h] *= sc;
        st.dh[h] *= sc;
        f4_scale(acc1[h], sc);
        if constexpr (ADJ) f4_scale(acc2[h], sc);
#pragma unroll
        for (int o = XL; o < 32; o <<= 1) {
          st.lh[h] += __shfl_xor_sync(kFullMask, st.lh[h], o);
          st.dh[h] += __shfl_xor_sync(kFullMask, st.dh[h], o);
          f4_add(acc1[h], f4_shfl_xor(acc1[h], o));
          if constexpr (ADJ) f4_add(acc2[h], f4_shfl_xor(acc2[h], o));
        }
      }
    }
  }

  // f (diffusion epilogue) and dq of the row from the merged state; executed by one warp.
  // u_s: [H][4*XL] floats of shared memory private to the warp.
  __device__ __forceinline__ void emit(const NlArgs& a, int row, int lane, const NlStats<H>& st, float alpha,
                                       float beta, float* u_s, const float* wk_s, int ldw) {
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
#pragma unroll
      for (int h = 0; h < H; ++h) {
        const float dot = st.dh[h] * inv[h] * inv_h;
        const float c2 = inv[h] * inv_h, c1 = inv[h] * dot;
        float4 uh;
        uh.x = acc2[h].x * c2 - acc1[h].x * c1;
        uh.y = acc2[h].y * c2 - acc1[h].y * c1;
        uh.z = acc2[h].z * c2 - acc1[h].z * c1;
        uh.w = acc2[h].w * c2 - acc1[h].w * c1;
        if (es == 0) *reinterpret_cast<float4*>(u_s + h * (4 * XL) + 4 * xl) = uh;
      }
      __syncwarp();
      const float scale = 1.0f / sqrtf((float)a.dk);
      for (int j = lane; j < a.adim; j += 32) {              // dq[i, j] = <Wk[j, :], u_{head(j)}> / sqrt(dk)
        const float* wr = wk_s + (size_t)j * ldw;
        const float* ur = u_s + (j / a.dk) * (4 * XL);
        float2 acc = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll 4
        for (int c = 0; c < a.dv; ++c) {
          const float4 wv = *reinterpret_cast<const float4*>(wr + 4 * c), uv = *reinterpret_cast<const float4*>(ur + 4 * c);
          acc = fma2(make_float2(wv.x, wv.y), make_float2(uv.x, uv.y), acc);
          acc1 = fma2(make_float2(wv.z, wv.w), make_float2(uv.z, uv.w), acc1);
        }
        a.dq[(size_t)row * a.lddq + j] = ((acc.x + acc1.x) + (acc.y + acc1.y)) * scale;
      }
    }
  }
};

// second sweep of the adjoint over the row's own slots [first, end) with stride `step` (in (edge, head) pairs,
// multiples of 32): parked raw score / dabar -> ds[e,h] = att (dabar/H - dot_h), abar[e] = mean_h att
template <int H>
__device__ __forceinline__ void nl_sweep_ds(const NlArgs& a, int start, int end, int lane, const NlStats<H>& st,
                                            int first, int step) {
  const int hsel = lane % H;
  float msel = st.mh[0], lsel = st.lh[0], dsel = st.dh[0];
#pragma unroll
  for (int h = 1; h < H; ++h)
    if (hsel == h) { msel = st.mh[h]; lsel = st.lh[h]; dsel = st.dh[h]; }
  const float inv = lsel > 0.f ? 1.0f / lsel : 0.f;
  const float inv_h = 1.0f / (float)H;
  const float dot = dsel * inv * inv_h;
  const int pairs = (end - start) * H;
  float* dsr = a.ds + (size_t)start * H;
  for (int p0 = first; p0 < pairs; p0 += step) {
    const int pp = p0 + lane;
    const bool ok = pp < pairs;
    const int e = pp / H;
    float att = 0.f, dab = 0.f;
    if (ok) {
      att = expf(dsr[pp] - msel) * inv;
      dab = a.dabar[start + e];
    }
    float tot = att;
#pragma unroll
    for (int o = 1; o < H; o <<= 1) tot += __shfl_xor_sync(kFullMask, tot, o);
    if (ok) {
      dsr[pp] = att * (dab * inv_h - dot);
      if (hsel == 0) a.abar[start + e] = tot * inv_h;
    }
  }
}

// shared-memory layout of both kernels: [A][ldw] staged Wk | per-warp q [W][A] | per-warp u [W][H][4*XL]
__host__ __device__ inline int nl_ldw(int dv) { return 4 * dv + 4; }

template <int XL, int H, bool ADJ>
inline size_t nl_smem_rows(int adim, int dv) {
  constexpr int kNlWarps = kNlWarpsOf<ADJ>;
  return sizeof(float) * ((size_t)adim * nl_ldw(dv) + (size_t)kNlWarps * adim + (size_t)kNlWarps * H * 4 * XL);
}
__device__ __forceinline__ void nl_stage_wk(const NlArgs& a, float* wk_s, int ldw) {
  const int total = a.adim * a.dv;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int j = i / a.dv, c = i - j * a.dv;
    *reinterpret_cast<float4*>(wk_s + (size_t)j * ldw + 4 * c) = ldg4(a.wk + (size_t)j * a.ldwk + 4 * c);
  }
}

// A long row (more edges than a.long_threshold) is worked by the whole CTA: warps stride the batches, then the
// partial online-softmax states are merged in warp order.  No extra shared memory: the per-warp statistics go to
// the q slots of warps 1.. ([W][3][H] <= 7 A floats), the accumulators are summed warp after warp into the u slots
// of warps 0-1 ([2][H][4 XL]).  Called by every thread of the CTA.
template <int XL, int H, bool ADJ>
__device__ __forceinline__ void nl_long_row(const NlArgs& a, int row, float* wk_s, int ldw, float alpha, float beta) {
  using Row = NlRow<XL, H, ADJ>;
  constexpr int kNlWarps = kNlWarpsOf<ADJ>;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q_s = wk_s + (size_t)a.adim * ldw;                                    // warp 0's q slot
  float* pst = q_s + a.adim;                                                   // [W][3][H]
  float* u_s = wk_s + (size_t)a.adim * ldw + (size_t)kNlWarps * a.adim;        // warp 0's u slot
  float* pacc = u_s;                                                           // [2][H][4*XL] (u slots of warps 0-1)
  const int start = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
  for (int j = threadIdx.x; j < a.adim; j += blockDim.x) q_s[j] = __ldg(a.q + (size_t)row * a.ldq + j);
  __syncthreads();
  Row r;
  r.init_lane(lane);
  r.load_p(a, wk_s, ldw, q_s);
  if constexpr (ADJ) r.gi = r.xl < a.dv ? ldg4(a.g + (size_t)row * a.ldg + 4 * r.xl) : f4_zero();
  r.reset();
  {
    int cols = r.load_cols(a, start + warp * Row::BATCH, end, lane);
    for (int base = start + warp * Row::BATCH; base < end; base += kNlWarps * Row::BATCH) {
      const int cnext = r.load_cols(a, base + kNlWarps * Row::BATCH, end, lane);
      r.prefetch_rows(a, cnext, lane);
      r.batch(a, base, end, lane, alpha, cols);
      cols = cnext;
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
  // every warp forms the same merged statistics
  float scw[H];                                               // this warp's rescale to the merged maximum
#pragma unroll
  for (int h = 0; h < H; ++h) {
    float mx = -CUDART_INF_F;
    for (int w = 0; w < kNlWarps; ++w) mx = fmaxf(mx, pst[(w * 3 + 0) * H + h]);
    float ls = 0.f, dsum = 0.f;
    for (int w = 0; w < kNlWarps; ++w) {
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
  // accumulators: warp 0 stores, warps 1.. add in turn (fixed order -> deterministic)
  for (int w = 0; w < kNlWarps; ++w) {
    if (warp == w && r.es == 0) {
#pragma unroll
      for (int h = 0; h < H; ++h) {
        float4* d1 = reinterpret_cast<float4*>(pacc + (size_t)h * 4 * XL + 4 * r.xl);
        float4 t1 = w == 0 ? f4_zero() : *d1;
        fma4(t1, scw[h], r.acc1[h]);
        *d1 = t1;
        if constexpr (ADJ) {
          float4* d2 = reinterpret_cast<float4*>(pacc + (size_t)(H + h) * 4 * XL + 4 * r.xl);
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
      r.acc1[h] = *reinterpret_cast<const float4*>(pacc + (size_t)h * 4 * XL + 4 * r.xl);
      if constexpr (ADJ) r.acc2[h] = *reinterpret_cast<const float4*>(pacc + (size_t)(H + h) * 4 * XL + 4 * r.xl);
    }
    __syncwarp();                                            // emit() reuses the same shared memory for u
    r.emit(a, row, lane, st, alpha, beta, u_s, wk_s, ldw);
  }
  if constexpr (ADJ) nl_sweep_ds<H>(a, start, end, lane, st, warp * 32, kNlWarps * 32);
  __syncthreads();                                          // the q / u slots go back to their warps
}

// persistent kernel: warps pull chunks of rows from an atomic counter
template <int XL, int H, bool ADJ>
__global__ void __launch_bounds__(kNlWarpsOf<ADJ> * 32, ADJ ? kNlCtasAdj : kNlCtasFwd) nl_rows_kernel(const NlArgs a) {
  extern __shared__ __align__(16) float smem[];
  using Row = NlRow<XL, H, ADJ>;
  constexpr int kNlWarps = kNlWarpsOf<ADJ>;
  const int ldw = nl_ldw(a.dv);
  float* wk_s = smem;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q_s = wk_s + (size_t)a.adim * ldw + (size_t)warp * a.adim;
  float* u_s = wk_s + (size_t)a.adim * ldw + (size_t)kNlWarps * a.adim + (size_t)warp * H * 4 * XL;
  nl_stage_wk(a, wk_s, ldw);
  __syncthreads();
  Row r;
  r.init_lane(lane);
  const float alpha = load_alpha(a.sp);
  const float beta = a.sp.beta_train != nullptr ? __ldg(a.sp.beta_train) : 0.f;
  const int nrows = a.row_hi - a.row_lo;
  // hub rows first, one CTA each (CTA b takes list entries b, b + grid, ...): they finish long before the
  // scheduler loop below runs dry, so they cost no tail and no second launch
  for (int li = blockIdx.x; li < a.n_long; li += gridDim.x) {
    const int row = __ldg(a.long_list + li);
    if (row < a.row_lo || row >= a.row_hi) continue;
    if (__ldg(a.rowptr + row + 1) - __ldg(a.rowptr + row) <= a.long_threshold) continue;   // a warp takes it below
    nl_long_row<XL, H, ADJ>(a, row, wk_s, ldw, alpha, beta);
  }
  for (;;) {
    int r0 = 0;
    if (lane == 0) r0 = atomicAdd(a.counter, kNlChunkRows);
    r0 = __shfl_sync(kFullMask, r0, 0);
    if (r0 >= nrows) break;
    const int r1 = min(r0 + kNlChunkRows, nrows);
    // the chunk's row pointers in one load (lanes 0 .. kNlChunkRows), handed out by shuffles
    const int rp = (lane <= kNlChunkRows && r0 + lane <= nrows) ? __ldg(a.rowptr + a.row_lo + r0 + lane) : 0;
    for (int rr = r0; rr < r1; ++rr) {
      const int row = a.row_lo + rr;
      const int start = __shfl_sync(kFullMask, rp, rr - r0), end = __shfl_sync(kFullMask, rp, rr - r0 + 1);
      if (end - start > a.long_threshold) continue;          // done CTA-wide above
      int cols = r.load_cols(a, start, end, lane);           // in flight while q is staged and p is formed
      for (int j = lane; j < a.adim; j += 32) q_s[j] = __ldg(a.q + (size_t)row * a.ldq + j);
      __syncwarp();
      r.load_p(a, wk_s, ldw, q_s);
      if constexpr (ADJ)
        r.gi = r.xl < a.dv ? ldg4(a.g + (size_t)row * a.ldg + 4 * r.xl) : f4_zero();
      r.reset();
      int cnext = r.load_cols(a, start + Row::BATCH, end, lane);
      for (int base = start; base < end; base += Row::BATCH) {
        const int cnn = r.load_cols(a, base + 2 * Row::BATCH, end, lane);
        r.prefetch_rows(a, cnext, lane);
        r.batch(a, base, end, lane, alpha, cols);
        cols = cnext;
        cnext = cnn;
      }
      NlStats<H> st;
      r.warp_merge(st);
      r.emit(a, row, lane, st, alpha, beta, u_s, wk_s, ldw);
      if constexpr (ADJ) {
        __syncwarp();
        nl_sweep_ds<H>(a, start, end, lane, st, 0, 32);
      }
      __syncwarp();                                          // q_s / u_s are reused by the next row
    }
  }
}

template <int XL, int H, bool ADJ>
int nl_launch(NlArgs a, int n_long, cudaStream_t stream) {
  static bool configured = false;
  constexpr int kNlWarps = kNlWarpsOf<ADJ>;
  const size_t smem = nl_smem_rows<XL, H, ADJ>(a.adim, a.dv);
  if (smem > kNlSmemLimit) {
    set_error("nl kernels: attention_dim x d too large for shared memory (%zu bytes)", smem);
    return GB_ERR_UNSUPPORTED;
  }
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(nl_rows_kernel<XL, H, ADJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNlSmemLimit));
    configured = true;
  }
  const int nrows = a.row_hi - a.row_lo;
  if (nrows <= 0) return GB_OK;
  a.n_long = n_long;
  GB_CUDA(cudaMemsetAsync(a.counter, 0, sizeof(int32_t), stream));
  const int want = (nrows + kNlWarps * kNlChunkRows - 1) / (kNlWarps * kNlChunkRows);
  const int blocks = min_int(want, (ADJ ? kNlCtasAdj : kNlCtasFwd) * sm_count());
  {
    LaunchScope ls(ADJ ? K_NL_ADJ_ROWS : K_NL_FWD, stream);
    nl_rows_kernel<XL, H, ADJ><<<blocks, kNlWarps * 32, smem, stream>>>(a);
  }
  return check_launch("nl_rows_kernel");
}

template <bool ADJ>
int nl_dispatch(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
#define GB_NL_CASE(XLV, HV) \
  if (xl == XLV && heads == HV) return nl_launch<XLV, HV, ADJ>(a, n_long, stream);
  const int xl = a.dv <= 16 ? 16 : 32;
  GB_NL_CASE(32, 4) GB_NL_CASE(32, 8) GB_NL_CASE(32, 2) GB_NL_CASE(32, 1)
  GB_NL_CASE(16, 4) GB_NL_CASE(16, 8) GB_NL_CASE(16, 2) GB_NL_CASE(16, 1)
#undef GB_NL_CASE
  set_error("nl kernels: unsupported heads=%d", heads);
  return GB_ERR_UNSUPPORTED;
}

// defined in nl_fwd.cu / nl_adj.cu (one translation unit per direction: they compile in parallel)
int nl_dispatch_fwd(const NlArgs& a, int heads, int n_long, cudaStream_t stream);
int nl_dispatch_adj(const NlArgs& a, int heads, int n_long, cudaStream_t stream);

}  // namespace gb
