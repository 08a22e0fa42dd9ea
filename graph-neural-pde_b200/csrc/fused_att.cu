// fused_att.cu - warp-per-row fused kernels of the GRAND-nl evaluation (ODEFuncTransformerAtt,
// src/function_transformer_attention.py:44-59) for the default grouping attention_norm_idx = 0.
//
//   att_row_fwd:  one warp owns one softmax segment (a CSR row, or a CSC column for norm_idx 1):
//     phase 1  gather the other endpoint's projection rows (LPE lanes per edge, 32/LPE edges per
//              instruction), per-head dot products by segmented butterflies, raw scores -> shared memory
//     phase 2  lane-parallel over (edge, head) pairs: exact max, sum of exp, normalise
//              (utils.softmax op order, src/utils.py:116-127), head mean
//     phase 3  (FUSE) the SpMM  sum_e abar_e x[col_e,:]  with the diffusion epilogue
//              alpha (Ax - x) + beta x0 (+ Runge-Kutta stage combination)
//   so one evaluation is the projection GEMM plus ONE graph kernel; att / abar are written to HBM only
//   when a backward pass will need them.  Segments longer than the shared-memory capacity (256
//   edges) are left to the CTA-per-segment kernels of attention.cu / spmm.cu.
//
// Roofline: HBM.  Algorithmic bytes per fused forward launch: 8NA + 4E + 12Nd + 4N (+ 4EH + 4E when
// att/abar are stored).
#include "att_common.cuh"
#include "spmm_common.cuh"
#include "nl_kernels.cuh"

namespace gb {
int nl_fill_args(NlArgs& a, const gb_graph* g, int row_lo, int row_hi, const float* x, int ldx, const float* q,
                 int ldq, const float* wk, int ldwk, int heads, int dk, int d, const float* x0, int ldx0,
                 const float* alpha_train, int alpha_sigmoid, const float* beta_train, float* out, int ldo,
                 const gb_stage* stage, void* workspace, const char* who);


constexpr int kFusedWarps = 4;
constexpr int kFusedThreads = kFusedWarps * 32;
constexpr int kFusedCap = 256;          // max edges per segment handled here (== plan long threshold)
constexpr int kColChunk = 64;           // att_col_bwd_kernel stages this many edges at a time: 6 KB of shared memory per
                                        // CTA instead of 24 KB, so occupancy is set by registers (12 CTAs/SM), not by smem
constexpr int kColCap = 256;            // gb_nl_adjoint: longer columns go to the CTA-per-column kernels (measured: a single warp
                                        // on a 4096-edge column costs more than it saves on the C4 and RMAT shapes)

struct FusedFwdArgs {
  AttArgs att;            // segment structure, projections, outputs (att / abar / prods may be null)
  SpmmArgs sp;            // x, epilogue (used when FUSE)
  int32_t lh;             // lanes per head = dk / V (power of two)
};

__device__ __forceinline__ float head_sum_rt(float v, int lh) {
  for (int o = lh >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int V, int LPE, int XL, int XCPL, bool FUSE>
__global__ void __launch_bounds__(kFusedThreads) att_row_fwd_kernel(const FusedFwdArgs fa) {
  extern __shared__ __align__(16) float smem[];
  const AttArgs& a = fa.att;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seg = blockIdx.x * kFusedWarps + warp;
  if (seg >= a.n) return;
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  const int deg = end - start;
  if (deg > kFusedCap) return;                       // CTA-per-segment kernels take these
  const int H = a.heads;
  float* sS = smem + (size_t)warp * kFusedCap * (H + 2);     // [deg][H] raw scores -> attention
  float* sA = sS + kFusedCap * H;                             // [deg] head-mean attention
  int* sC = reinterpret_cast<int*>(sA + kFusedCap);           // [deg] other endpoint

  constexpr int EPI = 32 / LPE;                      // edges per instruction in phase 1
  const int el = lane / LPE, cl = lane % LPE;
  const int width = H * a.dk;
  const bool active = cl * V < width;
  const int my_head = active ? (cl * V) / a.dk : 0;
  const bool leader = active && ((cl * V) % a.dk == 0);
  const float inv_sqrt_dk = 1.0f / sqrtf((float)a.dk);
  float ov2 = 1.f, inv_2l2 = 0.f;
  if (a.att_type == GB_ATT_EXP_KERNEL) {
    const float ov = __ldg(a.kparams), ls = __ldg(a.kparams + 1);
    ov2 = ov * ov;
    inv_2l2 = 1.0f / (2.0f * ls * ls);
  }
  const int lh = fa.lh;
  const Vec<V> qo = active ? Vec<V>::load(a.own + (size_t)seg * a.ldown + cl * V) : Vec<V>::zero();

  // ---- phase 1 ------------------------------------------------------------------------------
  constexpr int U = 4;
  for (int base = 0; base < deg; base += EPI * U) {
    Vec<V> kg[U];
    int slot[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int e = base + u * EPI + el;
      kg[u] = Vec<V>::zero();
      slot[u] = -1;
      if (e < deg) {
        const int other = ld_stream_i32(a.idx + start + e);
        slot[u] = a.perm ? ld_stream_i32(a.perm + start + e) : start + e;
        if (active) kg[u] = Vec<V>::load(a.oth + (size_t)other * a.ldoth + cl * V);
        if (cl == 0) sC[e] = other;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (base + u * EPI >= deg) break;              // warp-uniform
      float s = edge_score_hs<V>(a.att_type, a.dk, qo, kg[u], [lh](float v) { return head_sum_rt(v, lh); },
                                 inv_sqrt_dk, ov2, inv_2l2);
      if (slot[u] >= 0 && leader) {
        if (a.edge_w) s *= __ldg(a.edge_w + slot[u]);
        const int e = base + u * EPI + el;
        sS[e * H + my_head] = s;
        if (a.prods) a.prods[(size_t)slot[u] * H + my_head] = s;
      }
    }
  }
  __syncwarp();

  // ---- phase 2: lanes over (edge, head) pairs; 32 % H == 0 so a lane always sees the same head ---
  const int pairs = deg * H;
  float m = -INFINITY;
  for (int p = lane; p < pairs; p += 32) m = fmaxf(m, sS[p]);
  for (int o = H; o < 32; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float l = 0.f;
  for (int p = lane; p < pairs; p += 32) l += expf(sS[p] - m);
  for (int o = H; o < 32; o <<= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
  const float den = l + 1e-16f;
  const float hf = (float)H;
  const int head2 = lane % H;
  for (int p0 = 0; p0 < pairs; p0 += 32) {
    const int p = p0 + lane;
    const bool ok = p < pairs;
    const float av = ok ? expf(sS[p] - m) / den : 0.f;
    float tot = av;
    for (int o = 1; o < H; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (ok) {
      const int e = p / H;
      const int slot = a.perm ? __ldg(a.perm + start + e) : start + e;
      if (a.att) a.att[(size_t)slot * H + head2] = av;
      if (head2 == 0) {
        const float ab = tot / hf;
        sA[e] = ab;
        if (a.abar) a.abar[slot] = ab;
      }
    }
  }
  if (!FUSE) return;
  __syncwarp();

  // ---- phase 3: out[seg,:] = alpha (sum_e abar_e x[col_e,:] - x[seg,:]) + beta x0[seg,:] ---------
  const SpmmArgs& sp = fa.sp;
  constexpr int EPX = 32 / XL;
  const int xl = lane % XL, xg = lane / XL;
  Vec<V> acc[XCPL];
#pragma unroll
  for (int j = 0; j < XCPL; ++j) acc[j] = Vec<V>::zero();
  constexpr int UX = XCPL == 1 ? 4 : 2;
  for (int base = 0; base < deg; base += EPX * UX) {
    float w[UX];
    Vec<V> r[UX][XCPL];
#pragma unroll
    for (int u = 0; u < UX; ++u) {
      const int e = base + u * EPX + xg;
      const bool ok = e < deg;
      const int c = ok ? sC[e] : 0;
      w[u] = ok ? sA[e] : 0.f;
#pragma unroll
      for (int j = 0; j < XCPL; ++j) {
        const int ch = xl + j * XL;
        r[u][j] = (ok && ch < sp.dv) ? Vec<V>::load(sp.src + (size_t)c * sp.lds + ch * V) : Vec<V>::zero();
      }
    }
#pragma unroll
    for (int u = 0; u < UX; ++u)
#pragma unroll
      for (int j = 0; j < XCPL; ++j) acc[j].fma(w[u], r[u][j]);
  }
  // fixed-order combine of the EPX edge sub-groups
#pragma unroll
  for (int j = 0; j < XCPL; ++j)
    for (int o = XL; o < 32; o <<= 1)
      acc[j] = Vec<V>::map2(acc[j], acc[j], [o](float p, float) { return p + __shfl_xor_sync(0xffffffffu, p, o); });
  if (xg == 0) {
    const float alpha = load_alpha(sp);
    const float beta = sp.beta_train != nullptr ? __ldg(sp.beta_train) : 0.f;
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < XCPL; ++j) {
      const int ch = xl + j * XL;
      if (ch < sp.dv) row_epilogue<V, false>(sp, seg, ch * V, acc[j], alpha, beta, s1, s2);
    }
  }
}


// ================================ backward (norm_idx 0) =========================================
struct FusedBwdArgs {
  // graph
  const int32_t* rowptr; const int32_t* colidx;            // CSR
  const int32_t* colptr; const int32_t* rowidx; const int32_t* tperm;   // CSC
  int32_t n, heads, dk, lh;
  // saved forward tensors
  const float* q; int32_t ldq;
  const float* k; int32_t ldk;
  const float* att;        // [E,H]
  const float* abar;       // [E]
  // incoming gradient and state
  const float* g; int32_t ldg;
  const float* x; int32_t ldx;
  int32_t xdv;             // vector chunks per state row
  int32_t col_cap;         // columns up to this many edges are taken by att_col_bwd_kernel
  // outputs
  float* ds;               // [E,H] scratch: dL/d(raw score), CSR order
  float* dq; int32_t lddq;
  float* dk_out; int32_t lddk;
  SpmmArgs sp;             // transpose-SpMM epilogue (dx, partials, x, x0, alpha)
};

// one warp per CSR row: SDDMM -> softmax backward -> dq
template <int V, int LPE, int XL, int XCPL>
__global__ void __launch_bounds__(kFusedThreads) att_row_bwd_kernel(const FusedBwdArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seg = blockIdx.x * kFusedWarps + warp;
  if (seg >= a.n) return;
  const int start = __ldg(a.rowptr + seg), end = __ldg(a.rowptr + seg + 1);
  const int deg = end - start;
  if (deg > kFusedCap) return;
  const int H = a.heads;
  float* sS = smem + (size_t)warp * kFusedCap * (H + 2);     // ds
  float* sD = sS + kFusedCap * H;                             // dabar
  int* sC = reinterpret_cast<int*>(sD + kFusedCap);
  float alpha = __ldg(a.sp.alpha_train);
  if (a.sp.alpha_sigmoid) alpha = sigmoidf_(alpha);

  // ---- pass A: dabar_e = alpha <g_i, x[col_e]> ----------------------------------------------
  {
    constexpr int EPX = 32 / XL;
    const int xl = lane % XL, xg = lane / XL;
    Vec<V> gi[XCPL];
#pragma unroll
    for (int j = 0; j < XCPL; ++j) {
      const int ch = xl + j * XL;
      gi[j] = ch < a.xdv ? Vec<V>::load(a.g + (size_t)seg * a.ldg + ch * V) : Vec<V>::zero();
    }
    constexpr int UX = XCPL == 1 ? 4 : 2;
    for (int base = 0; base < deg; base += EPX * UX) {
      float part[UX];
#pragma unroll
      for (int u = 0; u < UX; ++u) {
        const int e = base + u * EPX + xg;
        part[u] = 0.f;
        if (e < deg) {
          const int c = ld_stream_i32(a.colidx + start + e);
          if (xl == 0) sC[e] = c;
#pragma unroll
          for (int j = 0; j < XCPL; ++j) {
            const int ch = xl + j * XL;
            if (ch < a.xdv) part[u] += gi[j].dot(Vec<V>::load(a.x + (size_t)c * a.ldx + ch * V));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UX; ++u) {
        float tot = part[u];
#pragma unroll
        for (int o = XL / 2; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        const int e = base + u * EPX + xg;
        if (xl == 0 && e < deg) sD[e] = alpha * tot;
      }
    }
  }
  __syncwarp();

  // ---- pass B: ds[e,h] = a (da - sum_g a da), da = dabar / H -----------------------------------
  {
    const int pairs = deg * H;
    const float inv_h = 1.0f / (float)H;
    float dot = 0.f;
    for (int p = lane; p < pairs; p += 32)
      dot = fmaf(__ldg(a.att + (size_t)start * H + p), sD[p / H] * inv_h, dot);
    for (int o = H; o < 32; o <<= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    for (int p = lane; p < pairs; p += 32) {
      const float v = __ldg(a.att + (size_t)start * H + p) * (sD[p / H] * inv_h - dot);
      sS[p] = v;
      a.ds[(size_t)start * H + p] = v;
    }
  }
  __syncwarp();

  // ---- pass C: dq[i,h,:] = scale * sum_e ds[e,h] k[col_e,h,:] -----------------------------------
  {
    constexpr int EPI = 32 / LPE;
    const int el = lane / LPE, cl = lane % LPE;
    const int width = H * a.dk;
    const bool active = cl * V < width;
    const int my_head = active ? (cl * V) / a.dk : 0;
    Vec<V> acc = Vec<V>::zero();
    constexpr int U = 4;
    for (int base = 0; base < deg; base += EPI * U) {
      float w[U];
      Vec<V> r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int e = base + u * EPI + el;
        const bool ok = e < deg && active;
        w[u] = ok ? sS[e * H + my_head] : 0.f;
        r[u] = ok ? Vec<V>::load(a.k + (size_t)sC[e] * a.ldk + cl * V) : Vec<V>::zero();
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc.fma(w[u], r[u]);
    }
    for (int o = LPE; o < 32; o <<= 1)
      acc = Vec<V>::map2(acc, acc, [o](float p, float) { return p + __shfl_xor_sync(0xffffffffu, p, o); });
    if (el == 0 && active) {
      const float scale = 1.0f / sqrtf((float)a.dk);
      acc = Vec<V>::map2(acc, acc, [scale](float p, float) { return p * scale; });
      acc.store(a.dq + (size_t)seg * a.lddq + cl * V);
    }
  }
}

// one warp per CSC column: transpose SpMM (dx, dalpha/dbeta partials) + dk.  The column's per-edge scalars are
// staged through shared memory in chunks of kColChunk edges; the accumulators live in registers across the
// chunks, so columns up to `a.col_cap` edges stay in this kernel (longer ones: CTA-per-column kernels).
template <int V, int LPE, int XL, int XCPL>
__global__ void __launch_bounds__(kFusedThreads) att_col_bwd_kernel(const FusedBwdArgs a) {
  extern __shared__ __align__(16) float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seg = blockIdx.x * kFusedWarps + warp;
  const int H = a.heads;
  float s1 = 0.f, s2 = 0.f;
  int start = 0, deg = -1;
  if (seg < a.n) {
    start = __ldg(a.colptr + seg);
    deg = __ldg(a.colptr + seg + 1) - start;
    if (deg > a.col_cap) deg = -1;                    // long column: handled elsewhere
  }
  if (deg >= 0) {
    float* sS = smem + (size_t)warp * kColChunk * (H + 2);   // ds of the chunk's edges (16-byte aligned rows)
    float* sA = sS + kColChunk * H;                           // abar
    int* sC = reinterpret_cast<int*>(sA + kColChunk);         // row of each edge
    const SpmmArgs& sp = a.sp;
    constexpr int EPX = 32 / XL;
    const int xl = lane % XL, xg = lane / XL;
    constexpr int EPI = 32 / LPE;
    const int el = lane / LPE, cl = lane % LPE;
    const int width = H * a.dk;
    const bool active = cl * V < width;
    const int my_head = active ? (cl * V) / a.dk : 0;
    Vec<V> acc[XCPL];
#pragma unroll
    for (int j = 0; j < XCPL; ++j) acc[j] = Vec<V>::zero();
    Vec<V> acck = Vec<V>::zero();
    const bool ds4 = H == 4 && (reinterpret_cast<uintptr_t>(a.ds) & 15) == 0;
    for (int c0 = 0; c0 < deg; c0 += kColChunk) {
      const int cn = min(kColChunk, deg - c0);
      // ---- stage per-edge scalars (coalesced over the column's CSC slots) -------------------------
      for (int e = lane; e < cn; e += 32) {
        const int slot = ld_stream_i32(a.tperm + start + c0 + e);
        sC[e] = ld_stream_i32(a.rowidx + start + c0 + e);
        sA[e] = __ldg(a.abar + slot);
        if (ds4) {                                      // one 16-byte load per edge
          *reinterpret_cast<float4*>(sS + e * 4) = __ldg(reinterpret_cast<const float4*>(a.ds + (size_t)slot * 4));
        } else {
          for (int h = 0; h < H; ++h) sS[e * H + h] = __ldg(a.ds + (size_t)slot * H + h);
        }
      }
      __syncwarp();
      // ---- t += sum_e abar_e g[row_e,:] ----------------------------------------------------------
      constexpr int UX = XCPL == 1 ? 4 : 2;
      for (int base = 0; base < cn; base += EPX * UX) {
        float w[UX];
        Vec<V> r[UX][XCPL];
#pragma unroll
        for (int u = 0; u < UX; ++u) {
          const int e = base + u * EPX + xg;
          const bool ok = e < cn;
          const int c = ok ? sC[e] : 0;
          w[u] = ok ? sA[e] : 0.f;
#pragma unroll
          for (int j = 0; j < XCPL; ++j) {
            const int ch = xl + j * XL;
            r[u][j] = (ok && ch < sp.dv) ? Vec<V>::load(sp.src + (size_t)c * sp.lds + ch * V) : Vec<V>::zero();
          }
        }
#pragma unroll
        for (int u = 0; u < UX; ++u)
#pragma unroll
          for (int j = 0; j < XCPL; ++j) acc[j].fma(w[u], r[u][j]);
      }
      // ---- dk[j,h,:] += sum_e ds[e,h] q[row_e,h,:] -------------------------------------------------
      constexpr int U = 4;
      for (int base = 0; base < cn; base += EPI * U) {
        float w[U];
        Vec<V> r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int e = base + u * EPI + el;
          const bool ok = e < cn && active;
          w[u] = ok ? sS[e * H + my_head] : 0.f;
          r[u] = ok ? Vec<V>::load(a.q + (size_t)sC[e] * a.ldq + cl * V) : Vec<V>::zero();
        }
#pragma unroll
        for (int u = 0; u < U; ++u) acck.fma(w[u], r[u]);
      }
      __syncwarp();                                   // the next chunk overwrites the staging buffers
    }
    // fixed-order combine of the edge sub-groups, epilogues
#pragma unroll
    for (int j = 0; j < XCPL; ++j)
      for (int o = XL; o < 32; o <<= 1)
        acc[j] = Vec<V>::map2(acc[j], acc[j], [o](float p, float) { return p + __shfl_xor_sync(0xffffffffu, p, o); });
    if (xg == 0) {
      const float alpha = load_alpha(sp);
#pragma unroll
      for (int j = 0; j < XCPL; ++j) {
        const int ch = xl + j * XL;
        if (ch < sp.dv) row_epilogue<V, true>(sp, seg, ch * V, acc[j], alpha, 0.f, s1, s2);
      }
    }
    for (int o = LPE; o < 32; o <<= 1)
      acck = Vec<V>::map2(acck, acck, [o](float p, float) { return p + __shfl_xor_sync(0xffffffffu, p, o); });
    if (el == 0 && active) {
      const float scale = 1.0f / sqrtf((float)a.dk);
      acck = Vec<V>::map2(acck, acck, [scale](float p, float) { return p * scale; });
      acck.store(a.dk_out + (size_t)seg * a.lddk + cl * V);
    }
  }
  write_partials(a.sp, s1, s2, a.sp.partial_base + blockIdx.x);
}

}  // namespace gb

using namespace gb;

namespace {

inline bool pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

// can the warp-per-row kernels take this problem?  (else the caller uses the general kernels)
int fused_supported(const void* q, int ldq, const void* k, int ldk, int heads, int dk, const char** why) {
  const int width = heads * dk;
  if (!(pow2(heads) && heads <= 8)) { *why = "heads must be a power of two <= 8"; return 0; }
  if (dk % 4 != 0 || !pow2(dk / 4)) { *why = "dk/4 must be a power of two"; return 0; }
  if (width / 4 > 32) { *why = "attention_dim > 128"; return 0; }
  if (vec_width(q, ldq, width) != 4 || vec_width(k, ldk, width) != 4) { *why = "q/k not 16-byte aligned"; return 0; }
  return 1;
}

inline size_t fused_smem(int heads) { return (size_t)kFusedWarps * kFusedCap * (heads + 2) * sizeof(float); }

template <int LPE, int XL, int XCPL, bool FUSE>
int launch_row_fwd(const FusedFwdArgs& fa, cudaStream_t stream) {
  const int blocks = (fa.att.n + kFusedWarps - 1) / kFusedWarps;
  LaunchScope ls(FUSE ? K_FUSED_ATT_SPMM_FWD : K_EDGE_SOFTMAX_FWD, stream);
  att_row_fwd_kernel<4, LPE, XL, XCPL, FUSE><<<blocks, kFusedThreads, fused_smem(fa.att.heads), stream>>>(fa);
  return check_launch("att_row_fwd");
}

template <int LPE, bool FUSE>
int dispatch_row_fwd_x(const FusedFwdArgs& fa, cudaStream_t stream) {
  if (!FUSE) return launch_row_fwd<LPE, 32, 1, false>(fa, stream);
  const RowShape rs = row_shape(fa.sp.dv);
  if (rs.lpr == 8 && rs.cpl == 1) return launch_row_fwd<LPE, 8, 1, FUSE>(fa, stream);
  if (rs.lpr == 16) return launch_row_fwd<LPE, 16, 1, FUSE>(fa, stream);
  if (rs.lpr == 8 && rs.cpl == 3) return launch_row_fwd<LPE, 8, 3, FUSE>(fa, stream);
  if (rs.cpl == 1) return launch_row_fwd<LPE, 32, 1, FUSE>(fa, stream);
  if (rs.cpl == 2) return launch_row_fwd<LPE, 32, 2, FUSE>(fa, stream);
  if (rs.cpl == 3) return launch_row_fwd<LPE, 32, 3, FUSE>(fa, stream);
  return launch_row_fwd<LPE, 32, 4, FUSE>(fa, stream);
}

template <bool FUSE>
int dispatch_row_fwd(const FusedFwdArgs& fa, cudaStream_t stream) {
  switch (lpr_for(fa.att.heads * fa.att.dk / 4)) {
    case 8: return dispatch_row_fwd_x<8, FUSE>(fa, stream);
    case 16: return dispatch_row_fwd_x<16, FUSE>(fa, stream);
    default: return dispatch_row_fwd_x<32, FUSE>(fa, stream);
  }
}

template <int LPE, int XL, int XCPL>
int launch_bwd(const FusedBwdArgs& a, cudaStream_t stream, bool col) {
  const int blocks = (a.n + kFusedWarps - 1) / kFusedWarps;
  LaunchScope ls(col ? K_FUSED_ATT_BWD_COL : K_FUSED_ATT_BWD_ROW, stream);
  if (col)
    att_col_bwd_kernel<4, LPE, XL, XCPL><<<blocks, kFusedThreads,
                                          (size_t)kFusedWarps * kColChunk * (a.heads + 2) * sizeof(float), stream>>>(a);
  else
    att_row_bwd_kernel<4, LPE, XL, XCPL><<<blocks, kFusedThreads, fused_smem(a.heads), stream>>>(a);
  return check_launch(col ? "att_col_bwd" : "att_row_bwd");
}

template <int LPE>
int dispatch_bwd_x(const FusedBwdArgs& a, cudaStream_t stream, bool col) {
  const RowShape rs = row_shape(a.xdv);
  if (rs.lpr == 8 && rs.cpl == 1) return launch_bwd<LPE, 8, 1>(a, stream, col);
  if (rs.lpr == 16) return launch_bwd<LPE, 16, 1>(a, stream, col);
  if (rs.lpr == 8 && rs.cpl == 3) return launch_bwd<LPE, 8, 3>(a, stream, col);
  if (rs.cpl == 1) return launch_bwd<LPE, 32, 1>(a, stream, col);
  if (rs.cpl == 2) return launch_bwd<LPE, 32, 2>(a, stream, col);
  if (rs.cpl == 3) return launch_bwd<LPE, 32, 3>(a, stream, col);
  return launch_bwd<LPE, 32, 4>(a, stream, col);
}

int dispatch_bwd(const FusedBwdArgs& a, cudaStream_t stream, bool col) {
  switch (lpr_for(a.heads * a.dk / 4)) {
    case 8: return dispatch_bwd_x<8>(a, stream, col);
    case 16: return dispatch_bwd_x<16>(a, stream, col);
    default: return dispatch_bwd_x<32>(a, stream, col);
  }
}

}  // namespace

extern "C" {

int gb_att_fused_supported(const float* q, int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk,
                           const float* x, int32_t ldx, int32_t d) {
  const char* why = "";
  if (!fused_supported(q, ldq, k, ldk, heads, dk, &why)) return 0;
  if (x != nullptr && (vec_width(x, ldx, d) != 4 || d > 512)) return 0;
  return 1;
}

int gb_att_fwd_fused(const gb_graph* g, int32_t norm_idx, const float* q, int32_t ldq, const float* k, int32_t ldk,
                     int32_t heads, int32_t dk, int32_t att_type, const float* edge_w, const float* kparams,
                     float* att, float* abar, float* prods, int32_t store,
                     const float* x, int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train,
                     int32_t alpha_sigmoid, const float* beta_train, int32_t d, float* out, int32_t ldo,
                     const gb_stage* stage, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(g && g->n > 0 && g->rowptr && g->colptr, "gb_att_fwd_fused: bad graph");
  GB_REQUIRE(q && k && heads > 0 && dk > 0, "gb_att_fwd_fused: bad arguments");
  GB_REQUIRE(norm_idx == 0 || norm_idx == 1, "gb_att_fwd_fused: attention_norm_idx must be 0 or 1");
  GB_REQUIRE(att_type >= GB_ATT_SCALED_DOT && att_type <= GB_ATT_EXP_KERNEL, "gb_att_fwd_fused: bad attention type");
  GB_REQUIRE(att_type != GB_ATT_EXP_KERNEL || kparams, "gb_att_fwd_fused: exp_kernel needs kparams");
  const bool fuse = x != nullptr;
  GB_REQUIRE(!fuse || norm_idx == 0, "gb_att_fwd_fused: the fused SpMM needs attention_norm_idx == 0");
  GB_REQUIRE(fuse || (att && abar), "gb_att_fwd_fused: attention-only mode needs att and abar outputs");
  GB_REQUIRE(!store || (att && abar), "gb_att_fwd_fused: store requested without att/abar buffers");
  const char* why = "";
  if (!fused_supported(q, ldq, k, ldk, heads, dk, &why)) {
    set_error("gb_att_fwd_fused: unsupported shape (%s)", why);
    return GB_ERR_UNSUPPORTED;
  }
  const bool use_long = g->long_threshold > 0;
  GB_REQUIRE(!use_long || g->long_threshold == kFusedCap, "gb_att_fwd_fused: plan long_threshold must be %d", kFusedCap);
  GB_REQUIRE(use_long || g->e <= kFusedCap, "gb_att_fwd_fused: plan without long-segment lists");
  const int n_long = use_long ? (norm_idx == 0 ? g->n_long_rows : g->n_long_cols) : 0;
  GB_REQUIRE(n_long == 0 || (att && abar), "gb_att_fwd_fused: long segments need att/abar scratch buffers");
  FusedFwdArgs fa{};
  AttArgs& a = fa.att;
  a.n = g->n; a.heads = heads; a.dk = dk; a.att_type = att_type; a.long_threshold = kFusedCap;
  a.edge_w = edge_w; a.kparams = kparams; a.prods = prods;
  a.att = (store || !fuse) ? att : nullptr;
  a.abar = (store || !fuse) ? abar : nullptr;
  if (norm_idx == 0) {
    a.ptr = g->rowptr; a.idx = g->colidx; a.perm = nullptr; a.own = q; a.ldown = ldq; a.oth = k; a.ldoth = ldk;
    a.long_list = g->long_rows;
  } else {
    a.ptr = g->colptr; a.idx = g->rowidx; a.perm = g->tperm; a.own = k; a.ldown = ldk; a.oth = q; a.ldoth = ldq;
    a.long_list = g->long_cols;
  }
  fa.lh = dk / 4;
  if (g->e == 0 && !fuse) return GB_OK;
  SpmmArgs& sp = fa.sp;
  if (fuse) {
    GB_REQUIRE(out && alpha_train && d > 0 && ldx >= d && ldo >= d, "gb_att_fwd_fused: bad SpMM arguments");
    const bool source = x0 != nullptr && beta_train != nullptr;
    int v = min_int(vec_width(x, ldx, d), vec_width(out, ldo, d));
    if (source) v = min_int(v, vec_width(x0, ldx0, d));
    if (stage) {
      GB_REQUIRE(stage->y && stage->ynext && stage->n_prev >= 0 && stage->n_prev <= 6 && stage->ld >= d,
                 "gb_att_fwd_fused: bad stage descriptor");
      v = min_int(v, min_int(vec_width(stage->y, stage->ld, d), vec_width(stage->ynext, stage->ld, d)));
      for (int j = 0; j < stage->n_prev; ++j) v = min_int(v, vec_width(stage->k[j], stage->ld, d));
    }
    if (v != 4 || d > 512) {
      set_error("gb_att_fwd_fused: state rows must be 16-byte aligned with d %% 4 == 0 and d <= 512");
      return GB_ERR_UNSUPPORTED;
    }
    sp.ptr = g->rowptr; sp.idx = g->colidx; sp.perm = nullptr; sp.vals = abar; sp.n = g->n; sp.dv = d / 4;
    sp.long_threshold = kFusedCap; sp.src = x; sp.lds = ldx; sp.alpha_train = alpha_train;
    sp.alpha_sigmoid = alpha_sigmoid; sp.x0 = source ? x0 : nullptr; sp.ldx0 = ldx0;
    sp.beta_train = source ? beta_train : nullptr; sp.out = out; sp.ldo = ldo; sp.long_list = g->long_rows;
    if (stage) { sp.has_stage = 1; sp.stage = *stage; }
    if (int rc = dispatch_row_fwd<true>(fa, stream)) return rc;
  } else {
    if (int rc = dispatch_row_fwd<false>(fa, stream)) return rc;
  }
  if (n_long > 0) {
    AttArgs la = a;
    la.att = att; la.abar = abar;                       // scratch / outputs for the long segments
    if (int rc = att_long_fwd(la, 4, n_long, stream)) return rc;
    if (fuse)
      if (int rc = spmm_long_only(sp, 4, false, n_long, stream)) return rc;
  }
  return GB_OK;
}

size_t gb_att_bwd_fused_partials(const gb_graph* g) {
  if (!g || g->n <= 0) return 0;
  const size_t slots = (size_t)(g->n + kFusedWarps - 1) / kFusedWarps + (size_t)g->n_long_cols;
  return slots * 2 + kMidFloats + 64;
}

int gb_att_bwd_fused(const gb_graph* g, const float* q, int32_t ldq, const float* k, int32_t ldk, int32_t heads,
                     int32_t dk, const float* att, const float* abar, const float* gout, int32_t ldg,
                     const float* x, int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train,
                     int32_t alpha_sigmoid, int32_t d, float* ds, float* dabar_scratch, float* dq, int32_t lddq,
                     float* dk_out, int32_t lddk, float* dx, int32_t lddx, float* partials, float* dparams,
                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(g && g->n > 0 && g->rowptr && g->colptr, "gb_att_bwd_fused: bad graph");
  GB_REQUIRE(q && k && att && abar && gout && x && alpha_train && ds && dq && dk_out && dx && partials,
             "gb_att_bwd_fused: null pointer");
  const char* why = "";
  if (!fused_supported(q, ldq, k, ldk, heads, dk, &why) || vec_width(dq, lddq, heads * dk) != 4 ||
      vec_width(dk_out, lddk, heads * dk) != 4) {
    set_error("gb_att_bwd_fused: unsupported shape (%s)", why);
    return GB_ERR_UNSUPPORTED;
  }
  int v = min_int(min_int(vec_width(gout, ldg, d), vec_width(x, ldx, d)), vec_width(dx, lddx, d));
  if (x0) v = min_int(v, vec_width(x0, ldx0, d));
  if (v != 4 || d > 512) {
    set_error("gb_att_bwd_fused: state rows must be 16-byte aligned with d %% 4 == 0 and d <= 512");
    return GB_ERR_UNSUPPORTED;
  }
  GB_REQUIRE(reinterpret_cast<uintptr_t>(partials) % 8 == 0, "gb_att_bwd_fused: partials must be 8-byte aligned");
  const bool use_long = g->long_threshold > 0;
  GB_REQUIRE(!use_long || g->long_threshold == kFusedCap, "gb_att_bwd_fused: plan long_threshold must be %d", kFusedCap);
  GB_REQUIRE(use_long || g->e <= kFusedCap, "gb_att_bwd_fused: plan without long-segment lists");
  const int nlr = use_long ? g->n_long_rows : 0, nlc = use_long ? g->n_long_cols : 0;
  GB_REQUIRE(nlr == 0 || dabar_scratch, "gb_att_bwd_fused: long rows need the dabar scratch buffer");

  FusedBwdArgs a{};
  a.rowptr = g->rowptr; a.colidx = g->colidx; a.colptr = g->colptr; a.rowidx = g->rowidx; a.tperm = g->tperm;
  a.n = g->n; a.heads = heads; a.dk = dk; a.lh = dk / 4;
  a.q = q; a.ldq = ldq; a.k = k; a.ldk = ldk; a.att = att; a.abar = abar;
  a.g = gout; a.ldg = ldg; a.x = x; a.ldx = ldx; a.xdv = d / 4;
  a.ds = ds; a.dq = dq; a.lddq = lddq; a.dk_out = dk_out; a.lddk = lddk; a.col_cap = kFusedCap;
  SpmmArgs& sp = a.sp;
  sp.ptr = g->colptr; sp.idx = g->rowidx; sp.perm = g->tperm; sp.vals = abar; sp.n = g->n; sp.dv = d / 4;
  sp.long_threshold = kFusedCap; sp.src = gout; sp.lds = ldg; sp.alpha_train = alpha_train;
  sp.alpha_sigmoid = alpha_sigmoid; sp.x0 = dparams ? x0 : nullptr; sp.ldx0 = ldx0;
  sp.x = dparams ? x : nullptr; sp.ldx = ldx; sp.out = dx; sp.ldo = lddx; sp.accumulate = 0;
  sp.partials = partials + kMidFloats; sp.partial_base = 0; sp.long_list = g->long_cols;

  // rows: SDDMM -> softmax backward -> dq
  if (int rc = dispatch_bwd(a, stream, false)) return rc;
  if (nlr > 0) {
    SddmmArgs sd{};
    sd.ptr = g->rowptr; sd.idx = g->colidx; sd.own = gout; sd.ldown = ldg; sd.oth = x; sd.ldoth = ldx; sd.n = g->n;
    sd.dv = d / 4; sd.long_threshold = kFusedCap; sd.accumulate = 0; sd.alpha_train = alpha_train;
    sd.alpha_sigmoid = alpha_sigmoid; sd.dabar = dabar_scratch; sd.long_list = g->long_rows;
    if (int rc = sddmm_long(sd, 4, nlr, stream)) return rc;
    SmBwdArgs sm{};
    sm.ptr = g->rowptr; sm.perm = nullptr; sm.att = att; sm.dabar = dabar_scratch; sm.datt = nullptr;
    sm.edge_w = nullptr; sm.ds = ds; sm.n = g->n; sm.heads = heads; sm.long_threshold = kFusedCap;
    sm.long_list = g->long_rows;
    if (int rc = softmax_bwd_long(sm, nlr, stream)) return rc;
    MhArgs mh{};
    mh.ptr = g->rowptr; mh.idx = g->colidx; mh.perm = nullptr; mh.ds = ds; mh.src = k; mh.lds = ldk; mh.out = dq;
    mh.ldo = lddq; mh.n = g->n; mh.heads = heads; mh.dk = dk; mh.long_threshold = kFusedCap;
    mh.scale = 1.0f / sqrtf((float)dk); mh.long_list = g->long_rows;
    if (int rc = mh_spmm_long(mh, 4, nlr, K_EDGE_SCORE_BWD_Q, stream)) return rc;
  }
  // columns: transpose SpMM (dx, partial sums) + dk
  if (int rc = dispatch_bwd(a, stream, true)) return rc;
  int slots = (g->n + kFusedWarps - 1) / kFusedWarps;
  if (nlc > 0) {
    SpmmArgs lsp = sp;
    lsp.partial_base = slots;
    if (int rc = spmm_long_only(lsp, 4, true, nlc, stream)) return rc;
    slots += nlc;
    MhArgs mh{};
    mh.ptr = g->colptr; mh.idx = g->rowidx; mh.perm = g->tperm; mh.ds = ds; mh.src = q; mh.lds = ldq;
    mh.out = dk_out; mh.ldo = lddk; mh.n = g->n; mh.heads = heads; mh.dk = dk; mh.long_threshold = kFusedCap;
    mh.scale = 1.0f / sqrtf((float)dk); mh.long_list = g->long_cols;
    if (int rc = mh_spmm_long(mh, 4, nlc, K_EDGE_SCORE_BWD_K, stream)) return rc;
  }
  if (dparams) return reduce_partials_launch(partials, slots, alpha_train, alpha_sigmoid, dparams, stream);
  return GB_OK;
}

}  // extern "C"

namespace gb {
// Column pass of the GRAND-nl VJP, shared by gb_nl_adjoint_phase (single gather) and gb_nlk_adjoint_phase (k gather):
// transpose SpMM (dx, dalpha / dbeta partial sums) + dk over the edges this rank holds, from the ds [E,H] / abar [E]
// the row pass left behind.
int nl_cols_pass(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
                 int32_t ldq, int32_t heads, int32_t dk, int32_t d, const float* x0, int32_t ldx0,
                 const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, const float* gout,
                 int32_t ldg, float* ds, float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk, float* dx,
                 int32_t lddx, float* partials, float* dparams, cudaStream_t stream) {
  const bool use_long = g->long_threshold > 0;
  // columns: transpose SpMM (dx, dalpha / dbeta partial sums) + dk over the edges this rank holds
  const bool part = !(row_lo == 0 && row_hi == g->n);
  FusedBwdArgs b{};
  b.rowptr = g->rowptr; b.colidx = g->colidx; b.colptr = g->colptr; b.rowidx = g->rowidx; b.tperm = g->tperm;
  b.n = g->n; b.heads = heads; b.dk = dk; b.lh = dk / 4;
  b.q = q; b.ldq = ldq; b.k = nullptr; b.ldk = 0; b.att = nullptr; b.abar = abar;
  b.g = gout; b.ldg = ldg; b.x = x; b.ldx = ldx; b.xdv = d / 4;
  b.ds = ds; b.dq = dq; b.lddq = lddq; b.dk_out = dk_out; b.lddk = lddk; b.col_cap = kColCap;
  SpmmArgs& sp = b.sp;
  sp.ptr = g->colptr; sp.idx = g->rowidx; sp.perm = g->tperm; sp.vals = abar; sp.n = g->n; sp.dv = d / 4;
  sp.long_threshold = kFusedCap; sp.src = gout; sp.lds = ldg; sp.alpha_train = alpha_train;
  sp.alpha_sigmoid = alpha_sigmoid; sp.x0 = (dparams && beta_train) ? x0 : nullptr; sp.ldx0 = ldx0;
  sp.x = dparams ? x : nullptr; sp.ldx = ldx; sp.out = dx; sp.ldo = lddx; sp.accumulate = 0;
  sp.partials = partials + kMidFloats; sp.partial_base = 0; sp.long_list = g->long_cols;
  // (0, 0) means "own every column" to row_epilogue: a rank WITHOUT rows (row_lo == row_hi, possibly both 0 under a
  // compact halo) must own none - an empty range that can never test as (0, 0)
  if (part) { sp.part_lo = row_lo; sp.part_hi = row_hi > row_lo ? row_hi : -1; }
  const char* why = "";
  if (!fused_supported(q, ldq, q, ldq, heads, dk, &why)) {
    set_error("gb_nl_adjoint: column pass does not support this shape (%s)", why);
    return GB_ERR_UNSUPPORTED;
  }
  int slots = (g->n + kFusedWarps - 1) / kFusedWarps;
  const int nlc = use_long ? g->n_long_cols : 0;
  const cudaStream_t side = nlc > 0 ? side_fork(stream) : stream;   // hub columns next to the main column pass
  if (nlc > 0) {
    SpmmArgs lsp = sp;
    lsp.partial_base = slots;
    lsp.long_threshold = kColCap;                         // shorter listed columns are done by the main kernel
    if (int rc = spmm_long_only(lsp, 4, true, nlc, side)) return rc;
    slots += nlc;
    MhArgs mh{};
    mh.ptr = g->colptr; mh.idx = g->rowidx; mh.perm = g->tperm; mh.ds = ds; mh.src = q; mh.lds = ldq;
    mh.out = dk_out; mh.ldo = lddk; mh.n = g->n; mh.heads = heads; mh.dk = dk; mh.long_threshold = kColCap;
    mh.scale = 1.0f / sqrtf((float)dk); mh.long_list = g->long_cols;
    if (int rc = mh_spmm_long(mh, 4, nlc, K_EDGE_SCORE_BWD_K, side)) return rc;
  }
  if (int rc = dispatch_bwd(b, stream, true)) return rc;
  if (int rc = side_join(stream, side)) return rc;
  if (dparams) return reduce_partials_launch(partials, slots, alpha_train, alpha_sigmoid, dparams, stream);
  return GB_OK;
}
}  // namespace gb

extern "C" {

int gb_nl_adjoint(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
                  int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk, int32_t d, const float* x0,
                  int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train,
                  const float* gout, int32_t ldg, float* out, int32_t ldo, float* ds, float* dabar_scratch,
                  float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk, float* dx, int32_t lddx,
                  float* partials, float* dparams, void* workspace, void* stream_) {
  return gb_nl_adjoint_phase(g, row_lo, row_hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train,
                             alpha_sigmoid, beta_train, gout, ldg, out, ldo, ds, dabar_scratch, abar, dq, lddq, dk_out,
                             lddk, dx, lddx, partials, dparams, workspace, GB_NL_PHASE_BOTH, stream_);
}

int gb_nl_adjoint_phase(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
                        int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk, int32_t d,
                        const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                        const float* beta_train, const float* gout, int32_t ldg, float* out, int32_t ldo, float* ds,
                        float* dabar_scratch, float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk,
                        float* dx, int32_t lddx, float* partials, float* dparams, void* workspace, int32_t phase,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(phase == GB_NL_PHASE_BOTH || phase == GB_NL_PHASE_ROWS || phase == GB_NL_PHASE_COLS,
             "gb_nl_adjoint_phase: bad phase");
  NlArgs a;
  if (int rc = nl_fill_args(a, g, row_lo, row_hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train,
                            alpha_sigmoid, beta_train, out, ldo, nullptr, workspace, "gb_nl_adjoint"))
    return rc;
  GB_REQUIRE(gout && (g->e == 0 || (ds && dabar_scratch && abar)) && dq && dk_out && dx && partials, "gb_nl_adjoint: null pointer");
  GB_REQUIRE(g->colptr && (g->e == 0 || (g->rowidx && g->tperm)), "gb_nl_adjoint: graph without CSC arrays");
  GB_REQUIRE(ldg >= d && lddx >= d && vec_width(gout, ldg, d) == 4 && vec_width(dx, lddx, d) == 4,
             "gb_nl_adjoint: gout / dx rows must be 16-byte aligned");
  const int adim = heads * dk;
  GB_REQUIRE(lddq >= adim && lddk >= adim && vec_width(dk_out, lddk, adim) == 4 && vec_width(q, ldq, adim) == 4,
             "gb_nl_adjoint: q / dk rows must be 16-byte aligned");
  GB_REQUIRE(reinterpret_cast<uintptr_t>(partials) % 8 == 0, "gb_nl_adjoint: partials must be 8-byte aligned");
  const bool use_long = g->long_threshold > 0;
  GB_REQUIRE(!use_long || g->long_threshold == kFusedCap, "gb_nl_adjoint: plan long_threshold must be %d", kFusedCap);
  GB_REQUIRE(use_long || g->e <= kFusedCap, "gb_nl_adjoint: plan without long-segment lists");
  a.write_f = out != nullptr;
  a.g = gout; a.ldg = ldg; a.ds = ds; a.dabar = dabar_scratch; a.abar = abar; a.dq = dq; a.lddq = lddq;
  // rows: evaluation + SDDMM + softmax backward + dq in one gather pass
  if (phase != GB_NL_PHASE_COLS) {
    if (int rc = nl_dispatch_adj(a, heads, use_long ? g->n_long_rows : 0, stream)) return rc;
  }
  if (phase == GB_NL_PHASE_ROWS) return GB_OK;

  return nl_cols_pass(g, row_lo, row_hi, x, ldx, q, ldq, heads, dk, d, x0, ldx0, alpha_train, alpha_sigmoid, beta_train,
                      gout, ldg, ds, abar, dq, lddq, dk_out, lddk, dx, lddx, partials, dparams, stream);
}

}  // extern "C"
