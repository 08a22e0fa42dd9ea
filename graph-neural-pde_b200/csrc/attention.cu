// attention.cu - K2+K4 (forward) and K5b (backward): per-edge attention scores fused with the segment
// softmax and the head mean, and the deterministic VJP kernels.
//
// Reference semantics (src/function_transformer_attention.py:224-266, src/utils.py:116-127):
//   q = Q(x), k = K(x) viewed [N, H, dk];  s[e,h] = score(q[row_e,h,:], k[col_e,h,:])
//   a[e,h] = exp(s - max_g s) / (sum_g exp(s - max_g s) + 1e-16),  g = edges sharing
//            edge_index[attention_norm_idx][e]  (rows for 0, columns for 1)
//   abar[e] = mean_h a[e,h]                      (function_transformer_attention.py:34)
// The reference runs ~20 library kernels with four [E,dk,H]-sized temporaries and float atomics
// (torch_scatter); here one kernel walks each softmax segment once: a sub-warp group of LPR lanes
// owns a segment, lane l holds V consecutive columns of the A = H*dk wide q/k rows, so every
// gathered row is one coalesced 128-bit-per-lane request and the per-head dot products are
// segmented butterflies inside the group.  The softmax is the exact three-phase form (max, sum,
// normalise) in the reference's op order so that ties give bit-identical results (K3 known answer:
// exactly 0.5).  Raw scores are parked in the output buffer between phases (L1/L2 resident).
//
// Roofline: HBM.  Algorithmic bytes per forward launch: 8NA (q,k read once) + 4E (index) + 4EH + 4E.
#include "att_common.cuh"

namespace gb {

// One softmax segment per LPR-lane group (LONG == false) or per CTA (LONG == true).
template <int V, int LPR, int LH, bool LONG>
__global__ void __launch_bounds__(kAttThreads) edge_softmax_fwd_kernel(const AttArgs a) {
  constexpr int G = kAttThreads / LPR;
  __shared__ float sh_red[LONG ? G * LPR : 1];
  const int gl = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const unsigned mask = group_mask<LPR>();
  const int width = a.heads * a.dk;
  const bool active = gl * V < width;
  const int my_head = active ? (gl * V) / a.dk : -1;
  const bool leader = active && ((gl * V) % a.dk == 0);
  const float inv_sqrt_dk = 1.0f / sqrtf((float)a.dk);
  float ov2 = 1.f, inv_2l2 = 0.f;
  if (a.att_type == GB_ATT_EXP_KERNEL) {
    const float ov = __ldg(a.kparams), ls = __ldg(a.kparams + 1);
    ov2 = ov * ov;
    inv_2l2 = 1.0f / (2.0f * ls * ls);
  }
  int seg;
  if (LONG) {
    seg = __ldg(a.long_list + blockIdx.x);
  } else {
    seg = blockIdx.x * G + grp;
    if (seg >= a.n) return;
  }
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  if (!LONG && end - start > a.long_threshold) return;
  if (LONG && end - start <= a.long_threshold) return;      // listed, but taken by a warp-level kernel (block-uniform)
  const int first = LONG ? start + grp : start;
  const int step = LONG ? G : 1;
  const Vec<V> qo = active ? Vec<V>::load(a.own + (size_t)seg * a.ldown + gl * V) : Vec<V>::zero();

  // ---- phase 1: raw scores (parked in att) and the running maximum ------------------------
  float m = -INFINITY;
  constexpr int U = 2;
  for (int p = first; p < end; p += U * step) {
    int slot[U];
    Vec<V> kg[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int pp = p + u * step;
      slot[u] = -1;
      kg[u] = Vec<V>::zero();
      if (pp < end) {
        const int other = ld_stream_i32(a.idx + pp);
        slot[u] = a.perm ? ld_stream_i32(a.perm + pp) : pp;
        if (active) kg[u] = Vec<V>::load(a.oth + (size_t)other * a.ldoth + gl * V);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (slot[u] < 0) continue;                    // uniform inside the group
      float s = edge_score<V, LPR, LH>(a, qo, kg[u], my_head, mask, inv_sqrt_dk, ov2, inv_2l2);
      if (a.edge_w) s *= __ldg(a.edge_w + slot[u]);
      if (leader) {
        a.att[(size_t)slot[u] * a.heads + my_head] = s;
        if (a.prods) a.prods[(size_t)slot[u] * a.heads + my_head] = s;
      }
      m = fmaxf(m, s);
    }
  }
  if (LONG) {
    sh_red[grp * LPR + gl] = m;
    __syncthreads();
    for (int g = 0; g < G; ++g) m = fmaxf(m, sh_red[g * LPR + gl]);
    __syncthreads();
  } else {
    __syncwarp(mask);
  }

  // ---- phase 2: denominator --------------------------------------------------------------
  float l = 0.f;
  if (active)
    for (int p = first; p < end; p += step) {
      const int slot = a.perm ? ld_stream_i32(a.perm + p) : p;
      l += expf(a.att[(size_t)slot * a.heads + my_head] - m);
    }
  if (LONG) {
    sh_red[grp * LPR + gl] = l;
    __syncthreads();
    l = 0.f;
    for (int g = 0; g < G; ++g) l += sh_red[g * LPR + gl];
  }
  const float den = l + 1e-16f;

  // ---- phase 3: normalise, head mean -----------------------------------------------------
  const int lh = a.dk / V;                          // lanes per head
  const float hf = (float)a.heads;
  for (int p = first; p < end; p += step) {
    const int slot = a.perm ? ld_stream_i32(a.perm + p) : p;
    float av = 0.f;
    if (active) av = expf(a.att[(size_t)slot * a.heads + my_head] - m) / den;
    __syncwarp(mask);                               // all lanes of the head have read the raw score
    if (leader) a.att[(size_t)slot * a.heads + my_head] = av;
    float tot = 0.f;
    for (int h = 0; h < a.heads; ++h) tot += __shfl_sync(mask, av, h * lh, LPR);
    if (gl == 0) a.abar[slot] = tot / hf;
  }
}


// ================================ backward ====================================================

// ---- (1) SDDMM: dabar[s] = alpha * <gout[row_s,:], x[col_s,:]>  over CSR rows ------------------

template <int V, int LPR, int CPL, bool LONG>
__global__ void __launch_bounds__(kAttThreads) sddmm_kernel(const SddmmArgs a) {
  constexpr int G = kAttThreads / LPR;
  const int gl = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const unsigned mask = group_mask<LPR>();
  int row;
  if (LONG) {
    row = __ldg(a.long_list + blockIdx.x);
  } else {
    row = blockIdx.x * G + grp;
    if (row >= a.n) return;
  }
  const int start = __ldg(a.ptr + row), end = __ldg(a.ptr + row + 1);
  if (!LONG && end - start > a.long_threshold) return;
  float alpha = __ldg(a.alpha_train);
  if (a.alpha_sigmoid) alpha = sigmoidf_(alpha);
  Vec<V> g[CPL];
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = gl + j * LPR;
    g[j] = ch < a.dv ? Vec<V>::load(a.own + (size_t)row * a.ldown + ch * V) : Vec<V>::zero();
  }
  const int first = LONG ? start + grp : start;
  const int step = LONG ? G : 1;
  constexpr int U = CPL == 1 ? 4 : 2;
  for (int p = first; p < end; p += U * step) {
    float part[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int pp = p + u * step;
      part[u] = 0.f;
      if (pp < end) {
        const int c = ld_stream_i32(a.idx + pp);
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
          const int ch = gl + j * LPR;
          if (ch < a.dv) part[u] += g[j].dot(Vec<V>::load(a.oth + (size_t)c * a.ldoth + ch * V));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int pp = p + u * step;
      if (pp >= end) continue;
      float tot = part[u];
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) tot += __shfl_xor_sync(mask, tot, o, LPR);
      if (gl == 0) {
        const float v = alpha * tot;
        a.dabar[pp] = a.accumulate ? a.dabar[pp] + v : v;
      }
    }
  }
}

// ---- (2) softmax backward over the softmax segments --------------------------------------------

__device__ __forceinline__ float sm_da(const SmBwdArgs& a, int slot, int h, float inv_h) {
  float v = 0.f;
  if (a.dabar) v = __ldg(a.dabar + slot) * inv_h;
  if (a.datt) v += __ldg(a.datt + (size_t)slot * a.heads + h);
  return v;
}

// one thread per (segment, head); two sequential passes over the segment (fixed order)
__global__ void __launch_bounds__(256) softmax_bwd_kernel(const SmBwdArgs a) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int seg = (int)(t / a.heads), h = (int)(t % a.heads);
  if (seg >= a.n) return;
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  if (end - start > a.long_threshold) return;
  const float inv_h = 1.0f / (float)a.heads;
  float dot = 0.f;
  for (int p = start; p < end; ++p) {
    const int slot = a.perm ? __ldg(a.perm + p) : p;
    dot = fmaf(__ldg(a.att + (size_t)slot * a.heads + h), sm_da(a, slot, h, inv_h), dot);
  }
  for (int p = start; p < end; ++p) {
    const int slot = a.perm ? __ldg(a.perm + p) : p;
    float v = __ldg(a.att + (size_t)slot * a.heads + h) * (sm_da(a, slot, h, inv_h) - dot);
    if (a.edge_w) v *= __ldg(a.edge_w + slot);
    a.ds[(size_t)slot * a.heads + h] = v;
  }
}

// one CTA per long segment
__global__ void __launch_bounds__(256) softmax_bwd_long_kernel(const SmBwdArgs a) {
  __shared__ float sh[256];
  __shared__ float sh_dot[64];
  const int seg = __ldg(a.long_list + blockIdx.x);
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  const int per = 256 / a.heads;                     // edge lanes per head (heads <= 64)
  const int h = threadIdx.x % a.heads, lane = threadIdx.x / a.heads;
  const bool on = lane < per;
  const float inv_h = 1.0f / (float)a.heads;
  float part = 0.f;
  if (on)
    for (int p = start + lane; p < end; p += per) {
      const int slot = a.perm ? __ldg(a.perm + p) : p;
      part = fmaf(__ldg(a.att + (size_t)slot * a.heads + h), sm_da(a, slot, h, inv_h), part);
    }
  sh[threadIdx.x] = on ? part : 0.f;
  __syncthreads();
  if (threadIdx.x < a.heads) {
    float tot = 0.f;
    for (int l = 0; l < per; ++l) tot += sh[l * a.heads + threadIdx.x];
    sh_dot[threadIdx.x] = tot;
  }
  __syncthreads();
  const float dot = sh_dot[h];
  if (on)
    for (int p = start + lane; p < end; p += per) {
      const int slot = a.perm ? __ldg(a.perm + p) : p;
      float v = __ldg(a.att + (size_t)slot * a.heads + h) * (sm_da(a, slot, h, inv_h) - dot);
      if (a.edge_w) v *= __ldg(a.edge_w + slot);
      a.ds[(size_t)slot * a.heads + h] = v;
    }
}

// ---- (3) multi-head SpMM: out[seg, c] = scale * sum_p ds[slot_p, head(c)] * src[idx_p, c] -------
// dq over CSR rows (src = k, perm = nullptr); dk over CSC columns (src = q, perm = tperm).

template <int V, int LPR, bool LONG>
__global__ void __launch_bounds__(kAttThreads) mh_spmm_kernel(const MhArgs a) {
  constexpr int G = kAttThreads / LPR;
  __shared__ __align__(16) float red[LONG ? kAttThreads * V : 4];
  const int gl = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const int width = a.heads * a.dk;
  const bool active = gl * V < width;
  const int my_head = active ? (gl * V) / a.dk : 0;
  int seg;
  if (LONG) {
    seg = __ldg(a.long_list + blockIdx.x);
  } else {
    seg = blockIdx.x * G + grp;
    if (seg >= a.n) return;
  }
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  if (!LONG && end - start > a.long_threshold) return;
  if (LONG && end - start <= a.long_threshold) return;      // listed, but taken by a warp-level kernel (block-uniform)
  const int first = LONG ? start + grp : start;
  const int step = LONG ? G : 1;
  Vec<V> acc = Vec<V>::zero();
  constexpr int U = 4;
  if (active)
    for (int p = first; p < end; p += U * step) {
      float w[U];
      Vec<V> r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int pp = p + u * step;
        w[u] = 0.f;
        r[u] = Vec<V>::zero();
        if (pp < end) {
          const int c = ld_stream_i32(a.idx + pp);
          const int slot = a.perm ? ld_stream_i32(a.perm + pp) : pp;
          w[u] = __ldg(a.ds + (size_t)slot * a.heads + my_head);
          r[u] = Vec<V>::load(a.src + (size_t)c * a.lds + gl * V);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) acc.fma(w[u], r[u]);
    }
  const float scale = a.scale;
  if (LONG) {
    acc.store(red + (size_t)threadIdx.x * V);
    __syncthreads();
    if (grp == 0 && active) {
      Vec<V> tot = Vec<V>::zero();
      for (int g = 0; g < G; ++g) {
        const Vec<V> part = *reinterpret_cast<const Vec<V>*>(red + ((size_t)g * LPR + gl) * V);
        tot = Vec<V>::map2(tot, part, [](float x, float y) { return x + y; });
      }
      tot = Vec<V>::map2(tot, tot, [scale](float x, float) { return x * scale; });
      tot.store(a.out + (size_t)seg * a.ldo + gl * V);
    }
  } else if (active) {
    acc = Vec<V>::map2(acc, acc, [scale](float x, float) { return x * scale; });
    acc.store(a.out + (size_t)seg * a.ldo + gl * V);
  }
}

}  // namespace gb

using namespace gb;

namespace {

int check_graph_att(const gb_graph* g, const char* who) {
  GB_REQUIRE(g != nullptr, "%s: null graph", who);
  GB_REQUIRE(g->n > 0 && g->e >= 0 && g->rowptr && g->colptr, "%s: bad graph", who);
  GB_REQUIRE(g->e == 0 || (g->colidx && g->rowidx && g->tperm), "%s: graph without index arrays", who);
  GB_REQUIRE(g->n_long_rows == 0 || g->long_rows, "%s: long_rows missing", who);
  GB_REQUIRE(g->n_long_cols == 0 || g->long_cols, "%s: long_cols missing", who);
  return GB_OK;
}

inline int eff_threshold(const gb_graph* g) { return g->long_threshold > 0 ? g->long_threshold : 0x7fffffff; }

// vector width usable for [N, heads*dk] matrices whose heads must not straddle a vector
int att_vec(const float* a, int lda, const float* b, int ldb, int width, int dk) {
  int v = min_int(vec_width(a, lda, width), vec_width(b, ldb, width));
  while (v > 1 && dk % v != 0) v /= 2;
  return v;
}

inline bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

template <int V, int LPR, int LH>
int launch_softmax_fwd(const AttArgs& a, int n_long, cudaStream_t stream) {
  constexpr int G = kAttThreads / LPR;
  {
    LaunchScope ls(K_EDGE_SOFTMAX_FWD, stream);
    edge_softmax_fwd_kernel<V, LPR, LH, false><<<(a.n + G - 1) / G, kAttThreads, 0, stream>>>(a);
  }
  if (n_long > 0) {
    LaunchScope ls(K_EDGE_SOFTMAX_FWD_LONG, stream);
    edge_softmax_fwd_kernel<V, LPR, LH, true><<<n_long, kAttThreads, 0, stream>>>(a);
  }
  return check_launch("edge_softmax_fwd");
}

template <int V, int LPR>
int dispatch_softmax_lh(const AttArgs& a, int lh, int n_long, cudaStream_t stream) {
  if (V == 4 && is_pow2(lh) && lh <= 8 && lh <= LPR) {
    switch (lh) {
      case 1: return launch_softmax_fwd<V, LPR, 1>(a, n_long, stream);
      case 2: return launch_softmax_fwd<V, LPR, 2>(a, n_long, stream);
      case 4: return launch_softmax_fwd<V, LPR, 4>(a, n_long, stream);
      case 8: return launch_softmax_fwd<V, LPR, 8>(a, n_long, stream);
    }
  }
  return launch_softmax_fwd<V, LPR, 0>(a, n_long, stream);
}

template <int V>
int dispatch_softmax(const AttArgs& a, int n_long, cudaStream_t stream) {
  const int chunks = a.heads * a.dk / V, lh = a.dk / V;
  switch (lpr_for(chunks)) {
    case 8: return dispatch_softmax_lh<V, 8>(a, lh, n_long, stream);
    case 16: return dispatch_softmax_lh<V, 16>(a, lh, n_long, stream);
    default: return dispatch_softmax_lh<V, 32>(a, lh, n_long, stream);
  }
}

template <int V, int LPR, int CPL>
int launch_sddmm(const SddmmArgs& a, int n_long, cudaStream_t stream) {
  constexpr int G = kAttThreads / LPR;
  LaunchScope ls(K_SDDMM, stream);
  sddmm_kernel<V, LPR, CPL, false><<<(a.n + G - 1) / G, kAttThreads, 0, stream>>>(a);
  if (n_long > 0) sddmm_kernel<V, LPR, CPL, true><<<n_long, kAttThreads, 0, stream>>>(a);
  return check_launch("sddmm");
}

template <int V>
int dispatch_sddmm(const SddmmArgs& a, int n_long, cudaStream_t stream) {
  const RowShape rs = row_shape(a.dv);
  if (rs.lpr == 8 && rs.cpl == 1) return launch_sddmm<V, 8, 1>(a, n_long, stream);
  if (rs.lpr == 16) return launch_sddmm<V, 16, 1>(a, n_long, stream);
  if (rs.lpr == 8 && rs.cpl == 3) return launch_sddmm<V, 8, 3>(a, n_long, stream);
  if (rs.cpl == 1) return launch_sddmm<V, 32, 1>(a, n_long, stream);
  if (rs.cpl == 2) return launch_sddmm<V, 32, 2>(a, n_long, stream);
  if (rs.cpl == 3) return launch_sddmm<V, 32, 3>(a, n_long, stream);
  return launch_sddmm<V, 32, 4>(a, n_long, stream);
}

template <int V, int LPR>
int launch_mh(const MhArgs& a, int n_long, int kid, cudaStream_t stream) {
  constexpr int G = kAttThreads / LPR;
  LaunchScope ls(kid, stream);
  mh_spmm_kernel<V, LPR, false><<<(a.n + G - 1) / G, kAttThreads, 0, stream>>>(a);
  if (n_long > 0) mh_spmm_kernel<V, LPR, true><<<n_long, kAttThreads, 0, stream>>>(a);
  return check_launch("mh_spmm");
}

template <int V>
int dispatch_mh(const MhArgs& a, int n_long, int kid, cudaStream_t stream) {
  switch (lpr_for(a.heads * a.dk / V)) {
    case 8: return launch_mh<V, 8>(a, n_long, kid, stream);
    case 16: return launch_mh<V, 16>(a, n_long, kid, stream);
    default: return launch_mh<V, 32>(a, n_long, kid, stream);
  }
}

}  // namespace


namespace gb {

int att_long_fwd(const AttArgs& a, int v, int n_long, cudaStream_t stream) {
  if (n_long <= 0) return GB_OK;
  const int chunks = a.heads * a.dk / v;
  LaunchScope ls(K_EDGE_SOFTMAX_FWD_LONG, stream);
#define GB_ATT_LONG(VV, LL) edge_softmax_fwd_kernel<VV, LL, 0, true><<<n_long, kAttThreads, 0, stream>>>(a)
  const int lpr = lpr_for(chunks);
  if (v == 4) { if (lpr == 8) GB_ATT_LONG(4, 8); else if (lpr == 16) GB_ATT_LONG(4, 16); else GB_ATT_LONG(4, 32); }
  else if (v == 2) { if (lpr == 8) GB_ATT_LONG(2, 8); else if (lpr == 16) GB_ATT_LONG(2, 16); else GB_ATT_LONG(2, 32); }
  else { if (lpr == 8) GB_ATT_LONG(1, 8); else if (lpr == 16) GB_ATT_LONG(1, 16); else GB_ATT_LONG(1, 32); }
#undef GB_ATT_LONG
  return check_launch("att_long_fwd");
}

template <int V>
static int sddmm_long_v(const SddmmArgs& a, int n_long, cudaStream_t stream) {
  const RowShape rs = row_shape(a.dv);
  if (rs.lpr == 8 && rs.cpl == 1) sddmm_kernel<V, 8, 1, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else if (rs.lpr == 16) sddmm_kernel<V, 16, 1, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else if (rs.lpr == 8 && rs.cpl == 3) sddmm_kernel<V, 8, 3, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else if (rs.cpl == 1) sddmm_kernel<V, 32, 1, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else if (rs.cpl == 2) sddmm_kernel<V, 32, 2, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else if (rs.cpl == 3) sddmm_kernel<V, 32, 3, true><<<n_long, kAttThreads, 0, stream>>>(a);
  else sddmm_kernel<V, 32, 4, true><<<n_long, kAttThreads, 0, stream>>>(a);
  return check_launch("sddmm_long");
}

int sddmm_long(const SddmmArgs& a, int v, int n_long, cudaStream_t stream) {
  if (n_long <= 0) return GB_OK;
  LaunchScope ls(K_SDDMM, stream);
  if (v == 4) return sddmm_long_v<4>(a, n_long, stream);
  if (v == 2) return sddmm_long_v<2>(a, n_long, stream);
  return sddmm_long_v<1>(a, n_long, stream);
}

int softmax_bwd_long(const SmBwdArgs& a, int n_long, cudaStream_t stream) {
  if (n_long <= 0) return GB_OK;
  LaunchScope ls(K_EDGE_SOFTMAX_BWD, stream);
  softmax_bwd_long_kernel<<<n_long, 256, 0, stream>>>(a);
  return check_launch("softmax_bwd_long");
}

int mh_spmm_long(const MhArgs& a, int v, int n_long, int kernel_id, cudaStream_t stream) {
  if (n_long <= 0) return GB_OK;
  LaunchScope ls(kernel_id, stream);
  const int lpr = lpr_for(a.heads * a.dk / v);
#define GB_MH_LONG(VV, LL) mh_spmm_kernel<VV, LL, true><<<n_long, kAttThreads, 0, stream>>>(a)
  if (v == 4) { if (lpr == 8) GB_MH_LONG(4, 8); else if (lpr == 16) GB_MH_LONG(4, 16); else GB_MH_LONG(4, 32); }
  else if (v == 2) { if (lpr == 8) GB_MH_LONG(2, 8); else if (lpr == 16) GB_MH_LONG(2, 16); else GB_MH_LONG(2, 32); }
  else { if (lpr == 8) GB_MH_LONG(1, 8); else if (lpr == 16) GB_MH_LONG(1, 16); else GB_MH_LONG(1, 32); }
#undef GB_MH_LONG
  return check_launch("mh_spmm_long");
}

}  // namespace gb

extern "C" {

int gb_edge_softmax_fwd(const gb_graph* g, int32_t norm_idx, const float* q, int32_t ldq, const float* k, int32_t ldk,
                        int32_t heads, int32_t dk, int32_t att_type, const float* edge_w, const float* kparams,
                        float* att, float* abar, float* prods, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_graph_att(g, "gb_edge_softmax_fwd")) return rc;
  GB_REQUIRE(q && k && att && abar && heads > 0 && dk > 0, "gb_edge_softmax_fwd: bad arguments");
  GB_REQUIRE(norm_idx == 0 || norm_idx == 1, "gb_edge_softmax_fwd: attention_norm_idx must be 0 or 1");
  GB_REQUIRE(att_type >= GB_ATT_SCALED_DOT && att_type <= GB_ATT_EXP_KERNEL, "gb_edge_softmax_fwd: bad attention type");
  GB_REQUIRE(att_type != GB_ATT_EXP_KERNEL || kparams, "gb_edge_softmax_fwd: exp_kernel needs kparams");
  const int width = heads * dk;
  GB_REQUIRE(ldq >= width && ldk >= width, "gb_edge_softmax_fwd: leading dimension too small");
  if (g->e == 0) return GB_OK;
  const int v = att_vec(q, ldq, k, ldk, width, dk);
  if (width / v > 32) {
    set_error("gb_edge_softmax_fwd: attention_dim %d with vector width %d exceeds 32 lanes per row", width, v);
    return GB_ERR_UNSUPPORTED;
  }
  AttArgs a{};
  a.n = g->n; a.heads = heads; a.dk = dk; a.att_type = att_type; a.long_threshold = eff_threshold(g);
  a.edge_w = edge_w; a.kparams = kparams; a.att = att; a.abar = abar; a.prods = prods;
  int n_long;
  if (norm_idx == 0) {
    a.ptr = g->rowptr; a.idx = g->colidx; a.perm = nullptr; a.own = q; a.ldown = ldq; a.oth = k; a.ldoth = ldk;
    a.long_list = g->long_rows; n_long = g->n_long_rows;
  } else {
    a.ptr = g->colptr; a.idx = g->rowidx; a.perm = g->tperm; a.own = k; a.ldown = ldk; a.oth = q; a.ldoth = ldq;
    a.long_list = g->long_cols; n_long = g->n_long_cols;
  }
  if (g->long_threshold <= 0) n_long = 0;
  if (v == 4) return dispatch_softmax<4>(a, n_long, stream);
  if (v == 2) return dispatch_softmax<2>(a, n_long, stream);
  return dispatch_softmax<1>(a, n_long, stream);
}

int gb_sddmm(const gb_graph* g, const float* gout, int32_t ldg, const float* x, int32_t ldx, int32_t d,
             const float* alpha_train, int32_t alpha_sigmoid, float* dabar, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_graph_att(g, "gb_sddmm")) return rc;
  GB_REQUIRE(gout && x && alpha_train && dabar && d > 0, "gb_sddmm: bad arguments");
  GB_REQUIRE(ldg >= d && ldx >= d, "gb_sddmm: leading dimension too small");
  if (g->e == 0) return GB_OK;
  const int v = min_int(vec_width(gout, ldg, d), vec_width(x, ldx, d));
  const int n_long = g->long_threshold > 0 ? g->n_long_rows : 0;
  const int max_cols = 128 * v;
  for (int c0 = 0; c0 < d; c0 += max_cols) {
    SddmmArgs a{};
    a.ptr = g->rowptr; a.idx = g->colidx; a.own = gout + c0; a.ldown = ldg; a.oth = x + c0; a.ldoth = ldx;
    a.n = g->n; a.dv = min_int(max_cols, d - c0) / v; a.long_threshold = eff_threshold(g); a.accumulate = c0 > 0;
    a.alpha_train = alpha_train; a.alpha_sigmoid = alpha_sigmoid; a.dabar = dabar; a.long_list = g->long_rows;
    int rc;
    if (v == 4) rc = dispatch_sddmm<4>(a, n_long, stream);
    else if (v == 2) rc = dispatch_sddmm<2>(a, n_long, stream);
    else rc = dispatch_sddmm<1>(a, n_long, stream);
    if (rc) return rc;
  }
  return GB_OK;
}

int gb_edge_softmax_bwd(const gb_graph* g, int32_t norm_idx, const float* att, const float* dabar, const float* datt,
                        int32_t heads, const float* edge_w, float* ds, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_graph_att(g, "gb_edge_softmax_bwd")) return rc;
  GB_REQUIRE(att && (dabar || datt) && ds && heads > 0 && heads <= 64, "gb_edge_softmax_bwd: bad arguments (heads <= 64)");
  GB_REQUIRE(norm_idx == 0 || norm_idx == 1, "gb_edge_softmax_bwd: attention_norm_idx must be 0 or 1");
  if (g->e == 0) return GB_OK;
  SmBwdArgs a{};
  a.att = att; a.dabar = dabar; a.datt = datt; a.edge_w = edge_w; a.ds = ds; a.n = g->n; a.heads = heads;
  a.long_threshold = eff_threshold(g);
  int n_long;
  if (norm_idx == 0) { a.ptr = g->rowptr; a.perm = nullptr; a.long_list = g->long_rows; n_long = g->n_long_rows; }
  else { a.ptr = g->colptr; a.perm = g->tperm; a.long_list = g->long_cols; n_long = g->n_long_cols; }
  if (g->long_threshold <= 0) n_long = 0;
  const int64_t threads = (int64_t)g->n * heads;
  LaunchScope ls(K_EDGE_SOFTMAX_BWD, stream);
  softmax_bwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(a);
  if (n_long > 0) softmax_bwd_long_kernel<<<n_long, 256, 0, stream>>>(a);
  return check_launch("edge_softmax_bwd");
}

int gb_edge_score_bwd(const gb_graph* g, const float* ds, const float* q, int32_t ldq, const float* k, int32_t ldk,
                      int32_t heads, int32_t dk, int32_t att_type, const float* kparams, const float* prods,
                      float* dq, int32_t lddq, float* dk_out, int32_t lddk, float* dkparams_partials,
                      float* dkparams, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  (void)kparams; (void)prods; (void)dkparams_partials; (void)dkparams;
  if (int rc = check_graph_att(g, "gb_edge_score_bwd")) return rc;
  GB_REQUIRE(ds && q && k && dq && dk_out && heads > 0 && dk > 0, "gb_edge_score_bwd: bad arguments");
  if (att_type != GB_ATT_SCALED_DOT) {
    set_error("gb_edge_score_bwd: backward of attention_type %d is not implemented (scaled_dot only)", att_type);
    return GB_ERR_UNSUPPORTED;
  }
  const int width = heads * dk;
  GB_REQUIRE(ldq >= width && ldk >= width && lddq >= width && lddk >= width, "gb_edge_score_bwd: leading dimension too small");
  int v = min_int(att_vec(q, ldq, k, ldk, width, dk), att_vec(dq, lddq, dk_out, lddk, width, dk));
  if (width / v > 32) {
    set_error("gb_edge_score_bwd: attention_dim %d with vector width %d exceeds 32 lanes per row", width, v);
    return GB_ERR_UNSUPPORTED;
  }
  MhArgs a{};
  a.ds = ds; a.n = g->n; a.heads = heads; a.dk = dk; a.long_threshold = eff_threshold(g);
  a.scale = 1.0f / sqrtf((float)dk);
  const bool use_long = g->long_threshold > 0;
  // dq[i] = scale * sum_{s in row i} ds[s,h] k[col_s]
  a.ptr = g->rowptr; a.idx = g->colidx; a.perm = nullptr; a.src = k; a.lds = ldk; a.out = dq; a.ldo = lddq;
  a.long_list = g->long_rows;
  int rc;
  const int nlr = use_long ? g->n_long_rows : 0, nlc = use_long ? g->n_long_cols : 0;
  if (v == 4) rc = dispatch_mh<4>(a, nlr, K_EDGE_SCORE_BWD_Q, stream);
  else if (v == 2) rc = dispatch_mh<2>(a, nlr, K_EDGE_SCORE_BWD_Q, stream);
  else rc = dispatch_mh<1>(a, nlr, K_EDGE_SCORE_BWD_Q, stream);
  if (rc) return rc;
  // dk[j] = scale * sum_{p in col j} ds[tperm[p],h] q[row_p]
  a.ptr = g->colptr; a.idx = g->rowidx; a.perm = g->tperm; a.src = q; a.lds = ldq; a.out = dk_out; a.ldo = lddk;
  a.long_list = g->long_cols;
  if (v == 4) return dispatch_mh<4>(a, nlc, K_EDGE_SCORE_BWD_K, stream);
  if (v == 2) return dispatch_mh<2>(a, nlc, K_EDGE_SCORE_BWD_K, stream);
  return dispatch_mh<1>(a, nlc, K_EDGE_SCORE_BWD_K, stream);
}

}  // extern "C"
