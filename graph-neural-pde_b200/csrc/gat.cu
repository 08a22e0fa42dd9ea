// gat.cu - segment softmax of GAT scores (function_GAT_attention.py:117-141 + utils.py:116-127).
//
// The reference gathers both endpoints' [dk, H] projections per edge, multiplies them with the shared vector `a`
// and sums over dk.  The sum factorises: with per-node scalars
//     s_src[i,h] = sum_c a[c]      * (x W)[i, h, c]          s_dst[j,h] = sum_c a[dk + c] * (x W)[j, h, c]
// (a [N,2H] projection the host computes once per evaluation) the raw score of edge e = (row, col) is
//     z[e,h] = s_src[row_e, h] + s_dst[col_e, h],      prods[e,h] = leaky_relu(z[e,h], slope)
// so the per-edge gather is H floats instead of 2 dk H.  One warp per softmax segment (row for
// attention_norm_idx 0, column for 1), lanes over the segment's edges, three passes (max, sum of exp, normalise);
// segments on the plan's long lists (more than long_threshold edges) are taken by one CTA each.  HBM-bound: 4H bytes gathered + 8H + 4
// bytes written per edge.
#include "common.cuh"

namespace gb {

constexpr int kGatMaxHeads = 8;
constexpr int kGatThreads = 256;

struct GatArgs {
  const int32_t* ptr;      // rowptr (norm 0) / colptr (norm 1)
  const int32_t* idx;      // other endpoint of slot p in segment order
  const int32_t* perm;     // nullptr or tperm (CSC slot -> CSR slot)
  const float* own;        // scalars indexed by the segment id: s_src (norm 0) / s_dst (norm 1)
  const float* oth;        // scalars gathered through idx
  const float* scores;     // nullptr, or ready-made raw scores [E,H] in slot order (gb_segment_softmax_fwd)
  int32_t lds, n, heads, long_threshold;
  const int32_t* long_list;
  float slope;
  float* att;              // [E,H] CSR order
  float* abar;             // [E]
  float* prods;            // [E,H] or nullptr
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One segment per warp (CTA == false) or per CTA (CTA == true: the listed long segments).
template <bool CTA>
__global__ void __launch_bounds__(kGatThreads) gat_softmax_kernel(const GatArgs a) {
  constexpr int W = kGatThreads / 32;
  __shared__ float red[CTA ? W * kGatMaxHeads : 1];
  __shared__ float bc[CTA ? kGatMaxHeads : 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int seg;
  if constexpr (CTA) {
    seg = __ldg(a.long_list + blockIdx.x);
  } else {
    seg = blockIdx.x * W + warp;
    if (seg >= a.n) return;
  }
  const int start = __ldg(a.ptr + seg), end = __ldg(a.ptr + seg + 1);
  const bool is_long = end - start > a.long_threshold;
  if (CTA != is_long) return;                          // block-uniform for the CTA kernel
  const int first = start + (CTA ? threadIdx.x : lane);
  const int step = CTA ? kGatThreads : 32;
  const int H = a.heads;

  float own[kGatMaxHeads], m[kGatMaxHeads], sum[kGatMaxHeads];
#pragma unroll
  for (int h = 0; h < kGatMaxHeads; ++h) {
    own[h] = (h < H && a.own) ? __ldg(a.own + (size_t)seg * a.lds + h) : 0.f;
    m[h] = -INFINITY;
    sum[h] = 0.f;
  }

  // reduce v[h] over the segment's lanes, result in every lane
  auto reduce = [&](float* v, bool is_max) {
#pragma unroll
    for (int h = 0; h < kGatMaxHeads; ++h) {
      if (h >= H) break;
      v[h] = is_max ? warp_max(v[h]) : warp_sum(v[h]);
    }
    if constexpr (CTA) {
      __syncthreads();                                 // red / bc free again
      if (lane == 0) {
#pragma unroll
        for (int h = 0; h < kGatMaxHeads; ++h)
          if (h < H) red[warp * kGatMaxHeads + h] = v[h];
      }
      __syncthreads();
      if (threadIdx.x < H) {
        float r = red[threadIdx.x];
        for (int w = 1; w < W; ++w) {
          const float o = red[w * kGatMaxHeads + threadIdx.x];
          r = is_max ? fmaxf(r, o) : r + o;            // fixed order: deterministic
        }
        bc[threadIdx.x] = r;
      }
      __syncthreads();
#pragma unroll
      for (int h = 0; h < kGatMaxHeads; ++h)
        if (h < H) v[h] = bc[h];
    }
  };

  // ---- pass 1: raw scores (parked in att, kept in prods) and the segment maximum -----------------------------
  for (int p = first; p < end; p += step) {
    const int other = a.scores ? 0 : ld_stream_i32(a.idx + p);
    const int slot = a.perm ? ld_stream_i32(a.perm + p) : p;
#pragma unroll
    for (int h = 0; h < kGatMaxHeads; ++h) {
      if (h >= H) break;
      float s;
      if (a.scores) {
        s = ld_stream_f32(a.scores + (size_t)slot * H + h);
      } else {
        const float z = own[h] + __ldg(a.oth + (size_t)other * a.lds + h);
        s = z > 0.f ? z : a.slope * z;
      }
      a.att[(size_t)slot * H + h] = s;
      if (a.prods) a.prods[(size_t)slot * H + h] = s;
      m[h] = fmaxf(m[h], s);
    }
  }
  reduce(m, true);

  // ---- pass 2: exp(s - max), parked in att, and the segment sum ------------------------------------------------
  for (int p = first; p < end; p += step) {
    const int slot = a.perm ? ld_stream_i32(a.perm + p) : p;
#pragma unroll
    for (int h = 0; h < kGatMaxHeads; ++h) {
      if (h >= H) break;
      const float e = expf(a.att[(size_t)slot * H + h] - m[h]);
      a.att[(size_t)slot * H + h] = e;
      sum[h] += e;
    }
  }
  reduce(sum, false);

  // ---- pass 3: normalise (utils.py:124-125: + 1e-16 in the denominator) and the head mean ------------------------
  const float inv_h = 1.0f / (float)H;
  for (int p = first; p < end; p += step) {
    const int slot = a.perm ? ld_stream_i32(a.perm + p) : p;
    float mean = 0.f;
#pragma unroll
    for (int h = 0; h < kGatMaxHeads; ++h) {
      if (h >= H) break;
      const float v = a.att[(size_t)slot * H + h] / (sum[h] + 1e-16f);
      a.att[(size_t)slot * H + h] = v;
      mean += v;
    }
    if (a.abar) a.abar[slot] = mean * inv_h;
  }
}

}  // namespace gb

using namespace gb;

extern "C" {

int gb_gat_softmax_fwd(const gb_graph* g, int32_t norm_idx, const float* s_src, const float* s_dst, int32_t lds,
                       int32_t heads, float slope, float* att, float* abar, float* prods, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(g != nullptr && g->n > 0 && g->e >= 0 && g->rowptr && g->colptr, "gb_gat_softmax_fwd: bad graph");
  GB_REQUIRE(g->e == 0 || (g->colidx && g->rowidx && g->tperm), "gb_gat_softmax_fwd: graph without index arrays");
  GB_REQUIRE(s_src && s_dst && att && abar, "gb_gat_softmax_fwd: null pointer");
  GB_REQUIRE(norm_idx == 0 || norm_idx == 1, "gb_gat_softmax_fwd: attention_norm_idx must be 0 or 1");
  GB_REQUIRE(lds >= heads, "gb_gat_softmax_fwd: leading dimension too small");
  if (heads < 1 || heads > kGatMaxHeads) {
    set_error("gb_gat_softmax_fwd: 1 <= heads <= %d supported, got %d", kGatMaxHeads, heads);
    return GB_ERR_UNSUPPORTED;
  }
  if (g->e == 0) return GB_OK;
  GatArgs a{};
  a.lds = lds; a.n = g->n; a.heads = heads; a.slope = slope; a.att = att; a.abar = abar; a.prods = prods;
  if (norm_idx == 0) { a.ptr = g->rowptr; a.idx = g->colidx; a.perm = nullptr; a.own = s_src; a.oth = s_dst; }
  else { a.ptr = g->colptr; a.idx = g->rowidx; a.perm = g->tperm; a.own = s_dst; a.oth = s_src; }
  int n_long = 0;
  a.long_threshold = g->long_threshold > 0 ? g->long_threshold : 0x7fffffff;
  if (g->long_threshold > 0) {
    a.long_list = norm_idx == 0 ? g->long_rows : g->long_cols;
    n_long = norm_idx == 0 ? g->n_long_rows : g->n_long_cols;
    GB_REQUIRE(n_long == 0 || a.long_list, "gb_gat_softmax_fwd: long-segment list missing");
  }
  const int warps = kGatThreads / 32;
  LaunchScope ls(K_GAT_SOFTMAX, stream);
  gat_softmax_kernel<false><<<(unsigned)((g->n + warps - 1) / warps), kGatThreads, 0, stream>>>(a);
  if (n_long > 0) gat_softmax_kernel<true><<<(unsigned)n_long, kGatThreads, 0, stream>>>(a);
  return check_launch("gat_softmax");
}


int gb_segment_softmax_fwd(const int32_t* ptr, const int32_t* perm, int32_t n, int32_t heads, const float* scores,
                           float* att, float* abar, const int32_t* long_list, int32_t n_long, int32_t long_threshold,
                           void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(ptr && scores && att && n > 0, "gb_segment_softmax_fwd: bad arguments");
  GB_REQUIRE(n_long == 0 || (long_list && long_threshold > 0), "gb_segment_softmax_fwd: long-segment list without threshold");
  if (heads < 1 || heads > kGatMaxHeads) {
    set_error("gb_segment_softmax_fwd: 1 <= heads <= %d supported, got %d", kGatMaxHeads, heads);
    return GB_ERR_UNSUPPORTED;
  }
  GatArgs a{};
  a.ptr = ptr; a.perm = perm; a.scores = scores; a.n = n; a.heads = heads; a.slope = 1.f; a.att = att; a.abar = abar;
  // groups on `long_list` (more than long_threshold members) get a CTA each, everything else a warp
  a.long_list = long_list;
  a.long_threshold = n_long > 0 ? long_threshold : 0x7fffffff;
  const int warps = kGatThreads / 32;
  LaunchScope ls(K_GAT_SOFTMAX, stream);
  gat_softmax_kernel<false><<<(unsigned)((n + warps - 1) / warps), kGatThreads, 0, stream>>>(a);
  if (n_long > 0) gat_softmax_kernel<true><<<(unsigned)n_long, kGatThreads, 0, stream>>>(a);
  return check_launch("segment_softmax");
}

}  // extern "C"
