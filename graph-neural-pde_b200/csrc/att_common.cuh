// att_common.cuh - pieces shared by attention.cu (general kernels) and fused_att.cu (warp-per-row fused kernels)
#pragma once
#include "vec.cuh"

namespace gb {

constexpr int kAttThreads = 256;
inline int lpr_for(int chunks) { return chunks <= 8 ? 8 : (chunks <= 16 ? 16 : 32); }

struct AttArgs {
  const int32_t* ptr;      // segment pointer (rowptr for norm_idx 0, colptr for 1)
  const int32_t* idx;      // the other endpoint of slot p in segment order
  const int32_t* perm;     // nullptr (segment order == CSR order) or tperm (CSC slot -> CSR slot)
  const float* own;        // matrix indexed by the segment id   (q for norm 0, k for norm 1)
  int32_t ldown;
  const float* oth;        // matrix gathered through idx        (k for norm 0, q for norm 1)
  int32_t ldoth;
  int32_t n, heads, dk, att_type, long_threshold;
  const float* edge_w;     // optional [E] (CSR order) multiplier of the raw score
  const float* kparams;    // [output_var, lengthscale] for exp_kernel
  float* att;              // [E,H] CSR order
  float* abar;             // [E]
  float* prods;            // optional [E,H]
  const int32_t* long_list;
};

template <int LPR>
__device__ __forceinline__ unsigned group_mask() {
  if (LPR == 32) return 0xffffffffu;
  const unsigned base = (threadIdx.x & 31) / LPR * LPR;
  return ((1u << (LPR & 31)) - 1u) << base;
}

// Sum `v` over the lanes of one head.  LH > 0: heads are aligned blocks of LH (power of two) lanes ->
// segmented butterfly.  LH == 0: generic layout -> one masked group reduction per head.
template <int LPR, int LH>
__device__ __forceinline__ float head_sum(float v, int my_head, int heads, unsigned mask) {
  if (LH > 0) {
#pragma unroll
    for (int o = LH / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, LPR);
    return v;
  }
  float mine = 0.f;
  for (int h = 0; h < heads; ++h) {
    float c = my_head == h ? v : 0.f;
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) c += __shfl_xor_sync(mask, c, o, LPR);
    if (my_head == h) mine = c;
  }
  return mine;
}

// raw score of one (edge, head) from the lane-local chunks; result valid in every lane of the head
template <int V, class HS>
__device__ __forceinline__ float edge_score_hs(int att_type, int dk, const Vec<V>& qo, const Vec<V>& kg, HS hs,
                                               float inv_sqrt_dk, float ov2, float inv_2l2) {
  if (att_type == GB_ATT_SCALED_DOT) {
    // sum(src * dst_k) / sqrt(dk)
    return hs(qo.dot(kg)) * inv_sqrt_dk;
  }
  if (att_type == GB_ATT_EXP_KERNEL) {
    const Vec<V> df = Vec<V>::map2(qo, kg, [](float p, float q) { return p - q; });
    const float d2 = hs(df.dot(df));
    return ov2 * expf(-d2 * inv_2l2);
  }
  Vec<V> u = qo, w = kg;
  if (att_type == GB_ATT_PEARSON) {
    const float inv_dk = 1.0f / (float)dk;
    const float mu = hs(qo.sum()) * inv_dk;
    const float mw = hs(kg.sum()) * inv_dk;
    const Vec<V> mu_v = Vec<V>::map2(qo, qo, [mu](float, float) { return mu; });
    const Vec<V> mw_v = Vec<V>::map2(kg, kg, [mw](float, float) { return mw; });
    u = Vec<V>::map2(qo, mu_v, [](float p, float q) { return p - q; });
    w = Vec<V>::map2(kg, mw_v, [](float p, float q) { return p - q; });
  }
  // torch.nn.CosineSimilarity(eps=1e-5): (x1 / max(|x1|, eps)) . (x2 / max(|x2|, eps))
  const float uu = hs(u.dot(u));
  const float ww = hs(w.dot(w));
  const float uw = hs(u.dot(w));
  const float nu = fmaxf(sqrtf(uu), 1e-5f), nw = fmaxf(sqrtf(ww), 1e-5f);
  return uw / (nu * nw);
}

template <int V, int LPR, int LH>
__device__ __forceinline__ float edge_score(const AttArgs& a, const Vec<V>& qo, const Vec<V>& kg, int my_head,
                                            unsigned mask, float inv_sqrt_dk, float ov2, float inv_2l2) {
  const int heads = a.heads;
  return edge_score_hs<V>(a.att_type, a.dk, qo, kg,
                          [=](float v) { return head_sum<LPR, LH>(v, my_head, heads, mask); }, inv_sqrt_dk, ov2,
                          inv_2l2);
}

struct SddmmArgs {
  const int32_t* ptr;
  const int32_t* idx;
  const float* own;        // gout (indexed by the row)
  int32_t ldown;
  const float* oth;        // x (gathered through colidx)
  int32_t ldoth;
  int32_t n, dv, long_threshold, accumulate;
  const float* alpha_train;
  int32_t alpha_sigmoid;
  float* dabar;
  const int32_t* long_list;
};

struct SmBwdArgs {
  const int32_t* ptr;
  const int32_t* perm;     // nullptr or tperm
  const float* att;        // [E,H]
  const float* dabar;      // optional [E]   dL/d(mean_h att)
  const float* datt;       // optional [E,H] dL/d att (CSR order), added to dabar/H
  const float* edge_w;     // optional
  float* ds;               // [E,H] dL/d(raw score)
  int32_t n, heads, long_threshold;
  const int32_t* long_list;
};

struct MhArgs {
  const int32_t* ptr;
  const int32_t* idx;
  const int32_t* perm;
  const float* ds;         // [E,H] CSR order
  const float* src;
  int32_t lds;
  float* out;
  int32_t ldo;
  int32_t n, heads, dk, long_threshold;
  float scale;
  const int32_t* long_list;
};

// ---- long-segment launchers shared with fused_att.cu (defined in attention.cu / spmm.cu) ----------
int att_long_fwd(const AttArgs& a, int v, int n_long, cudaStream_t stream);
int sddmm_long(const SddmmArgs& a, int v, int n_long, cudaStream_t stream);
int softmax_bwd_long(const SmBwdArgs& a, int n_long, cudaStream_t stream);
int mh_spmm_long(const MhArgs& a, int v, int n_long, int kernel_id, cudaStream_t stream);

}  // namespace gb
