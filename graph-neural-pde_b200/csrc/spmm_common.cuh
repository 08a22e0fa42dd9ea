// spmm_common.cuh - argument block and fused diffusion epilogue shared by spmm.cu and fused_att.cu
#pragma once
#include "vec.cuh"

namespace gb {

constexpr int kSpmmThreads = 256;
constexpr int kRowTiles = 4;   // consecutive row tiles per CTA
constexpr int kMidFloats = 2 * 1024 * 2;   // level-1 reduction results: 2 x 1024 doubles

struct SpmmArgs {
  const int32_t* ptr;
  const int32_t* idx;
  const int32_t* perm;     // nullptr: vals[p]; else vals[perm[p]]
  const float* vals;
  int32_t n, dv, long_threshold;
  const float* src;        // gathered matrix; also the "- self" term
  int32_t lds;
  const float* alpha_train;
  int32_t alpha_sigmoid;
  // forward epilogue
  const float* x0;
  int32_t ldx0;
  const float* beta_train;
  float* out;
  int32_t ldo;
  int32_t has_stage;
  gb_stage stage;
  // backward epilogue
  const float* x;
  int32_t ldx;
  int32_t accumulate;
  float* partials;         // [blocks][2]
  const int32_t* long_list;
  int32_t partial_base;    // first partial slot of this launch
  // 1-D row partition (backward only): segments outside [part_lo, part_hi) are columns this rank does not
  // own - it contributes sum_e a_e g[row_e] but not the "- g[j]" self term nor <g_j, x0_j>.  0,0 = own all.
  int32_t part_lo, part_hi;
  int32_t row0;            // first segment processed by the launch (forward row partition); segments [row0, n)
};

__device__ __forceinline__ float load_alpha(const SpmmArgs& a) {
  const float v = __ldg(a.alpha_train);
  return a.alpha_sigmoid ? sigmoidf_(v) : v;
}

// epilogue of one vector chunk of one row
// EARLY (forward only): every operand of the epilogue is requested before the first is used - one global round trip per
// row instead of one per operand (a load followed by its fma stalls the warp before the next load is issued).  Same
// arithmetic in the same order.  Measured (profiles/r02x_epilogue_ab.log): k-gather forward rows 0.385 -> 0.360 ms on the
// ogbn-arxiv shape, 9.18 -> 8.85 ms on ogbn-products; the adjoint row kernel and the SpMM kernels lose (registers), so
// only the forward row kernel asks for it.
template <int V, bool BWD, bool EARLY = false>
__device__ __forceinline__ void row_epilogue(const SpmmArgs& a, int row, int col, const Vec<V>& acc, float alpha,
                                             float beta, float& s1, float& s2) {
  const bool owned = !BWD || a.part_hi == 0 || (row >= a.part_lo && row < a.part_hi);
  const Vec<V> self = owned ? Vec<V>::load_stream(a.src + (size_t)row * a.lds + col) : Vec<V>::zero();
  if (!BWD && EARLY) {
    Vec<V> s = Vec<V>::zero(), y = Vec<V>::zero(), kj[3];
    if (a.x0 != nullptr) s = Vec<V>::load_stream(a.x0 + (size_t)row * a.ldx0 + col);
    const size_t off = a.has_stage ? (size_t)row * a.stage.ld + col : 0;
    if (a.has_stage) {
      y = Vec<V>::load_stream(a.stage.y + off);
#pragma unroll
      for (int j = 0; j < 3; ++j)                              // rk4: at most three earlier stages
        kj[j] = (j < a.stage.n_prev && a.stage.c[j] != 0.f) ? Vec<V>::load_stream(a.stage.k[j] + off) : Vec<V>::zero();
    }
    Vec<V> t = Vec<V>::map2(acc, self, [](float p, float q) { return p - q; });
    Vec<V> f = Vec<V>::map2(t, t, [alpha](float p, float) { return alpha * p; });
    if (a.x0 != nullptr) f = Vec<V>::map2(f, s, [beta](float p, float q) { return fmaf(beta, q, p); });
    if (!a.has_stage || a.stage.write_f) f.store_stream(a.out + (size_t)row * a.ldo + col);
    if (a.has_stage) {                                         // y + sum_j c_j k_j + c_last f, in this order
#pragma unroll
      for (int j = 0; j < 3; ++j)
        if (j < a.stage.n_prev && a.stage.c[j] != 0.f) y.fma(a.stage.c[j], kj[j]);
      for (int j = 3; j < a.stage.n_prev; ++j) {
        const float c = a.stage.c[j];
        if (c != 0.f) y.fma(c, Vec<V>::load_stream(a.stage.k[j] + off));
      }
      y.fma(a.stage.c[a.stage.n_prev], f);
      y.store_stream(a.stage.ynext + off);
    }
  } else if (!BWD) {
    Vec<V> t = Vec<V>::map2(acc, self, [](float p, float q) { return p - q; });
    Vec<V> f = Vec<V>::map2(t, t, [alpha](float p, float) { return alpha * p; });
    if (a.x0 != nullptr) {
      const Vec<V> s = Vec<V>::load_stream(a.x0 + (size_t)row * a.ldx0 + col);
      f = Vec<V>::map2(f, s, [beta](float p, float q) { return fmaf(beta, q, p); });
    }
    if (!a.has_stage || a.stage.write_f) f.store_stream(a.out + (size_t)row * a.ldo + col);
    if (a.has_stage) {
      const size_t off = (size_t)row * a.stage.ld + col;
      Vec<V> y = Vec<V>::load_stream(a.stage.y + off);
      for (int j = 0; j < a.stage.n_prev; ++j) {
        const float c = a.stage.c[j];
        if (c != 0.f) y.fma(c, Vec<V>::load_stream(a.stage.k[j] + off));
      }
      y.fma(a.stage.c[a.stage.n_prev], f);
      y.store_stream(a.stage.ynext + off);
    }
  } else {
    Vec<V> t = Vec<V>::map2(acc, self, [](float p, float q) { return p - q; });
    float* dxp = a.out + (size_t)row * a.ldo + col;
    Vec<V> dx = Vec<V>::map2(t, t, [alpha](float p, float) { return alpha * p; });
    if (a.accumulate) {
      const Vec<V> old = Vec<V>::load_rw(dxp);
      dx = Vec<V>::map2(dx, old, [](float p, float q) { return p + q; });
    }
    dx.store_stream(dxp);
    if (a.x != nullptr) s1 += Vec<V>::load_stream(a.x + (size_t)row * a.ldx + col).dot(t);
    if (a.x0 != nullptr && owned) s2 += Vec<V>::load_stream(a.x0 + (size_t)row * a.ldx0 + col).dot(self);
  }
}

static __device__ __forceinline__ void write_partials(const SpmmArgs& a, float s1, float s2, int slot) {
  __shared__ float red[64];
  const float r1 = block_sum(s1, red);
  const float r2 = block_sum(s2, red + 32);
  if (threadIdx.x == 0) {
    a.partials[2 * (size_t)slot] = r1;
    a.partials[2 * (size_t)slot + 1] = r2;
  }
}

int reduce_partials_launch(float* partials, int slots, const float* alpha_train, int alpha_sigmoid, float* dparams,
                           cudaStream_t stream);
// launches only the CTA-per-row kernel over `a.long_list` (used by fused_att.cu); defined in spmm.cu
int spmm_long_only(const SpmmArgs& a, int v, bool bwd, int n_long, cudaStream_t stream);

}  // namespace gb
