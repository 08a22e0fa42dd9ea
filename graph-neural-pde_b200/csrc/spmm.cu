// spmm.cu - K1 / K5a: row-segmented SpMM over CSR (forward) and CSC (transpose, backward) with the
// diffusion epilogue fused in.
//
//   forward  (LaplacianODEFunc.forward, src/function_laplacian_diffusion.py:60-77):
//       out[i,:] = alpha * (sum_{s in row i} a_s x[col_s,:] - x[i,:]) + beta * x0[i,:]
//       (+ optional Runge-Kutta stage combination  ynext = y + sum c_j k_j + c_s out)
//   backward (autograd of the above; the reference relies on torch autograd / odeint_adjoint):
//       t[j,:]  = sum_{p in col j} a_{tperm[p]} g[row_p,:] - g[j,:] ;  dx[j,:] = alpha * t[j,:]
//       dalpha_train = alpha' * sum_j <x[j,:], t[j,:]>      (== sum g.(Ax - x), re-associated so that
//       dbeta_train  = sum_j <g[j,:], x0[j,:]>               Ax never has to be stored or recomputed)
//
// Work decomposition: a row is owned by a sub-warp group of LPR lanes, each lane holding CPL vector
// chunks (128-bit when d % 4 == 0) of the accumulator; the edge loop is unrolled so that every lane
// keeps U*CPL independent row gathers in flight.  Rows longer than plan.long_threshold are skipped
// here and handled by one CTA each (spmm_long), whose warps stride the edge list and combine through
// shared memory in a fixed order.  No atomics anywhere: results are bit-reproducible run to run.
//
// Roofline: HBM.  Algorithmic bytes per forward launch: 8E + 12Nd + 4N.
#include <mutex>
#include <stdlib.h>
#include "spmm_common.cuh"
#include "spmm2_kernels.cuh"

namespace gb {

template <int V, int LPR, int CPL, bool BWD>
__global__ void __launch_bounds__(kSpmmThreads) spmm_rows(const SpmmArgs a) {
  constexpr int RPT = kSpmmThreads / LPR;          // rows per tile
  constexpr int U = CPL == 1 ? 4 : 2;              // edge unroll
  const int gl = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const float alpha = load_alpha(a);
  const float beta = (!BWD && a.beta_train != nullptr) ? __ldg(a.beta_train) : 0.f;
  float s1 = 0.f, s2 = 0.f;
#pragma unroll 1
  for (int tile = 0; tile < kRowTiles; ++tile) {
    const int row = a.row0 + (blockIdx.x * kRowTiles + tile) * RPT + grp;
    if (row >= a.n) break;
    const int start = __ldg(a.ptr + row), end = __ldg(a.ptr + row + 1);
    if (end - start > a.long_threshold) continue;    // handled by spmm_long
    Vec<V> acc[CPL];
#pragma unroll
    for (int j = 0; j < CPL; ++j) acc[j] = Vec<V>::zero();
    for (int p = start; p < end; p += U) {
      int c[U];
      float w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int pp = p + u;
        const bool ok = pp < end;
        c[u] = ok ? ld_stream_i32(a.idx + pp) : -1;
        if (BWD)
          w[u] = ok ? __ldg(a.vals + ld_stream_i32(a.perm + pp)) : 0.f;
        else
          w[u] = ok ? ld_stream_f32(a.vals + pp) : 0.f;
      }
      Vec<V> r[U][CPL];
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int j = 0; j < CPL; ++j) {
          const int ch = gl + j * LPR;
          r[u][j] = (c[u] >= 0 && ch < a.dv) ? Vec<V>::load(a.src + (size_t)c[u] * a.lds + ch * V) : Vec<V>::zero();
        }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int j = 0; j < CPL; ++j) acc[j].fma(w[u], r[u][j]);
    }
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      const int ch = gl + j * LPR;
      if (ch < a.dv) row_epilogue<V, BWD>(a, row, ch * V, acc[j], alpha, beta, s1, s2);
    }
  }
  if (BWD) write_partials(a, s1, s2, a.partial_base + blockIdx.x);
}

// one CTA per long row: groups stride the edge list, fixed-order combine through shared memory
template <int V, int LPR, int CPL, bool BWD>
__global__ void __launch_bounds__(kSpmmThreads) spmm_long(const SpmmArgs a) {
  constexpr int G = kSpmmThreads / LPR;
  __shared__ __align__(16) float red[kSpmmThreads * CPL * V];
  const int gl = threadIdx.x % LPR;
  const int grp = threadIdx.x / LPR;
  const int row = __ldg(a.long_list + blockIdx.x);
  const int start = __ldg(a.ptr + row), end = __ldg(a.ptr + row + 1);
  if (end - start <= a.long_threshold) {                    // listed, but taken by a warp-level kernel (block-uniform)
    if (BWD) write_partials(a, 0.f, 0.f, a.partial_base + blockIdx.x);
    return;
  }
  const float alpha = load_alpha(a);
  const float beta = (!BWD && a.beta_train != nullptr) ? __ldg(a.beta_train) : 0.f;
  Vec<V> acc[CPL];
#pragma unroll
  for (int j = 0; j < CPL; ++j) acc[j] = Vec<V>::zero();
  constexpr int U = 2;
  for (int p = start + grp; p < end; p += U * G) {
    int c[U];
    float w[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int pp = p + u * G;
      const bool ok = pp < end;
      c[u] = ok ? ld_stream_i32(a.idx + pp) : -1;
      if (BWD)
        w[u] = ok ? __ldg(a.vals + ld_stream_i32(a.perm + pp)) : 0.f;
      else
        w[u] = ok ? ld_stream_f32(a.vals + pp) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < CPL; ++j) {
        const int ch = gl + j * LPR;
        if (c[u] >= 0 && ch < a.dv) acc[j].fma(w[u], Vec<V>::load(a.src + (size_t)c[u] * a.lds + ch * V));
      }
  }
  // red[grp][ch][V]
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    const int ch = gl + j * LPR;
    acc[j].store(red + ((size_t)grp * LPR * CPL + ch) * V);
  }
  __syncthreads();
  float s1 = 0.f, s2 = 0.f;
  for (int ch = threadIdx.x; ch < a.dv; ch += kSpmmThreads) {
    Vec<V> tot = Vec<V>::zero();
    for (int g = 0; g < G; ++g) {
      Vec<V> part;
      part = *reinterpret_cast<const Vec<V>*>(red + ((size_t)g * LPR * CPL + ch) * V);
      tot = Vec<V>::map2(tot, part, [](float p, float q) { return p + q; });
    }
    row_epilogue<V, BWD>(a, row, ch * V, tot, alpha, beta, s1, s2);
  }
  if (BWD) write_partials(a, s1, s2, a.partial_base + blockIdx.x);
}

// fixed-order two-level reduction of the [count][2] partials into dparams[2]
__global__ void __launch_bounds__(256) reduce_partials_l1(const float* __restrict__ partials, int count, int per_block,
                                                        double* __restrict__ mid) {
  __shared__ double sh[2][256];
  const int b0 = blockIdx.x * per_block;
  const int b1 = min(count, b0 + per_block);
  double a1 = 0.0, a2 = 0.0;
  for (int i = b0 + threadIdx.x; i < b1; i += 256) {
    a1 += (double)partials[2 * (size_t)i];
    a2 += (double)partials[2 * (size_t)i + 1];
  }
  sh[0][threadIdx.x] = a1;
  sh[1][threadIdx.x] = a2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    mid[2 * blockIdx.x] = sh[0][0];
    mid[2 * blockIdx.x + 1] = sh[1][0];
  }
}

__global__ void __launch_bounds__(256) reduce_partials_l2(const double* __restrict__ mid, int count,
                                                        const float* __restrict__ alpha_train, int alpha_sigmoid,
                                                        float* __restrict__ dparams) {
  __shared__ double sh[2][256];
  double a1 = 0.0, a2 = 0.0;
  for (int i = threadIdx.x; i < count; i += 256) {
    a1 += mid[2 * i];
    a2 += mid[2 * i + 1];
  }
  sh[0][threadIdx.x] = a1;
  sh[1][threadIdx.x] = a2;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    float scale = 1.f;
    if (alpha_sigmoid) {
      const float s = sigmoidf_(__ldg(alpha_train));
      scale = s * (1.f - s);
    }
    dparams[0] = (float)(sh[0][0] * (double)scale);
    dparams[1] = (float)sh[1][0];
  }
}

}  // namespace gb

namespace gb {

// partials = [kMidFloats floats of level-1 doubles][slots][2]; fixed-order two-level reduction -> dparams[2]
int reduce_partials_launch(float* partials, int slots, const float* alpha_train, int alpha_sigmoid, float* dparams,
                           cudaStream_t stream) {
  double* mid = reinterpret_cast<double*>(partials);
  int l1_blocks = min_int(1024, (slots + 255) / 256);
  if (l1_blocks < 1) l1_blocks = 1;
  const int per_block = (slots + l1_blocks - 1) / l1_blocks;
  LaunchScope ls(K_REDUCE_PARTIALS, stream);
  reduce_partials_l1<<<l1_blocks, 256, 0, stream>>>(partials + kMidFloats, slots, per_block, mid);
  reduce_partials_l2<<<1, 256, 0, stream>>>(mid, l1_blocks, alpha_train, alpha_sigmoid, dparams);
  return check_launch("reduce_partials");
}

}  // namespace gb


using namespace gb;

namespace {

template <int V, int LPR, int CPL, bool BWD>
int launch_shape(const SpmmArgs& a, int n_long, cudaStream_t stream) {
  constexpr int RPT = kSpmmThreads / LPR;
  const int blocks = (a.n - a.row0 + RPT * kRowTiles - 1) / (RPT * kRowTiles);
  // hub rows (> a.long_threshold edges, one CTA each) FIRST and on the forked high-priority side stream: their few
  // CTAs are placed before the main kernel's fill every SM and run next to it (profiles/r01q: as a second serial
  // launch they cost more than the main kernel on R-MAT graphs)
  cudaStream_t side = stream;
  if (n_long > 0) {
    side = side_fork(stream);
    SpmmArgs b = a;
    b.partial_base = a.partial_base + blocks;
    LaunchScope ls(BWD ? K_SPMM_BWD_LONG : K_SPMM_FWD_LONG, side);
    spmm_long<V, LPR, CPL, BWD><<<n_long, kSpmmThreads, 0, side>>>(b);
  }
  if (blocks > 0) {
    LaunchScope ls(BWD ? K_SPMM_BWD : K_SPMM_FWD, stream);
    spmm_rows<V, LPR, CPL, BWD><<<blocks, kSpmmThreads, 0, stream>>>(a);
  }
  if (side != stream) side_join(stream, side);
  return blocks + n_long;
}

template <int V, bool BWD>
int launch_v(const SpmmArgs& a, int n_long, cudaStream_t stream) {
  const RowShape rs = row_shape(a.dv);
  if (rs.lpr == 8 && rs.cpl == 1) return launch_shape<V, 8, 1, BWD>(a, n_long, stream);
  if (rs.lpr == 16) return launch_shape<V, 16, 1, BWD>(a, n_long, stream);
  if (rs.lpr == 8 && rs.cpl == 3) return launch_shape<V, 8, 3, BWD>(a, n_long, stream);
  if (rs.cpl == 1) return launch_shape<V, 32, 1, BWD>(a, n_long, stream);
  if (rs.cpl == 2) return launch_shape<V, 32, 2, BWD>(a, n_long, stream);
  if (rs.cpl == 3) return launch_shape<V, 32, 3, BWD>(a, n_long, stream);
  return launch_shape<V, 32, 4, BWD>(a, n_long, stream);
}

// number of partial slots one column slice uses (upper bound over all row shapes: LPR = 32)
inline int slice_blocks(int n, int n_long) {
  const int rpt = kSpmmThreads / 32;
  return (n + rpt * kRowTiles - 1) / (rpt * kRowTiles) + n_long;
}

// splits the d columns into slices of at most 128 vector chunks and launches each
// ---- large graphs: the dynamically scheduled kernels of spmm2_kernels.cuh ------------------------------------------
// Scheduler counters live in a file-scope device array (no workspace argument in gb_spmm_*): a ring of 64, one per
// launch in turn, zeroed on the launching stream before the kernel; launches that overlap on the GPU use different
// counters unless more than 64 are in flight.
static __device__ int32_t g_spmm2_counters[64];
constexpr int kSpmm2MinRows = 262144;     // below this the statically tiled kernels keep every SM busy with less set-up

// GRANDB200_SPMM_IMPL=1 keeps the statically tiled kernels; GRANDB200_SPMM2_MIN_ROWS overrides the size threshold
// (tests run the dynamic kernels on small graphs with it).  Read per call: a getenv is nothing next to a launch.
bool spmm2_enabled() {
  const char* e = getenv("GRANDB200_SPMM_IMPL");
  return !(e != nullptr && e[0] == '1');
}
int spmm2_min_rows() {
  const char* e = getenv("GRANDB200_SPMM2_MIN_ROWS");
  return e != nullptr && e[0] ? atoi(e) : kSpmm2MinRows;
}

int spmm2_counter(int32_t** out) {
  static int32_t* base[64] = {};
  static std::mutex mu;
  static unsigned next = 0;
  int dev = -1;
  GB_CUDA(cudaGetDevice(&dev));
  GB_REQUIRE(dev >= 0 && dev < 64, "spmm2: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (base[dev] == nullptr) {
    void* p = nullptr;
    GB_CUDA(cudaGetSymbolAddress(&p, g_spmm2_counters));
    base[dev] = static_cast<int32_t*>(p);
  }
  *out = base[dev] + (next++ % 64);
  return GB_OK;
}

template <bool BWD>
int run_spmm2(const SpmmArgs& a, int n_long, cudaStream_t stream, int* slots_used) {
  int32_t* counter = nullptr;
  if (int rc = spmm2_counter(&counter)) return rc;
  if (a.dv <= 16) return spmm2_launch<16, BWD, 24>(a, n_long, counter, stream, slots_used);
  return spmm2_launch<32, BWD, 24>(a, n_long, counter, stream, slots_used);
}

template <bool BWD>
int run_spmm(SpmmArgs a, int d, int v, int n_long, cudaStream_t stream, int* slots_used) {
  if (v == 4 && d <= 128 && a.n - a.row0 >= spmm2_min_rows() && spmm2_enabled()) {
    a.dv = d / 4;
    a.partial_base = 0;
    return run_spmm2<BWD>(a, n_long, stream, slots_used);
  }
  const int max_cols = 128 * v;
  int slot = 0;
  for (int c0 = 0; c0 < d; c0 += max_cols) {
    const int cols = min_int(max_cols, d - c0);
    SpmmArgs s = a;
    s.dv = cols / v;
    s.src = a.src + c0;
    s.out = a.out + c0;
    if (a.x0) s.x0 = a.x0 + c0;
    if (a.x) s.x = a.x + c0;
    if (a.has_stage) {
      s.stage.y = a.stage.y + c0;
      s.stage.ynext = a.stage.ynext + c0;
      for (int j = 0; j < a.stage.n_prev; ++j) s.stage.k[j] = a.stage.k[j] + c0;
    }
    s.partial_base = slot;
    int used;
    if (v == 4) used = launch_v<4, BWD>(s, n_long, stream);
    else if (v == 2) used = launch_v<2, BWD>(s, n_long, stream);
    else used = launch_v<1, BWD>(s, n_long, stream);
    slot += used;
  }
  if (slots_used) *slots_used = slot;
  return check_launch(BWD ? "spmm_bwd" : "spmm_fwd");
}

// rows longer than the plan's list threshold (256) go to spmm_long.  Raising this to 2048 for the statically tiled
// spmm_rows was measured and rejected (profiles/r02p: C5 GRAND-l 644 -> 718 ms/step, C3 BLEND 8.9 -> 16.2 ms: a 16-lane
// group walking 2000 edges stalls its whole CTA); large graphs take the dynamically scheduled spmm2 kernels instead
int spmm_long_threshold(const gb_graph* g) { return g->long_threshold > 0 ? g->long_threshold : 0x7fffffff; }

int check_graph(const gb_graph* g, const char* who) {
  GB_REQUIRE(g != nullptr, "%s: null graph", who);
  GB_REQUIRE(g->n > 0 && g->e >= 0, "%s: bad graph sizes n=%d e=%d", who, g->n, g->e);
  GB_REQUIRE(g->rowptr && g->colptr, "%s: graph without rowptr/colptr", who);
  GB_REQUIRE(g->e == 0 || (g->colidx && g->rowidx && g->tperm), "%s: graph without index arrays", who);
  GB_REQUIRE(g->n_long_rows == 0 || g->long_rows, "%s: long_rows missing", who);
  GB_REQUIRE(g->n_long_cols == 0 || g->long_cols, "%s: long_cols missing", who);
  return GB_OK;
}

}  // namespace


namespace gb {

template <int V, bool BWD>
static int spmm_long_v(const SpmmArgs& a, int n_long, cudaStream_t stream) {
  const RowShape rs = row_shape(a.dv);
  if (rs.lpr == 8 && rs.cpl == 1) spmm_long<V, 8, 1, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else if (rs.lpr == 16) spmm_long<V, 16, 1, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else if (rs.lpr == 8 && rs.cpl == 3) spmm_long<V, 8, 3, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else if (rs.cpl == 1) spmm_long<V, 32, 1, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else if (rs.cpl == 2) spmm_long<V, 32, 2, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else if (rs.cpl == 3) spmm_long<V, 32, 3, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  else spmm_long<V, 32, 4, BWD><<<n_long, kSpmmThreads, 0, stream>>>(a);
  return check_launch("spmm_long_only");
}

int spmm_long_only(const SpmmArgs& a, int v, bool bwd, int n_long, cudaStream_t stream) {
  if (n_long <= 0) return GB_OK;
  LaunchScope ls(bwd ? K_SPMM_BWD_LONG : K_SPMM_FWD_LONG, stream);
  if (v == 4) return bwd ? spmm_long_v<4, true>(a, n_long, stream) : spmm_long_v<4, false>(a, n_long, stream);
  if (v == 2) return bwd ? spmm_long_v<2, true>(a, n_long, stream) : spmm_long_v<2, false>(a, n_long, stream);
  return bwd ? spmm_long_v<1, true>(a, n_long, stream) : spmm_long_v<1, false>(a, n_long, stream);
}

}  // namespace gb

extern "C" {

int gb_spmm_fwd(const gb_graph* g, const float* vals, const float* x, int32_t ldx, const float* x0, int32_t ldx0,
                const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, int32_t d, float* out,
                int32_t ldo, const gb_stage* stage, void* stream_) {
  return gb_spmm_fwd_part(g, 0, g ? g->n : 0, vals, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, beta_train, d, out,
                          ldo, stage, stream_);
}

int gb_spmm_fwd_part(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* vals, const float* x,
                     int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                     const float* beta_train, int32_t d, float* out, int32_t ldo, const gb_stage* stage,
                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_graph(g, "gb_spmm_fwd")) return rc;
  GB_REQUIRE(row_lo >= 0 && row_lo <= row_hi && row_hi <= g->n, "gb_spmm_fwd: bad row range [%d, %d)", row_lo, row_hi);
  GB_REQUIRE(x && out && alpha_train && d > 0, "gb_spmm_fwd: null pointer or d <= 0");
  GB_REQUIRE(g->e == 0 || vals, "gb_spmm_fwd: null vals");
  GB_REQUIRE(ldx >= d && ldo >= d, "gb_spmm_fwd: leading dimension smaller than d");
  SpmmArgs a{};
  a.ptr = g->rowptr; a.idx = g->colidx; a.perm = nullptr; a.vals = vals;
  a.n = row_hi; a.row0 = row_lo; a.long_threshold = spmm_long_threshold(g);
  a.src = x; a.lds = ldx; a.alpha_train = alpha_train; a.alpha_sigmoid = alpha_sigmoid;
  const bool source = (x0 != nullptr && beta_train != nullptr);
  a.x0 = source ? x0 : nullptr; a.ldx0 = ldx0; a.beta_train = source ? beta_train : nullptr;
  a.out = out; a.ldo = ldo;
  int v = min_int(vec_width(x, ldx, d), vec_width(out, ldo, d));
  if (source) v = min_int(v, vec_width(x0, ldx0, d));
  if (stage) {
    GB_REQUIRE(stage->y && stage->ynext && stage->n_prev >= 0 && stage->n_prev <= 6 && stage->ld >= d,
               "gb_spmm_fwd: bad stage descriptor");
    a.has_stage = 1; a.stage = *stage;
    v = min_int(v, min_int(vec_width(stage->y, stage->ld, d), vec_width(stage->ynext, stage->ld, d)));
    for (int j = 0; j < stage->n_prev; ++j) {
      GB_REQUIRE(stage->k[j] != nullptr, "gb_spmm_fwd: null stage k[%d]", j);
      v = min_int(v, vec_width(stage->k[j], stage->ld, d));
    }
  }
  const int n_long = g->long_threshold > 0 ? g->n_long_rows : 0;
  a.long_list = g->long_rows;
  return run_spmm<false>(a, d, v, n_long, stream, nullptr);
}

size_t gb_spmm_bwd_partials(const gb_graph* g, int32_t d) {
  if (!g || g->n <= 0 || d <= 0) return 0;
  const int slices = (d + 127) / 128;                       // worst case: scalar path, 128 columns per slice
  size_t per_slice = (size_t)slice_blocks(g->n, g->n_long_cols);
  if ((size_t)spmm2_slots(g->n, g->n_long_cols) > per_slice) per_slice = (size_t)spmm2_slots(g->n, g->n_long_cols);
  const size_t slots = (size_t)slices * per_slice;
  return slots * 2 + kMidFloats + 64;                       // in floats
}

int gb_spmm_bwd(const gb_graph* g, const float* vals, const float* gout, int32_t ldg, const float* x, int32_t ldx,
                const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid, int32_t d, float* dx,
                int32_t lddx, int32_t accumulate, float* partials, float* dparams, void* stream_) {
  return gb_spmm_bwd_part(g, 0, 0, vals, gout, ldg, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, d, dx, lddx,
                          accumulate, partials, dparams, stream_);
}

int gb_spmm_bwd_part(const gb_graph* g, int32_t own_lo, int32_t own_hi, const float* vals, const float* gout,
                     int32_t ldg, const float* x, int32_t ldx, const float* x0, int32_t ldx0,
                     const float* alpha_train, int32_t alpha_sigmoid, int32_t d, float* dx, int32_t lddx,
                     int32_t accumulate, float* partials, float* dparams, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (int rc = check_graph(g, "gb_spmm_bwd")) return rc;
  GB_REQUIRE(gout && dx && alpha_train && d > 0, "gb_spmm_bwd: null pointer or d <= 0");
  GB_REQUIRE(g->e == 0 || vals, "gb_spmm_bwd: null vals");
  GB_REQUIRE(ldg >= d && lddx >= d, "gb_spmm_bwd: leading dimension smaller than d");
  GB_REQUIRE(dparams == nullptr || partials != nullptr, "gb_spmm_bwd: dparams requested without partials workspace");
  SpmmArgs a{};
  a.ptr = g->colptr; a.idx = g->rowidx; a.perm = g->tperm; a.vals = vals;
  a.n = g->n; a.long_threshold = spmm_long_threshold(g);
  a.src = gout; a.lds = ldg; a.alpha_train = alpha_train; a.alpha_sigmoid = alpha_sigmoid;
  a.x0 = dparams ? x0 : nullptr; a.ldx0 = ldx0;
  a.x = dparams ? x : nullptr; a.ldx = ldx;
  a.out = dx; a.ldo = lddx; a.accumulate = accumulate;
  a.part_lo = own_lo; a.part_hi = own_hi;
  int v = min_int(vec_width(gout, ldg, d), vec_width(dx, lddx, d));
  if (a.x) v = min_int(v, vec_width(x, ldx, d));
  if (a.x0) v = min_int(v, vec_width(x0, ldx0, d));
  // workspace layout: [2*1024 doubles of level-1 results][slots][2] floats
  GB_REQUIRE(partials != nullptr, "gb_spmm_bwd: partials workspace is required");
  GB_REQUIRE(reinterpret_cast<uintptr_t>(partials) % 8 == 0, "gb_spmm_bwd: partials must be 8-byte aligned");
  double* mid = reinterpret_cast<double*>(partials);
  a.partials = partials + kMidFloats;
  const int n_long = g->long_threshold > 0 ? g->n_long_cols : 0;
  a.long_list = g->long_cols;
  int slots = 0;
  if (int rc = run_spmm<true>(a, d, v, n_long, stream, &slots)) return rc;
  if (dparams) return reduce_partials_launch(partials, slots, alpha_train, alpha_sigmoid, dparams, stream);
  return GB_OK;
}

// ---- allocation-free variant of gb_spmm_bwd (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops) ---------------------
// The reduction workspace (`partials` of gb_spmm_bwd) lives in a file-scope device array: every GPU that loads
// the library holds GB_SPMM_BWD_STATIC_MAX_FLOATS floats (4 MiB) until the process exits.  In host code the
// array's NAME is a host-side stand-in, not a device address; the address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static __device__ __align__(256) float g_spmm_bwd_scratch[GB_SPMM_BWD_STATIC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float* g_scratch_ptr[kMaxDevices] = {};                  // per device, filled on first use
std::mutex g_scratch_mutex;

int static_scratch(float** out) {
  int dev = -1;
  GB_CUDA(cudaGetDevice(&dev));
  GB_REQUIRE(dev >= 0 && dev < kMaxDevices, "gb_spmm_bwd_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    GB_CUDA(cudaGetSymbolAddress(&p, g_spmm_bwd_scratch));   // the CURRENT device's copy of the array
    g_scratch_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return GB_OK;
}
}  // namespace

int gb_spmm_bwd_static(const gb_graph* g, const float* vals, const float* gout, int32_t ldg, const float* x,
                       int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                       int32_t d, float* dx, int32_t lddx, int32_t accumulate, float* dparams, void* stream_) {
  if (int rc = check_graph(g, "gb_spmm_bwd_static")) return rc;
  GB_REQUIRE(d > 0, "gb_spmm_bwd_static: d <= 0");
  const size_t need = gb_spmm_bwd_partials(g, d);
  if (need > (size_t)GB_SPMM_BWD_STATIC_MAX_FLOATS) {
    set_error("gb_spmm_bwd_static: n=%d, d=%d needs %zu floats of reduction scratch, the built-in array holds %d; "
              "use gb_spmm_bwd with a caller-provided workspace", g->n, d, need, (int)GB_SPMM_BWD_STATIC_MAX_FLOATS);
    return GB_ERR_WORKSPACE;
  }
  float* scratch = nullptr;
  if (int rc = static_scratch(&scratch)) return rc;
  // the array keeps the previous call's contents: clear the part this call reduces over
  GB_CUDA(cudaMemsetAsync(scratch, 0, need * sizeof(float), (cudaStream_t)stream_));
  return gb_spmm_bwd_part(g, 0, 0, vals, gout, ldg, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, d, dx, lddx,
                          accumulate, scratch, dparams, stream_);
}

}  // extern "C"
