// plan.cu - graph plan construction: COO -> CSR/CSC (stable radix sorts), ordered degree sums,
// random-walk / GCN normalisation, per-edge permutation helpers.
//
// Replaces, for the hot path, the reference's per-evaluation COO -> dense conversion
// (src/function_laplacian_diffusion.py:39-58) and the scatter_add based normalisation of
// src/utils.py:177-233.  Ordering contract: oracle/graph.py (coo_to_csr, csr_to_csc).
// Runs once per graph, so the sorts use CUB's device radix sort (stable); everything per-evaluation
// is hand-written (spmm.cu, attention.cu, linear.cu).
#include <cub/cub.cuh>
#include "common.cuh"

namespace gb {

static inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

static int key_bits(int64_t n) {
  // keys are a*n + b < n*n
  int bits = 1;
  unsigned __int128 lim = (unsigned __int128)n * (unsigned __int128)n;
  while (bits < 64 && (((unsigned __int128)1) << bits) < lim) ++bits;
  return bits;
}

__global__ void make_keys_coo(const int64_t* __restrict__ row, const int64_t* __restrict__ col, int64_t e, int64_t n,
                              uint64_t* __restrict__ key, int32_t* __restrict__ val) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
    key[i] = (uint64_t)row[i] * (uint64_t)n + (uint64_t)col[i];
    val[i] = (int32_t)i;
  }
}

// sorted key = major*n + minor.  Writes minor[] and the segment pointer of major (boundary fill:
// ptr[r] = first slot whose major >= r), optionally the next-level key minor*n + major.
__global__ void split_keys(const uint64_t* __restrict__ key, int64_t e, int64_t n, int32_t* __restrict__ minor_out,
                           int32_t* __restrict__ ptr, uint64_t* __restrict__ next_key, int32_t* __restrict__ next_val) {
  for (int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; s < e; s += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t kcur = key[s];
    const int64_t major = (int64_t)(kcur / (uint64_t)n);
    const int64_t minor = (int64_t)(kcur - (uint64_t)major * (uint64_t)n);
    minor_out[s] = (int32_t)minor;
    const int64_t prev = s == 0 ? -1 : (int64_t)(key[s - 1] / (uint64_t)n);
    for (int64_t r = prev + 1; r <= major; ++r) ptr[r] = (int32_t)s;
    if (s == e - 1)
      for (int64_t r = major + 1; r <= n; ++r) ptr[r] = (int32_t)e;
    if (next_key) {
      next_key[s] = (uint64_t)minor * (uint64_t)n + (uint64_t)major;
      next_val[s] = (int32_t)s;
    }
  }
}

__global__ void fill_i32(int32_t* p, int64_t count, int32_t v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

static inline int grid_for(int64_t work, int block = 256) {
  int64_t g = (work + block - 1) / block;
  int64_t cap = (int64_t)sm_count() * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

struct PlanWs {
  uint64_t* key_a;
  uint64_t* key_b;
  int32_t* val_a;
  void* cub_tmp;
  size_t cub_bytes;
  size_t total;
};

static PlanWs carve(void* base, int32_t n, int64_t e) {
  PlanWs w;
  size_t cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, e, 0, 64);
  size_t sel_bytes = 0;
  cub::DeviceSelect::Flagged(nullptr, sel_bytes, (const int32_t*)nullptr, (const char*)nullptr, (int32_t*)nullptr,
                             (int32_t*)nullptr, (int64_t)n);
  if (sel_bytes > cub_bytes) cub_bytes = sel_bytes;
  size_t off = 0;
  char* b = (char*)base;
  w.key_a = (uint64_t*)(b + off);
  off += align_up((size_t)e * 8);
  w.key_b = (uint64_t*)(b + off);
  off += align_up((size_t)e * 8);
  w.val_a = (int32_t*)(b + off);
  off += align_up(((size_t)e > (size_t)n + 1 ? (size_t)e : (size_t)n + 1) * 4);
  w.cub_tmp = (void*)(b + off);
  w.cub_bytes = cub_bytes;
  off += align_up(cub_bytes);
  w.total = off + 256;
  return w;
}

// ---- ordered degree ------------------------------------------------------------------------
__global__ void make_keys_idx(const int64_t* __restrict__ idx, int64_t e, uint64_t* __restrict__ key,
                              int32_t* __restrict__ val) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
    key[i] = (uint64_t)idx[i];
    val[i] = (int32_t)i;
  }
}

__global__ void seg_ptr_from_sorted(const uint64_t* __restrict__ key, int64_t e, int64_t n, int32_t* __restrict__ ptr) {
  for (int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; s < e; s += (int64_t)gridDim.x * blockDim.x) {
    const int64_t cur = (int64_t)key[s];
    const int64_t prev = s == 0 ? -1 : (int64_t)key[s - 1];
    for (int64_t r = prev + 1; r <= cur; ++r) ptr[r] = (int32_t)s;
    if (s == e - 1)
      for (int64_t r = cur + 1; r <= n; ++r) ptr[r] = (int32_t)e;
  }
}

// one thread per segment, strictly sequential fp32 sum in original edge order (matches the
// order of torch's CPU scatter_add, hence the oracle's bits)
__global__ void ordered_segment_sum(const int32_t* __restrict__ ptr, const int32_t* __restrict__ perm,
                                    const float* __restrict__ w, int32_t n, float* __restrict__ deg) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int p = ptr[i]; p < ptr[i + 1]; ++p) acc = __fadd_rn(acc, w ? w[perm[p]] : 1.0f);
    deg[i] = acc;
  }
}

__global__ void rw_normalize_kernel(const int64_t* __restrict__ idx, const float* __restrict__ w,
                                    const float* __restrict__ deg, int64_t e, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
    const float inv = __frcp_rn(deg[idx[i]]);                 // deg.pow_(-1)
    out[i] = __fmul_rn(w ? w[i] : 1.0f, inv);
  }
}

__global__ void gcn_normalize_kernel(const int64_t* __restrict__ row, const int64_t* __restrict__ col,
                                     const float* __restrict__ w, const float* __restrict__ deg, int64_t e,
                                     float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
    float dr = __frcp_rn(__fsqrt_rn(deg[row[i]]));            // deg.pow_(-0.5)
    float dc = __frcp_rn(__fsqrt_rn(deg[col[i]]));
    if (isinf(dr)) dr = 0.f;
    if (isinf(dc)) dc = 0.f;
    out[i] = __fmul_rn(__fmul_rn(dr, w ? w[i] : 1.0f), dc);
  }
}

// ---- per-edge plumbing ---------------------------------------------------------------------
__global__ void gather_f32_kernel(const float* __restrict__ src, const int32_t* __restrict__ perm, int64_t e,
                                  float* __restrict__ dst) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = src[perm[i]];
}

__global__ void head_mean_gather_kernel(const float* __restrict__ att, int heads, const int32_t* __restrict__ perm,
                                        int64_t e, float* __restrict__ out) {
  const float hf = (float)heads;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
    const float* a = att + (int64_t)perm[i] * heads;
    float acc = 0.f;
    for (int h = 0; h < heads; ++h) acc += a[h];
    out[i] = acc / hf;      // torch mean = sum / H
  }
}

__global__ void scatter_rows_kernel(const float* __restrict__ src, int width, const int32_t* __restrict__ perm,
                                    int64_t e, float* __restrict__ dst) {
  const int64_t total = e * width;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = i / width;
    const int c = (int)(i - s * width);
    dst[(int64_t)perm[s] * width + c] = src[i];
  }
}

__global__ void flag_long_segments(const int32_t* __restrict__ ptr, int32_t n, int32_t threshold,
                                   char* __restrict__ flags, int32_t* __restrict__ ids) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    flags[i] = (ptr[i + 1] - ptr[i]) > threshold ? 1 : 0;
    ids[i] = i;
  }
}

}  // namespace gb

using namespace gb;

extern "C" {

size_t gb_plan_workspace_bytes(int32_t n, int64_t e) {
  if (n <= 0 || e < 0) return 0;
  PlanWs w = carve(nullptr, n, e > 0 ? e : 1);
  // flags + ids for gb_plan_long_segments also fit: n bytes + 4n bytes <= key_a region when e >= n; be safe
  size_t extra = align_up((size_t)n) + align_up((size_t)n * 4) + 64;
  return w.total + extra;
}

int gb_plan_build(const int64_t* row, const int64_t* col, int64_t e, int32_t n, int32_t* rowptr, int32_t* colidx,
                  int32_t* eperm, int32_t* colptr, int32_t* rowidx, int32_t* tperm, void* workspace,
                  size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(n > 0, "gb_plan_build: n must be positive");
  GB_REQUIRE(e >= 0 && e < (int64_t)0x7fffffff, "gb_plan_build: edge count %lld does not fit int32 slots", (long long)e);
  GB_REQUIRE(rowptr && colptr, "gb_plan_build: null output");
  if (e == 0) {
    LaunchScope ls(K_PLAN, stream);
    fill_i32<<<grid_for(n + 1), 256, 0, stream>>>(rowptr, n + 1, 0);
    fill_i32<<<grid_for(n + 1), 256, 0, stream>>>(colptr, n + 1, 0);
    return check_launch("fill_i32");
  }
  GB_REQUIRE(row && col && colidx && eperm && rowidx && tperm && workspace, "gb_plan_build: null pointer");
  if (workspace_bytes < gb_plan_workspace_bytes(n, e)) {
    set_error("gb_plan_build: workspace too small (%zu < %zu)", workspace_bytes, gb_plan_workspace_bytes(n, e));
    return GB_ERR_WORKSPACE;
  }
  PlanWs w = carve(workspace, n, e);
  const int bits = key_bits(n);
  LaunchScope ls(K_PLAN, stream);
  make_keys_coo<<<grid_for(e), 256, 0, stream>>>(row, col, e, n, w.key_a, w.val_a);
  size_t tmp = w.cub_bytes;
  GB_CUDA(cub::DeviceRadixSort::SortPairs(w.cub_tmp, tmp, w.key_a, w.key_b, w.val_a, eperm, e, 0, bits, stream));
  // key_b sorted by (row, col): colidx, rowptr, and the transpose keys (col, row) -> key_a, val_a
  split_keys<<<grid_for(e), 256, 0, stream>>>(w.key_b, e, n, colidx, rowptr, w.key_a, w.val_a);
  tmp = w.cub_bytes;
  GB_CUDA(cub::DeviceRadixSort::SortPairs(w.cub_tmp, tmp, w.key_a, w.key_b, w.val_a, tperm, e, 0, bits, stream));
  split_keys<<<grid_for(e), 256, 0, stream>>>(w.key_b, e, n, rowidx, colptr, nullptr, nullptr);
  return check_launch("gb_plan_build");
}

int gb_plan_long_segments(const int32_t* ptr, int32_t n, int32_t threshold, int32_t* out, int32_t cap,
                          int32_t* count_host, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(ptr && count_host && workspace && n > 0, "gb_plan_long_segments: null pointer");
  size_t sel_bytes = 0;
  cub::DeviceSelect::Flagged(nullptr, sel_bytes, (const int32_t*)nullptr, (const char*)nullptr, (int32_t*)nullptr,
                             (int32_t*)nullptr, (int64_t)n);
  char* b = (char*)workspace;
  size_t off = 0;
  char* flags = b + off;
  off += align_up((size_t)n);
  int32_t* ids = (int32_t*)(b + off);
  off += align_up((size_t)n * 4);
  int32_t* sel = (int32_t*)(b + off);
  off += align_up((size_t)n * 4);
  int32_t* d_count = (int32_t*)(b + off);
  off += 256;
  void* tmp = b + off;
  off += align_up(sel_bytes);
  if (off > workspace_bytes) {
    set_error("gb_plan_long_segments: workspace too small (%zu < %zu)", workspace_bytes, off);
    return GB_ERR_WORKSPACE;
  }
  LaunchScope ls(K_PLAN, stream);
  flag_long_segments<<<grid_for(n), 256, 0, stream>>>(ptr, n, threshold, flags, ids);
  GB_CUDA(cub::DeviceSelect::Flagged(tmp, sel_bytes, ids, flags, sel, d_count, (int64_t)n, stream));
  int32_t count = 0;
  GB_CUDA(cudaMemcpyAsync(&count, d_count, sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  GB_CUDA(cudaStreamSynchronize(stream));
  *count_host = count;
  if (count > 0 && out) {
    GB_REQUIRE(count <= cap, "gb_plan_long_segments: %d long segments exceed capacity %d", count, cap);
    GB_CUDA(cudaMemcpyAsync(out, sel, (size_t)count * 4, cudaMemcpyDeviceToDevice, stream));
  }
  return check_launch("gb_plan_long_segments");
}

struct DegWs {
  uint64_t *key_a, *key_b;
  int32_t *val_a, *val_b, *ptr;
  void* cub_tmp;
  size_t cub_bytes, total;
};

static DegWs carve_deg(void* base, int32_t n, int64_t e) {
  DegWs w;
  size_t cub_bytes = 0;
  cub::DoubleBuffer<uint64_t> dk(nullptr, nullptr);
  cub::DoubleBuffer<int32_t> dv(nullptr, nullptr);
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, dk, dv, e, 0, 64);
  char* b = (char*)base;
  size_t off = 0;
  w.key_a = (uint64_t*)(b + off); off += align_up((size_t)e * 8);
  w.key_b = (uint64_t*)(b + off); off += align_up((size_t)e * 8);
  w.val_a = (int32_t*)(b + off);  off += align_up((size_t)e * 4);
  w.val_b = (int32_t*)(b + off);  off += align_up((size_t)e * 4);
  w.ptr = (int32_t*)(b + off);    off += align_up(((size_t)n + 1) * 4);
  w.cub_tmp = (void*)(b + off);   off += align_up(cub_bytes);
  w.cub_bytes = cub_bytes;
  w.total = off + 256;
  return w;
}

size_t gb_degree_workspace_bytes(int32_t n, int64_t e) {
  if (n <= 0 || e <= 0) return 256;
  return carve_deg(nullptr, n, e).total;
}

int gb_degree(const int64_t* idx, const float* w, int64_t e, int32_t n, float* deg, void* workspace,
              size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(n > 0 && deg, "gb_degree: bad arguments");
  GB_REQUIRE(e >= 0 && e < (int64_t)0x7fffffff, "gb_degree: edge count out of range");
  if (e == 0) {
    GB_CUDA(cudaMemsetAsync(deg, 0, (size_t)n * 4, stream));
    return GB_OK;
  }
  GB_REQUIRE(idx && workspace, "gb_degree: null pointer");
  if (workspace_bytes < gb_degree_workspace_bytes(n, e)) {
    set_error("gb_degree: workspace too small (%zu < %zu)", workspace_bytes, gb_degree_workspace_bytes(n, e));
    return GB_ERR_WORKSPACE;
  }
  DegWs ws = carve_deg(workspace, n, e);
  int bits = 1;
  while (bits < 63 && ((int64_t)1 << bits) < (int64_t)n) ++bits;
  LaunchScope ls(K_PLAN, stream);
  make_keys_idx<<<grid_for(e), 256, 0, stream>>>(idx, e, ws.key_a, ws.val_a);
  cub::DoubleBuffer<uint64_t> dk(ws.key_a, ws.key_b);
  cub::DoubleBuffer<int32_t> dv(ws.val_a, ws.val_b);
  size_t tmp = ws.cub_bytes;
  // radix sort is stable: inside one segment the edges keep their original relative order
  GB_CUDA(cub::DeviceRadixSort::SortPairs(ws.cub_tmp, tmp, dk, dv, e, 0, bits, stream));
  seg_ptr_from_sorted<<<grid_for(e), 256, 0, stream>>>(dk.Current(), e, n, ws.ptr);
  ordered_segment_sum<<<grid_for(n, 128), 128, 0, stream>>>(ws.ptr, dv.Current(), w, n, deg);
  return check_launch("gb_degree");
}

int gb_rw_normalize(const int64_t* idx, const float* w, const float* deg, int64_t e, float* w_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (e == 0) return GB_OK;
  GB_REQUIRE(idx && deg && w_out, "gb_rw_normalize: null pointer");
  LaunchScope ls(K_PLAN, stream);
  rw_normalize_kernel<<<grid_for(e), 256, 0, stream>>>(idx, w, deg, e, w_out);
  return check_launch("rw_normalize_kernel");
}

int gb_gcn_normalize(const int64_t* row, const int64_t* col, const float* w, const float* deg, int64_t e,
                     float* w_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (e == 0) return GB_OK;
  GB_REQUIRE(row && col && deg && w_out, "gb_gcn_normalize: null pointer");
  LaunchScope ls(K_PLAN, stream);
  gcn_normalize_kernel<<<grid_for(e), 256, 0, stream>>>(row, col, w, deg, e, w_out);
  return check_launch("gcn_normalize_kernel");
}

int gb_gather_f32(const float* src, const int32_t* perm, int64_t e, float* dst, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (e == 0) return GB_OK;
  GB_REQUIRE(src && perm && dst, "gb_gather_f32: null pointer");
  LaunchScope ls(K_EDGE_UTIL, stream);
  gather_f32_kernel<<<grid_for(e), 256, 0, stream>>>(src, perm, e, dst);
  return check_launch("gather_f32_kernel");
}

int gb_head_mean_gather(const float* att_coo, int32_t heads, const int32_t* eperm, int64_t e, float* abar_csr,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (e == 0) return GB_OK;
  GB_REQUIRE(att_coo && eperm && abar_csr && heads > 0, "gb_head_mean_gather: bad arguments");
  LaunchScope ls(K_EDGE_UTIL, stream);
  head_mean_gather_kernel<<<grid_for(e), 256, 0, stream>>>(att_coo, heads, eperm, e, abar_csr);
  return check_launch("head_mean_gather_kernel");
}

int gb_scatter_rows_f32(const float* src_csr, int32_t width, const int32_t* eperm, int64_t e, float* dst_coo,
                        void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (e == 0) return GB_OK;
  GB_REQUIRE(src_csr && eperm && dst_coo && width > 0, "gb_scatter_rows_f32: bad arguments");
  LaunchScope ls(K_EDGE_UTIL, stream);
  scatter_rows_kernel<<<grid_for(e * width), 256, 0, stream>>>(src_csr, width, eperm, e, dst_coo);
  return check_launch("scatter_rows_kernel");
}

}  // extern "C"
