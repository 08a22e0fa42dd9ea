// linear.cu - K3: the Q/K projection (nn.Linear, src/function_transformer_attention.py:224-225) and
// its VJP, as exact-fp32 CUDA-core tiled GEMMs.  This is the reference-accuracy path (plain FFMA,
// fp32 accumulate) that every other projection path (tcgen05 3xTF32, linear_tc.cu) is validated
// against; it also serves shapes the tensor-core path does not take (odd d, tiny graphs).
//
//   fwd   : out[i,j] = sum_c x[i,c] w[j,c] + b[j]           M = n, N = nout, K = d
//   bwd_x : dx[i,c] (+)= sum_j dy[i,j] w[j,c]               M = n, N = d,    K = nout
//   bwd_w : dw[j,c] = sum_i dy[i,j] x[i,c], db[j] = sum_i dy[i,j]   split over i, fixed-order combine
//
// Roofline: HBM for the streaming operand (x / dy, read once) at these K; CUDA-core FFMA otherwise.
#include "common.cuh"

namespace gb {

constexpr int BM = 64, BN = 64, BK = 16, GT = 256;

// C[m,j] (+)= sum_k A[m*lda + k] * B[k*sbk + j*sbj] (+ bias[j])
__global__ void __launch_bounds__(GT) sgemm_rowA(const float* __restrict__ A, int lda, int M, int K,
                                                 const float* __restrict__ B, int sbk, int sbj, int N,
                                                 const float* __restrict__ bias, float* __restrict__ C, int ldc,
                                                 int accumulate) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += BK) {
    // A tile: 64 rows x 16 k ; thread -> (row = t/16 + 16 i, k = t%16)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = threadIdx.x / 16 + 16 * i, kk = threadIdx.x % 16;
      const int gm = m0 + r, gk = k0 + kk;
      As[kk][r] = (gm < M && gk < K) ? __ldg(A + (size_t)gm * lda + gk) : 0.f;
    }
    // B tile: 16 k x 64 j ; orient the thread mapping along the contiguous axis of B
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int kk, jj;
      if (sbj == 1) {
        jj = threadIdx.x % 64;
        kk = threadIdx.x / 64 + 4 * i;
      } else {
        kk = threadIdx.x % 16;
        jj = threadIdx.x / 16 + 16 * i;
      }
      const int gk = k0 + kk, gj = n0 + jj;
      Bs[kk][jj] = (gk < K && gj < N) ? __ldg(B + (size_t)gk * sbk + (size_t)gj * sbj) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int gm = m0 + ty * 4 + r;
    if (gm >= M) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int gj = n0 + tx * 4 + c;
      if (gj >= N) continue;
      float v = acc[r][c];
      if (bias) v += __ldg(bias + gj);
      float* p = C + (size_t)gm * ldc + gj;
      *p = accumulate ? (*p + v) : v;
    }
  }
}

// split-K over the node axis: part[z][j][c] = sum_{i in split z} dy[i,j] x[i,c]; pdb[z][j] = sum dy[i,j]
__global__ void __launch_bounds__(GT) sgemm_tn_split(const float* __restrict__ dy, int lddy, const float* __restrict__ x,
                                                     int ldx, int n, int nout, int d, int rows_per_split,
                                                     float* __restrict__ part, float* __restrict__ pdb) {
  __shared__ __align__(16) float As[BK][BM + 4];   // dy tile  [i][j]
  __shared__ __align__(16) float Bs[BK][BN + 4];   // x tile   [i][c]
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int j0 = blockIdx.x * BM, c0 = blockIdx.y * BN;
  const int i_begin = blockIdx.z * rows_per_split;
  const int i_end = min(n, i_begin + rows_per_split);
  float acc[4][4] = {};
  float colsum[4] = {};
  for (int i0 = i_begin; i0 < i_end; i0 += BK) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int jj = threadIdx.x % 64, kk = threadIdx.x / 64 + 4 * t;
      const int gi = i0 + kk;
      As[kk][jj] = (gi < i_end && j0 + jj < nout) ? __ldg(dy + (size_t)gi * lddy + j0 + jj) : 0.f;
      Bs[kk][jj] = (gi < i_end && c0 + jj < d) ? __ldg(x + (size_t)gi * ldx + c0 + jj) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        colsum[r] += av[r];
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
      }
    }
    __syncthreads();
  }
  float* out = part + (size_t)blockIdx.z * nout * d;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int gj = j0 + ty * 4 + r;
    if (gj >= nout) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int gc = c0 + tx * 4 + c;
      if (gc < d) out[(size_t)gj * d + gc] = acc[r][c];
    }
    if (blockIdx.y == 0 && tx == 0) pdb[(size_t)blockIdx.z * nout + gj] = colsum[r];
  }
}

__global__ void reduce_splits(const float* __restrict__ part, const float* __restrict__ pdb, int splits, int nout, int d,
                              float* __restrict__ dw, float* __restrict__ db) {
  const int total = nout * d;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total + nout; i += gridDim.x * blockDim.x) {
    float acc = 0.f;
    if (i < total) {
      for (int z = 0; z < splits; ++z) acc += part[(size_t)z * total + i];
      dw[i] = acc;
    } else if (db != nullptr) {
      const int j = i - total;
      for (int z = 0; z < splits; ++z) acc += pdb[(size_t)z * nout + j];
      db[j] = acc;
    }
  }
}

static int choose_splits(int n, int nout, int d) {
  const int tiles = ((nout + BM - 1) / BM) * ((d + BN - 1) / BN);
  int s = (2 * sm_count() + tiles - 1) / tiles;
  const int max_s = (n + 4 * BK - 1) / (4 * BK);
  if (s > max_s) s = max_s;
  if (s < 1) s = 1;
  return s;
}

}  // namespace gb

namespace gb {
struct TcArgs {
  const float* a; int64_t lda; const float* w; int64_t ldw; const float* bias; float* out; int64_t ldo;
  int32_t rows, no, k; int32_t trans_w, accumulate, stages; int32_t k_valid, a_vec, w_vec;
};
bool tc_gemm_sliced_supported(const void* out, int64_t ldo, int no);
int tc_gemm_sliced_launch(const float* x, int64_t ldx, int rows, int k, const float* w, int64_t ldw, const float* bias,
                          int no, float* out, int64_t ldo, int kernel_id, cudaStream_t stream);
bool tc_gemm_supported(const void* a, int64_t lda, const void* w, int64_t ldw, const void* out, int64_t ldo, int no,
                       int k, int trans_w);
int tc_gemm_launch(const TcArgs& a, int kernel_id, cudaStream_t stream);
bool tc_gemm_tn_supported(const void* dy, int64_t lddy, const void* x, int64_t ldx, int nout, int d);
size_t tc_gemm_tn_workspace_bytes(int d);
int tc_gemm_tn_launch(const float* dy, int64_t lddy, const float* x, int64_t ldx, int rows, int d, float* dw, float* db,
                      void* workspace, cudaStream_t stream);
}  // namespace gb

using namespace gb;

extern "C" {

int gb_linear_tc_eligible(const float* a, int32_t lda, const float* w, int32_t ldw, const float* out, int32_t ldo,
                          int32_t nout, int32_t k, int32_t trans_w) {
  return tc_gemm_supported(a, lda, w, ldw, out, ldo, nout, k, trans_w) ? 1 : 0;
}

int gb_linear_fwd(const float* x, int32_t ldx, int32_t n, int32_t d, const float* w, const float* b, int32_t nout,
                  float* out, int32_t ldo, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(x && w && out && n > 0 && d > 0 && nout > 0, "gb_linear_fwd: bad arguments");
  GB_REQUIRE(ldx >= d && ldo >= nout, "gb_linear_fwd: leading dimension too small");
  if (tc_gemm_supported(x, ldx, w, d, out, ldo, nout, d, 0)) {       // tcgen05 3xTF32 path (linear_tc.cu)
    TcArgs t{x, ldx, w, d, b, out, ldo, n, nout, d, 0, 0, 0};
    return tc_gemm_launch(t, K_LINEAR_FWD, stream);
  }
  // wide / unaligned reductions (the encoder, src/GNN.py:30-38: [N, 1433] x [1433, d]): K slices on the same kernel
  // (from 12 288 rows: below that the slices are launch-bound - 22 row tiles on 148 SMs for Cora - and the CUDA-core tiles win,
  // measured 122 vs 158 us for [2708, 1433] x [1433, 64], 309 vs 415 us for Citeseer's [3327, 3703] x [3703, 80])
  if (d >= 32 && n >= 12288 && tc_gemm_sliced_supported(out, ldo, nout))
    return tc_gemm_sliced_launch(x, ldx, n, d, w, d, b, nout, out, ldo, K_LINEAR_FWD, stream);
  dim3 grid((n + BM - 1) / BM, (nout + BN - 1) / BN);
  LaunchScope ls(K_LINEAR_FWD, stream);
  sgemm_rowA<<<grid, GT, 0, stream>>>(x, ldx, n, d, w, 1, d, nout, b, out, ldo, 0);
  return check_launch("gb_linear_fwd");
}

int gb_linear_bwd_x(const float* dy, int32_t lddy, int32_t n, int32_t nout, const float* w, int32_t d, float* dx,
                    int32_t lddx, int32_t accumulate, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(dy && w && dx && n > 0 && d > 0 && nout > 0, "gb_linear_bwd_x: bad arguments");
  GB_REQUIRE(lddy >= nout && lddx >= d, "gb_linear_bwd_x: leading dimension too small");
  if (tc_gemm_supported(dy, lddy, w, d, dx, lddx, d, nout, 1)) {      // dx = dy . W: W [nout, d] is B^T, staged transposed
    TcArgs t{dy, lddy, w, d, nullptr, dx, lddx, n, d, nout, 1, accumulate != 0, 0};
    return tc_gemm_launch(t, K_LINEAR_BWD_X, stream);
  }
  dim3 grid((n + BM - 1) / BM, (d + BN - 1) / BN);
  LaunchScope ls(K_LINEAR_BWD_X, stream);
  sgemm_rowA<<<grid, GT, 0, stream>>>(dy, lddy, n, nout, w, d, 1, d, nullptr, dx, lddx, accumulate);
  return check_launch("gb_linear_bwd_x");
}

size_t gb_linear_bwd_w_workspace_bytes(int32_t n, int32_t d, int32_t nout) {
  if (n <= 0 || d <= 0 || nout <= 0) return 0;
  const size_t s = (size_t)choose_splits(n, nout, d);
  size_t bytes = s * ((size_t)nout * d + nout) * sizeof(float) + 256;
  if (nout == 128 && d % 32 == 0 && d <= 128) {               // tensor-core path: one partial per SM
    const size_t tc = tc_gemm_tn_workspace_bytes(d);
    if (tc > bytes) bytes = tc;
  }
  return bytes;
}

int gb_linear_bwd_w(const float* dy, int32_t lddy, const float* x, int32_t ldx, int32_t n, int32_t d, int32_t nout,
                    float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(dy && x && dw && workspace && n > 0 && d > 0 && nout > 0, "gb_linear_bwd_w: bad arguments");
  if (workspace_bytes < gb_linear_bwd_w_workspace_bytes(n, d, nout)) {
    set_error("gb_linear_bwd_w: workspace too small");
    return GB_ERR_WORKSPACE;
  }
  if (db != nullptr && tc_gemm_tn_supported(dy, lddy, x, ldx, nout, d))      // tcgen05 3xTF32 path (linear_tc.cu)
    return tc_gemm_tn_launch(dy, lddy, x, ldx, n, d, dw, db, workspace, stream);
  const int splits = choose_splits(n, nout, d);
  int rows = (n + splits - 1) / splits;
  rows = (rows + BK - 1) / BK * BK;
  float* part = (float*)workspace;
  float* pdb = part + (size_t)splits * nout * d;
  dim3 grid((nout + BM - 1) / BM, (d + BN - 1) / BN, splits);
  {
    LaunchScope ls(K_LINEAR_BWD_W, stream);
    sgemm_tn_split<<<grid, GT, 0, stream>>>(dy, lddy, x, ldx, n, nout, d, rows, part, pdb);
  }
  {
    LaunchScope ls(K_LINEAR_BWD_W_REDUCE, stream);
    const int total = nout * d + nout;
    reduce_splits<<<(total + 255) / 256, 256, 0, stream>>>(part, pdb, splits, nout, d, dw, db);
  }
  return check_launch("gb_linear_bwd_w");
}

}  // extern "C"
