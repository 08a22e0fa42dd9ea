// common.cuh - shared host/device helpers of libgrandb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "../../include/grandb200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libgrandb200 is written for sm_100a (B200) only"
#endif

namespace gb {

// ---- error reporting -----------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t err, const char* what, const char* file, int line);

#define GB_CUDA(call)                                                      \
  do {                                                                     \
    cudaError_t _e = (call);                                               \
    if (_e != cudaSuccess) return gb::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define GB_REQUIRE(cond, ...)                 \
  do {                                        \
    if (!(cond)) {                            \
      gb::set_error(__VA_ARGS__);             \
      return GB_ERR_INVALID;                  \
    }                                         \
  } while (0)

// ---- kernel ids for profiling / launch accounting ----------------------------------------
enum KernelId {
  K_SPMM_FWD = 0,
  K_SPMM_FWD_LONG,
  K_SPMM_BWD,
  K_SPMM_BWD_LONG,
  K_REDUCE_PARTIALS,
  K_LINEAR_FWD,
  K_LINEAR_BWD_X,
  K_LINEAR_BWD_W,
  K_LINEAR_BWD_W_REDUCE,
  K_EDGE_SOFTMAX_FWD,
  K_EDGE_SOFTMAX_FWD_LONG,
  K_SDDMM,
  K_EDGE_SOFTMAX_BWD,
  K_EDGE_SCORE_BWD_Q,
  K_EDGE_SCORE_BWD_K,
  K_PLAN,
  K_EDGE_UTIL,
  K_FUSED_ATT_SPMM_FWD,
  K_FUSED_ATT_BWD_ROW,
  K_FUSED_ATT_BWD_COL,
  K_LINEAR_TC,
  K_STAGE_AXPY,
  K_NL_FWD,
  K_NL_FWD_LONG,
  K_NL_ADJ_ROWS,
  K_NL_ADJ_LONG,
  K_GAT_SOFTMAX,
  K_DECODE_EVAL,
  K_NUM_KERNELS
};

// RAII-ish bracket around a launch: counts it, and (when profiling is on) records a pair of
// CUDA events on the launching stream.
struct LaunchScope {
  LaunchScope(int kernel_id, cudaStream_t stream);
  ~LaunchScope();
  int id;
  cudaStream_t stream;
  void* rec;
};

int check_launch(const char* name);   // cudaGetLastError -> gb_status

int sm_count();
// SMs the persistent row kernels occupy: sm_count() minus GRANDB200_SM_RESERVE (default 0).  A row-partitioned run
// that overlaps NCCL collectives with these kernels leaves a few SMs to NCCL's kernels this way.
int persistent_sms();

// fork / join of the per-device side stream (runtime.cu); side_fork returns `main` when unavailable
cudaStream_t side_fork(cudaStream_t main);
int side_join(cudaStream_t main, cudaStream_t side);

// ---- device helpers --------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void fma4(float4& acc, float a, const float4& v) {
  acc.x = fmaf(a, v.x, acc.x);
  acc.y = fmaf(a, v.y, acc.y);
  acc.z = fmaf(a, v.z, acc.z);
  acc.w = fmaf(a, v.w, acc.w);
}
__device__ __forceinline__ float dot4(const float4& a, const float4& b) {
  return fmaf(a.w, b.w, fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)));
}
__device__ __forceinline__ float sigmoidf_(float v) { return 1.0f / (1.0f + expf(-v)); }

// streaming (read-once / write-once) accesses: keep them out of L1 so gathered rows stay cached
__device__ __forceinline__ void st_stream4(float* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 ld_stream4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ int ld_stream_i32(const int* p) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream_f32(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// sum over the LPR lanes of an aligned sub-warp group, result in every lane of the group
template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// fixed-order block sum of one float per thread; result valid in thread 0. smem: >= 32 floats.
__device__ __forceinline__ float block_sum(float v, float* smem) {
  v = group_sum<32>(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    r = lane < nw ? smem[lane] : 0.f;
    r = group_sum<32>(r);
  }
  return r;
}

}  // namespace gb
