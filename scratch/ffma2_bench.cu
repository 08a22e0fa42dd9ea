// microbenchmark: FFMA vs FFMA2 issue throughput on sm_100a (scratch, not part of the library)
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ void ffma2(float2& acc, float2 a, float b) {
  float2 bb = make_float2(b, b);
  unsigned long long xa = *reinterpret_cast<unsigned long long*>(&a), yb = *reinterpret_cast<unsigned long long*>(&bb),
                     zc = *reinterpret_cast<unsigned long long*>(&acc), r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(xa), "l"(yb), "l"(zc));
  acc = *reinterpret_cast<float2*>(&r);
}
template <int MODE> __global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
  float2 acc[8];
  for (int j = 0; j < 8; ++j) acc[j] = make_float2(threadIdx.x * 0.001f + j, j);
  float2 m = make_float2(1.0001f, 0.9999f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 0) { acc[j].x = fmaf(acc[j].x, m.x, s); acc[j].y = fmaf(acc[j].y, m.y, s); }
        else ffma2(acc[j], m, s);
      }
  }
  float t = 0; for (int j = 0; j < 8; ++j) t += acc[j].x + acc[j].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int mode = 0; mode < 2; ++mode) for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(a);
    if (mode == 0) k<0><<<148 * 8, 256>>>(out, 20000, 0.5f); else k<1><<<148 * 8, 256>>>(out, 20000, 0.5f);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double fma = 148.0 * 8 * 256 * 20000.0 * 4 * 8 * 2;
    printf("mode %d: %.3f ms  %.1f TFLOP/s (fp32 FMA = 2 flop)\n", mode, ms, fma * 2 / ms / 1e9);
  }
  return 0;
}
