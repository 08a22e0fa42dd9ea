// vec.cuh - 1/2/4-wide fp32 vector helpers so every row kernel has one body for the 128-bit path
// (d % 4 == 0, 16-byte aligned), the 64-bit path and the scalar path (odd d).
#pragma once
#include "common.cuh"

namespace gb {

template <int V>
struct Vec;

template <>
struct Vec<4> {
  float4 v;
  __device__ __forceinline__ static Vec zero() { return Vec{make_float4(0.f, 0.f, 0.f, 0.f)}; }
  __device__ __forceinline__ static Vec load(const float* p) { return Vec{__ldg(reinterpret_cast<const float4*>(p))}; }
  __device__ __forceinline__ static Vec load_stream(const float* p) { return Vec{ld_stream4(p)}; }
  __device__ __forceinline__ static Vec load_rw(const float* p) { return Vec{*reinterpret_cast<const float4*>(p)}; }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
  __device__ __forceinline__ void store_stream(float* p) const { st_stream4(p, v); }
  __device__ __forceinline__ void fma(float a, const Vec& b) { fma4(v, a, b.v); }
  __device__ __forceinline__ float dot(const Vec& b) const { return dot4(v, b.v); }
  __device__ __forceinline__ float sum() const { return (v.x + v.y) + (v.z + v.w); }
  template <class F>
  __device__ __forceinline__ static Vec map2(const Vec& a, const Vec& b, F f) {
    return Vec{make_float4(f(a.v.x, b.v.x), f(a.v.y, b.v.y), f(a.v.z, b.v.z), f(a.v.w, b.v.w))};
  }
};

template <>
struct Vec<2> {
  float2 v;
  __device__ __forceinline__ static Vec zero() { return Vec{make_float2(0.f, 0.f)}; }
  __device__ __forceinline__ static Vec load(const float* p) { return Vec{__ldg(reinterpret_cast<const float2*>(p))}; }
  __device__ __forceinline__ static Vec load_stream(const float* p) { return load(p); }
  __device__ __forceinline__ static Vec load_rw(const float* p) { return Vec{*reinterpret_cast<const float2*>(p)}; }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float2*>(p) = v; }
  __device__ __forceinline__ void store_stream(float* p) const { store(p); }
  __device__ __forceinline__ void fma(float a, const Vec& b) {
    v.x = fmaf(a, b.v.x, v.x);
    v.y = fmaf(a, b.v.y, v.y);
  }
  __device__ __forceinline__ float dot(const Vec& b) const { return fmaf(v.y, b.v.y, v.x * b.v.x); }
  __device__ __forceinline__ float sum() const { return v.x + v.y; }
  template <class F>
  __device__ __forceinline__ static Vec map2(const Vec& a, const Vec& b, F f) {
    return Vec{make_float2(f(a.v.x, b.v.x), f(a.v.y, b.v.y))};
  }
};

template <>
struct Vec<1> {
  float v;
  __device__ __forceinline__ static Vec zero() { return Vec{0.f}; }
  __device__ __forceinline__ static Vec load(const float* p) { return Vec{__ldg(p)}; }
  __device__ __forceinline__ static Vec load_stream(const float* p) { return load(p); }
  __device__ __forceinline__ static Vec load_rw(const float* p) { return Vec{*p}; }
  __device__ __forceinline__ void store(float* p) const { *p = v; }
  __device__ __forceinline__ void store_stream(float* p) const { *p = v; }
  __device__ __forceinline__ void fma(float a, const Vec& b) { v = fmaf(a, b.v, v); }
  __device__ __forceinline__ float dot(const Vec& b) const { return v * b.v; }
  __device__ __forceinline__ float sum() const { return v; }
  template <class F>
  __device__ __forceinline__ static Vec map2(const Vec& a, const Vec& b, F f) {
    return Vec{f(a.v, b.v)};
  }
};

// widest vector width usable for a row-major fp32 matrix (pointer alignment, leading dim, width)
inline int vec_width(const void* p, int ld, int d) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  if (a % 16 == 0 && ld % 4 == 0 && d % 4 == 0) return 4;
  if (a % 8 == 0 && ld % 2 == 0 && d % 2 == 0) return 2;
  return 1;
}
inline int min_int(int a, int b) { return a < b ? a : b; }

// lanes-per-row / chunks-per-lane choice for a row of `dv` vector chunks (dv <= 128)
struct RowShape {
  int lpr, cpl;
};
inline RowShape row_shape(int dv) {
  if (dv <= 8) return {8, 1};
  if (dv <= 16) return {16, 1};
  if (dv == 24) return {8, 3};
  if (dv <= 32) return {32, 1};
  if (dv <= 64) return {32, 2};
  if (dv <= 96) return {32, 3};
  return {32, 4};
}

}  // namespace gb
