// stage.cu - Runge-Kutta stage combination  out = y + sum_j c[j] * k[j]  in ONE pass over the state
// (torchdiffeq forms it as a chain of axpy kernels, one [N,d] read-modify-write per term).
// Roofline: HBM, (nk + 2) * 4 bytes per element.
#include "common.cuh"

namespace gb {

struct CombineArgs {
  const float* y;
  const float* k[7];
  float c[7];
  int32_t nk;
  float* out;
  int64_t count4;     // number of float4 elements
  int64_t count;      // total floats
};

template <int NK>
__global__ void __launch_bounds__(256) stage_combine_kernel(const CombineArgs a) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.count4; i += stride) {
    float4 acc = ld_stream4(a.y + 4 * i);
#pragma unroll
    for (int j = 0; j < NK; ++j) fma4(acc, a.c[j], ld_stream4(a.k[j] + 4 * i));
    st_stream4(a.out + 4 * i, acc);
  }
  // scalar tail (count % 4 elements)
  const int64_t tail0 = a.count4 * 4;
  const int64_t t = tail0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < a.count) {
    float acc = a.y[t];
#pragma unroll
    for (int j = 0; j < NK; ++j) acc = fmaf(a.c[j], a.k[j][t], acc);
    a.out[t] = acc;
  }
}

// The same combination written to up to 8 destinations: the caller's own buffer and the matching slab of the gathered
// state matrix in every peer's memory (NVLink peer pointers / symmetric memory) - the all-gather of the NEXT stage's
// state fused into the kernel that produces it (DESIGN.md "Multi-GPU", push-style exchange).  Loads are
// read once from local HBM, every destination gets one 16-byte store per element group; remote stores are posted
// writes, so the kernel runs at local-HBM speed while the copies drain over NVLink.
struct CombineMultiArgs {
  CombineArgs base;
  float* outs[8];
  int32_t n_out;
};

template <int NK>
__global__ void __launch_bounds__(256) stage_combine_multi_kernel(const CombineMultiArgs m) {
  const CombineArgs& a = m.base;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.count4; i += stride) {
    float4 acc = ld_stream4(a.y + 4 * i);
#pragma unroll
    for (int j = 0; j < NK; ++j) fma4(acc, a.c[j], ld_stream4(a.k[j] + 4 * i));
    for (int o = 0; o < m.n_out; ++o) st_stream4(m.outs[o] + 4 * i, acc);
  }
  const int64_t tail0 = a.count4 * 4;
  const int64_t t = tail0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < a.count) {
    float acc = a.y[t];
#pragma unroll
    for (int j = 0; j < NK; ++j) acc = fmaf(a.c[j], a.k[j][t], acc);
    for (int o = 0; o < m.n_out; ++o) m.outs[o][t] = acc;
  }
}

}  // namespace gb

using namespace gb;

extern "C" int gb_stage_combine(const float* y, const float* const* k_host, const float* c_host, int32_t nk,
                                int64_t count, float* out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(y && out && count >= 0 && nk >= 0 && nk <= 7, "gb_stage_combine: bad arguments");
  GB_REQUIRE(nk == 0 || (k_host && c_host), "gb_stage_combine: null k / c");
  if (count == 0) return GB_OK;
  CombineArgs a{};
  a.y = y; a.out = out; a.nk = nk; a.count = count;
  bool aligned = reinterpret_cast<uintptr_t>(y) % 16 == 0 && reinterpret_cast<uintptr_t>(out) % 16 == 0;
  for (int j = 0; j < nk; ++j) {
    GB_REQUIRE(k_host[j] != nullptr, "gb_stage_combine: null k[%d]", j);
    a.k[j] = k_host[j];
    a.c[j] = c_host[j];
    aligned = aligned && reinterpret_cast<uintptr_t>(k_host[j]) % 16 == 0;
  }
  a.count4 = aligned ? count / 4 : 0;
  const int64_t work = a.count4 > 0 ? a.count4 : count;
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (a.count4 > 0 && blocks > cap) blocks = cap;
  if (a.count4 == 0) blocks = (count + 255) / 256;           // unaligned: every element goes through the tail path
  LaunchScope ls(K_STAGE_AXPY, stream);
  switch (nk) {
    case 0: stage_combine_kernel<0><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 1: stage_combine_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 2: stage_combine_kernel<2><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 3: stage_combine_kernel<3><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 4: stage_combine_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 5: stage_combine_kernel<5><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 6: stage_combine_kernel<6><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    default: stage_combine_kernel<7><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
  }
  return check_launch("stage_combine");
}

extern "C" int gb_stage_combine_multi(const float* y, const float* const* k_host, const float* c_host, int32_t nk,
                                      int64_t count, float* const* outs_host, int32_t n_out, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(y && outs_host && count >= 0 && nk >= 0 && nk <= 7 && n_out >= 1 && n_out <= 8,
             "gb_stage_combine_multi: bad arguments (nk <= 7, 1 <= n_out <= 8)");
  GB_REQUIRE(nk == 0 || (k_host && c_host), "gb_stage_combine_multi: null k / c");
  if (count == 0) return GB_OK;
  CombineMultiArgs m{};
  CombineArgs& a = m.base;
  a.y = y; a.out = nullptr; a.nk = nk; a.count = count;
  m.n_out = n_out;
  bool aligned = reinterpret_cast<uintptr_t>(y) % 16 == 0;
  for (int o = 0; o < n_out; ++o) {
    GB_REQUIRE(outs_host[o] != nullptr, "gb_stage_combine_multi: null destination %d", o);
    m.outs[o] = outs_host[o];
    aligned = aligned && reinterpret_cast<uintptr_t>(outs_host[o]) % 16 == 0;
  }
  for (int j = 0; j < nk; ++j) {
    GB_REQUIRE(k_host[j] != nullptr, "gb_stage_combine_multi: null k[%d]", j);
    a.k[j] = k_host[j];
    a.c[j] = c_host[j];
    aligned = aligned && reinterpret_cast<uintptr_t>(k_host[j]) % 16 == 0;
  }
  a.count4 = aligned ? count / 4 : 0;
  const int64_t work = a.count4 > 0 ? a.count4 : count;
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (a.count4 > 0 && blocks > cap) blocks = cap;
  if (a.count4 == 0) blocks = (count + 255) / 256;
  LaunchScope ls(K_STAGE_AXPY, stream);
  switch (nk) {
    case 0: stage_combine_multi_kernel<0><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 1: stage_combine_multi_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 2: stage_combine_multi_kernel<2><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 3: stage_combine_multi_kernel<3><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 4: stage_combine_multi_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 5: stage_combine_multi_kernel<5><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    case 6: stage_combine_multi_kernel<6><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
    default: stage_combine_multi_kernel<7><<<(unsigned)blocks, 256, 0, stream>>>(m); break;
  }
  return check_launch("stage_combine_multi");
}

// ---- gb_rows_push: rows of a local matrix, optionally picked by an index list, stored into up to 8 destination
// matrices in consecutive segments - the sending half of a sparse halo exchange as ONE kernel over NVLink peer
// memory (no packing buffer, no collective): segment s = rows [seg_start[s], seg_start[s + 1]) of the request list
// goes to dst[s], row after row.  Posted 128-bit stores: the kernel runs at the speed of the local gather while the
// writes drain over the links.
namespace gb {
struct RowsPushArgs {
  const float* src;
  int64_t ld;
  const int64_t* idx;        // [total] local row of every request, or null: request r is row r
  int64_t seg_start[9];
  float* dst[8];
  int64_t dst_ld;
  int32_t nseg, dv;          // dv = row width in 128-bit chunks
};

__global__ void __launch_bounds__(256) rows_push_kernel(const RowsPushArgs a) {
  const int64_t total = a.seg_start[a.nseg] * a.dv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / a.dv;
    const int c = (int)(i - r * a.dv);
    int s = 0;
#pragma unroll
    for (int j = 1; j < 8; ++j) s += (j < a.nseg && r >= a.seg_start[j]) ? 1 : 0;
    const int64_t row = a.idx ? __ldg(a.idx + r) : r;
    const float4 v = ld_stream4(a.src + row * a.ld + 4 * c);
    *reinterpret_cast<float4*>(a.dst[s] + (r - a.seg_start[s]) * a.dst_ld + 4 * c) = v;
  }
}
}  // namespace gb

extern "C" int gb_rows_push(const float* src, int32_t ld, const int64_t* idx, int32_t nseg, const int64_t* seg_start_host,
                            float* const* dst_host, int32_t dst_ld, int32_t d, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(src && seg_start_host && dst_host && nseg >= 1 && nseg <= 8 && d > 0 && d % 4 == 0 && ld >= d && dst_ld >= d,
             "gb_rows_push: bad arguments (1 <= nseg <= 8, d % 4 == 0)");
  GB_REQUIRE(ld % 4 == 0 && dst_ld % 4 == 0 && reinterpret_cast<uintptr_t>(src) % 16 == 0,
             "gb_rows_push: rows must be 16-byte aligned");
  RowsPushArgs a{};
  a.src = src; a.ld = ld; a.idx = idx; a.nseg = nseg; a.dst_ld = dst_ld; a.dv = d / 4;
  for (int s = 0; s <= nseg; ++s) {
    a.seg_start[s] = seg_start_host[s];
    GB_REQUIRE(s == 0 ? seg_start_host[0] == 0 : seg_start_host[s] >= seg_start_host[s - 1], "gb_rows_push: segment starts must ascend from 0");
  }
  for (int s = 0; s < nseg; ++s) {
    const bool empty = seg_start_host[s + 1] == seg_start_host[s];
    GB_REQUIRE(empty || (dst_host[s] != nullptr && reinterpret_cast<uintptr_t>(dst_host[s]) % 16 == 0),
               "gb_rows_push: null / unaligned destination %d", s);
    a.dst[s] = dst_host[s];
  }
  const int64_t work = a.seg_start[nseg] * a.dv;
  if (work == 0) return GB_OK;
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  LaunchScope ls(K_STAGE_AXPY, stream);
  rows_push_kernel<<<(unsigned)blocks, 256, 0, stream>>>(a);
  return check_launch("rows_push");
}
