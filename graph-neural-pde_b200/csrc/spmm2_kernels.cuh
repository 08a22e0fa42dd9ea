// spmm2_kernels.cuh - dynamically scheduled CSR / CSC SpMM with the fused diffusion epilogue (LaplacianODEFunc.forward,
// src/function_laplacian_diffusion.py:39-77, and its VJP) for LARGE graphs; same arguments and results as the
// statically tiled spmm_rows / spmm_long pair of spmm.cu, which stays in charge of small graphs.
//
// Why: on skewed graphs the static version spends more time in its hub-row kernel (one CTA per row > 256 edges, a
// second launch) than in the main kernel (profiles/r01q, r02p: R-MAT scale 24, 14 ms against 7.5 ms), and raising
// the threshold makes a 16-lane group walk thousands of edges while the rest of its CTA waits.  Here
//   * ONE persistent launch: CTAs of W warps, one per SM; warps pull chunks of 16 consecutive rows from an atomic
//     counter (integer atomic: the schedule never changes a sum), a warp owns a row of up to 2048 edges, longer rows
//     are worked CTA-wide first (warps stride the batches, partial sums merged in warp order)
//   * the gathered x[col] rows are STAGED IN SHARED MEMORY BY THE TMA UNIT: one `cp.async.bulk` per edge into the
//     warp's ring of 2 stages, completion on an mbarrier, issued a batch ahead of the arithmetic and across row
//     boundaries inside a chunk; the edge values (backward: through the CSC -> CSR permutation) are fetched at issue
//     time too and ride in a register
//   * lanes span the columns (128-bit chunks; d <= 64: two edge groups per warp): one LDS.128 + one shuffle + two FFMA2
//     per edge and lane
//   * backward: the per-row terms of dalpha / dbeta are summed per CHUNK (fixed rows, fixed order -> one slot per
//     chunk, schedule independent) and reduced by the existing fixed-order two-level reduction
// Roofline: HBM; algorithmic bytes as spmm.cu (8E + 12Nd + 4N forward, 12E + 16Nd + 4N backward); gather model
// + 4 (E - N) d.
#pragma once
#include "nl2_kernels.cuh"

namespace gb {

constexpr int kSpmm2Stages = 2;
constexpr int kSpmm2ChunkRows = 16;      // rows per scheduler grab = rows per partial-sum slot (backward)
constexpr int kSpmm2LongRow = 2048;      // rows longer than this are worked by a whole CTA

struct Spmm2Args {
  SpmmArgs sp;
  int32_t* counter;        // scheduler counter, zeroed before the launch
  int32_t n_long;          // entries of sp.long_list (rows > 256 edges; only those > kSpmm2LongRow are CTA-wide)
  int32_t nchunks;         // ceil((n - row0) / kSpmm2ChunkRows): backward slot of chunk c = partial_base + c,
                           // of list entry li = partial_base + nchunks + li
};

template <int XL, int W>
__host__ __device__ constexpr int spmm2_smem_bytes() {
  // per warp: ring [stages][EB rows of 16 XL bytes] + mbarriers, padded to 128; CTA: + [32] float4 merge slots
  constexpr int EB = 8 * (32 / XL);
  constexpr int warp_bytes = ((kSpmm2Stages * EB * 16 * XL + 8 * kSpmm2Stages + 127) / 128) * 128;
  return W * warp_bytes + 32 * 16;
}

template <int XL, bool BWD>
struct Spmm2Warp {
  static constexpr int EB = 8 * (32 / XL);                  // edges per batch
  static constexpr int EPI = 32 / XL;                       // edge groups
  static constexpr int NS = kSpmm2Stages;
  static constexpr int RS = 16 * XL;                        // bytes between staged rows
  static constexpr int WARP_BYTES = ((NS * EB * RS + 8 * NS + 127) / 128) * 128;

  int xl, es;
  float4 acc;
  unsigned char* ring;
  uint32_t ring_a, bar_a;
  int cstage, cphase, pstage;

  __device__ __forceinline__ void init(unsigned char* base, int lane) {
    xl = lane % XL;
    es = lane / XL;
    ring = base;
    ring_a = smem_addr(base);
    bar_a = smem_addr(base + NS * EB * RS);
    cstage = cphase = pstage = 0;
    if (lane == 0)
      for (int s = 0; s < NS; ++s) mbar_init(bar_a + 8 * s, 1);
  }

  // issue the gather of `cnt` rows; lane e < cnt holds the column id of edge e
  __device__ __forceinline__ void issue(const SpmmArgs& a, int cols, int cnt, int lane) {
    const uint32_t bar = bar_a + 8 * pstage;
    const uint32_t row_bytes = 16u * a.dv;
    if (lane == 0) mbar_expect_tx(bar, row_bytes * cnt);
    __syncwarp();
    if (lane < cnt)
      bulk_row_g2s(ring_a + (uint32_t)((pstage * EB + lane) * RS), a.src + (size_t)cols * a.lds, row_bytes, bar);
    pstage = pstage + 1 == NS ? 0 : pstage + 1;
  }

  // column id and value of edge slot p (producer side; the value goes through the CSC -> CSR permutation backward)
  static __device__ __forceinline__ void load_edge(const SpmmArgs& a, int p, int& col, float& val) {
    col = ld_stream_i32(a.idx + p);
    if (BWD) val = __ldg(a.vals + ld_stream_i32(a.perm + p));
    else val = ld_stream_f32(a.vals + p);
  }

  // consume one batch of `cnt` edges; lane e < cnt holds the value of edge e in `vals`
  __device__ __forceinline__ void batch(const SpmmArgs& a, int cnt, float vals, int lane) {
    mbar_wait(bar_a + 8 * cstage, (uint32_t)cphase);
    const unsigned char* xcol = ring + cstage * (EB * RS) + 16 * (xl < a.dv ? xl : 0);
    if (++cstage == NS) { cstage = 0; cphase ^= 1; }
    if (cnt == EB) {
#pragma unroll
      for (int i = 0; i < EB / EPI; ++i) {
        const int e = i * EPI + es;
        fma4p(acc, __shfl_sync(kFullMask, vals, e), *reinterpret_cast<const float4*>(xcol + e * RS));
      }
    } else {
#pragma unroll 1
      for (int i = 0; i * EPI < cnt; ++i) {
        const int e = i * EPI + es;
        const float w = __shfl_sync(kFullMask, vals, e < cnt ? e : 0);
        if (e < cnt) fma4p(acc, w, *reinterpret_cast<const float4*>(xcol + e * RS));
      }
    }
    __syncwarp();                                          // the stage is free again
  }

  // sum the edge groups (d <= 64): afterwards the lanes of group 0 hold the row's sums for their 4 columns
  __device__ __forceinline__ void merge_groups() {
    if constexpr (EPI > 1) {
#pragma unroll
      for (int o = XL; o < 32; o <<= 1) f4_add(acc, f4_shfl_xor(acc, o));
    }
  }
};

// a hub row, worked by the whole CTA; returns through s1 / s2 (warp 0's lanes) the row's dalpha / dbeta terms
template <int XL, bool BWD, int W>
__device__ __forceinline__ void spmm2_long_row(const SpmmArgs& a, Spmm2Warp<XL, BWD>& r, int row, float4* pacc,
                                               float alpha, float beta, float& s1, float& s2) {
  using Wp = Spmm2Warp<XL, BWD>;
  constexpr int EB = Wp::EB, NS = Wp::NS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int start = __ldg(a.ptr + row), end = __ldg(a.ptr + row + 1);
  r.acc = f4_zero();
  {
    const int stride = W * EB;
    int poff = start + warp * EB, pcols = 0;
    float pvals = 0.f, cvals = 0.f;
    auto load = [&]() {
      pcols = 0; pvals = 0.f;
      if (poff < end && lane < min(EB, end - poff)) Wp::load_edge(a, poff + lane, pcols, pvals);
    };
    load();
    // NS == 2: the values of the batch in flight wait in `cvals`
    if (poff < end) { r.issue(a, pcols, min(EB, end - poff), lane); cvals = pvals; poff += stride; load(); }
#pragma unroll 1
    for (int base = start + warp * EB; base < end; base += stride) {
      const float use = cvals;
      if (poff < end) { r.issue(a, pcols, min(EB, end - poff), lane); cvals = pvals; poff += stride; load(); }
      r.batch(a, min(EB, end - base), use, lane);
    }
  }
  r.merge_groups();
  for (int w = 0; w < W; ++w) {                             // warp 0 stores, warps 1.. add in turn (fixed order)
    if (warp == w && r.es == 0) {
      float4 t = w == 0 ? f4_zero() : pacc[r.xl];
      f4_add(t, r.acc);
      pacc[r.xl] = t;
    }
    __syncthreads();
  }
  if (warp == 0 && r.es == 0 && r.xl < a.dv)
    row_epilogue<4, BWD>(a, row, 4 * r.xl, Vec<4>{pacc[r.xl]}, alpha, beta, s1, s2);
  __syncthreads();                                          // pacc is reused by the next hub row
}

template <int XL, bool BWD, int W>
__global__ void __launch_bounds__(W * 32, 1) spmm2_kernel(const Spmm2Args g) {
  extern __shared__ __align__(128) unsigned char spmm2_smem[];
  using Wp = Spmm2Warp<XL, BWD>;
  constexpr int EB = Wp::EB;
  static_assert(kSpmm2Stages == 2, "the value hand-over below keeps exactly one batch in flight");
  const SpmmArgs& a = g.sp;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Wp r;
  r.init(spmm2_smem + (size_t)warp * Wp::WARP_BYTES, lane);
  float4* pacc = reinterpret_cast<float4*>(spmm2_smem + (size_t)W * Wp::WARP_BYTES);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const float alpha = load_alpha(a);
  const float beta = (!BWD && a.beta_train != nullptr) ? __ldg(a.beta_train) : 0.f;
  const int nrows = a.n - a.row0;
  // hub rows first, one CTA each; every list entry gets its partial-sum slot written (zeros when a warp takes the row)
  for (int li = blockIdx.x; li < g.n_long; li += gridDim.x) {
    const int row = __ldg(a.long_list + li);
    float s1 = 0.f, s2 = 0.f;
    const bool mine = row >= a.row0 && row < a.n && __ldg(a.ptr + row + 1) - __ldg(a.ptr + row) > kSpmm2LongRow;
    if (mine) spmm2_long_row<XL, BWD, W>(a, r, row, pacc, alpha, beta, s1, s2);
    if (BWD && warp == 0) {
      s1 = group_sum<32>(s1);
      s2 = group_sum<32>(s2);
      if (lane == 0) {
        a.partials[2 * (size_t)(a.partial_base + g.nchunks + li)] = s1;
        a.partials[2 * (size_t)(a.partial_base + g.nchunks + li) + 1] = s2;
      }
    }
  }
  for (;;) {
    int r0 = 0;
    if (lane == 0) r0 = atomicAdd(g.counter, kSpmm2ChunkRows);
    r0 = __shfl_sync(kFullMask, r0, 0);
    if (r0 >= nrows) break;
    const int nch = min(kSpmm2ChunkRows, nrows - r0);
    const int rp = lane <= nch ? __ldg(a.ptr + a.row0 + r0 + lane) : 0;
    // producer cursor: one batch ahead of the arithmetic, across the rows of the chunk
    int pr = -1, poff = 0, pend = 0, pcols = 0;
    float pvals = 0.f, cvals = 0.f;
    auto next_row = [&]() {
      while (poff >= pend) {
        if (++pr >= nch) break;
        poff = __shfl_sync(kFullMask, rp, pr);
        pend = __shfl_sync(kFullMask, rp, pr + 1);
        if (pend - poff > kSpmm2LongRow) poff = pend;       // done CTA-wide above
      }
      pcols = 0; pvals = 0.f;
      if (pr < nch && lane < min(EB, pend - poff)) Wp::load_edge(a, poff + lane, pcols, pvals);
    };
    auto produce = [&]() {                                   // issue the next batch; its values wait in cvals
      if (pr < nch) {
        r.issue(a, pcols, min(EB, pend - poff), lane);
        cvals = pvals;
        poff += EB;
        next_row();
      }
    };
    next_row();
    produce();
    float s1 = 0.f, s2 = 0.f;
#pragma unroll 1
    for (int rr = 0; rr < nch; ++rr) {
      const int row = a.row0 + r0 + rr;
      const int start = __shfl_sync(kFullMask, rp, rr), end = __shfl_sync(kFullMask, rp, rr + 1);
      if (end - start > kSpmm2LongRow) continue;
      r.acc = f4_zero();
#pragma unroll 1
      for (int base = start; base < end; base += EB) {
        const float use = cvals;
        produce();
        r.batch(a, min(EB, end - base), use, lane);
      }
      r.merge_groups();
      if (r.es == 0 && r.xl < a.dv) row_epilogue<4, BWD>(a, row, 4 * r.xl, Vec<4>{r.acc}, alpha, beta, s1, s2);
    }
    if (BWD) {                                               // one slot per chunk: fixed rows, fixed order
      s1 = group_sum<32>(s1);
      s2 = group_sum<32>(s2);
      if (lane == 0) {
        const size_t slot = (size_t)a.partial_base + (size_t)(r0 / kSpmm2ChunkRows);
        a.partials[2 * slot] = s1;
        a.partials[2 * slot + 1] = s2;
      }
    }
  }
}

// slots of one column slice (backward partial sums)
inline int spmm2_slots(int nrows, int n_long) { return (nrows + kSpmm2ChunkRows - 1) / kSpmm2ChunkRows + n_long; }

template <int XL, bool BWD, int W>
int spmm2_launch(const SpmmArgs& a, int n_long, int32_t* counter, cudaStream_t stream, int* slots_used) {
  static bool configured = false;
  constexpr int smem = spmm2_smem_bytes<XL, W>();
  static_assert(smem <= 227 * 1024, "spmm2: ring too large");
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(spmm2_kernel<XL, BWD, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  const int nrows = a.n - a.row0;
  Spmm2Args g{};
  g.sp = a;
  g.counter = counter;
  g.n_long = n_long;
  g.nchunks = (nrows + kSpmm2ChunkRows - 1) / kSpmm2ChunkRows;
  if (slots_used) *slots_used = spmm2_slots(nrows, n_long);
  if (nrows <= 0) return GB_OK;
  GB_CUDA(cudaMemsetAsync(counter, 0, sizeof(int32_t), stream));
  const int want = (g.nchunks + W - 1) / W;
  const int blocks = min_int(want, persistent_sms());
  {
    LaunchScope ls(BWD ? K_SPMM_BWD : K_SPMM_FWD, stream);
    spmm2_kernel<XL, BWD, W><<<blocks, W * 32, smem, stream>>>(g);
  }
  return check_launch("spmm2_kernel");
}

}  // namespace gb
