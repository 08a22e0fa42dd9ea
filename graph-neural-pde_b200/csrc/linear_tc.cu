// linear_tc.cu - the Q/K projection GEMMs (nn.Linear forward / input gradient,
// src/function_transformer_attention.py:224-225) on the 5th-generation tensor cores: tcgen05.mma kind::tf32
// with the accumulator in TMEM, fp32-accurate through a 3-term split (3xTF32):
//
//     a = a_hi + a_lo (+ <= 2^-22 |a|),  a_hi = rna_tf32(a), a_lo = rna_tf32(a - a_hi)   (both exact tf32 values,
//                                        so the result does not depend on how the tensor core treats low bits)
//     a.b ~= a_hi.b_hi + a_hi.b_lo + a_lo.b_hi        (dropped: a_lo.b_lo <= 2^-22 |a||b|), fp32 accumulate
//
// out[rows, NO] (+)= A[rows, K] . B[NO, K]^T (+ bias),  rows ~ 10^6 (nodes), NO, K <= 128: a tall-skinny GEMM
// that is HBM-bound once the math is on the tensor pipe (C4: 1.25 GB in + 0.63 GB out per projection; the
// CUDA-core SGEMM it replaces ran at 0.75 TB/s).
//
// One persistent CTA per SM, warp-specialised (no TMA: every operand element has to pass through registers
// anyway to be split, so the producers write the canonical UMMA shared-memory layout themselves):
//   warps 0-7   producers: two groups of 4 warps alternate over 32-float K-blocks of the current 128-row tile:
//               coalesced 128-bit global loads -> hi / lo split -> st.shared into the K-major SWIZZLE_128B
//               layout (16-byte chunk index XOR (row & 7)) -> fence.proxy.async -> mbarrier arrive
//   warp 8      MMA issuer: one elected thread waits for a stage, issues 3 x (32/8) tcgen05.mma per K-block
//               into one of two TMEM accumulators, tcgen05.commit frees the stage / publishes the tile
//   warps 9-12  epilogue: tcgen05.ld (32 lanes x 16 columns per instruction), + bias (+ previous value), 128-bit
//               global stores; then hands the accumulator back
// B (the weights, <= 64 KB) is split once per CTA and stays resident in shared memory.
#include "common.cuh"

namespace gb {

constexpr int kTcProducerWarps = 8;
constexpr int kTcThreads = (kTcProducerWarps + 1 + 4) * 32;     // 416
constexpr int kTcTileM = 128;
constexpr int kTcKBlock = 32;                                   // floats per K-block = one 128-byte swizzle row
constexpr int kTcMaxStages = 4;

struct TcArgs {
  const float* a;        // [rows, K]
  int64_t lda;
  const float* w;        // trans_w == 0: [NO, K] (K contiguous); trans_w == 1: [K, NO] (NO contiguous)
  int64_t ldw;
  const float* bias;     // [NO] or null
  float* out;            // [rows, NO]
  int64_t ldo;
  int32_t rows, no, k;   // NO % 16 == 0, NO <= 128, K % 32 == 0, K <= 128
  int32_t trans_w, accumulate, stages;
  // K slices of a wider reduction (gb_linear_fwd on the encoder, src/GNN.py:30-38: F = 1433 input features):
  // the slice holds k_valid <= k real columns (the rest of the last K-block is zero-filled), and operands whose
  // rows are not 16-byte aligned are loaded element-wise.  k_valid == 0: k_valid = k, both operands vectorised.
  int32_t k_valid, a_vec, w_vec;
};

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, tf32 inputs, fp32 accumulate
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// shared-memory matrix descriptor, K-major, SWIZZLE_128B: rows of 128 bytes, 8-row groups 1024 bytes apart
// (cute::UMMA::SmemDescriptor: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46 | layout 2 << 61)
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = TF32, both K-major, M = 128
__device__ __forceinline__ uint32_t umma_idesc_tf32(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTcTileM >> 4) << 24);
}

__device__ __forceinline__ float rna_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));      // round to nearest: 13 low mantissa bits become zero
  return __uint_as_float(r);
}
__device__ __forceinline__ void split_tf32(const float4& v, float4& hi, float4& lo) {
  hi.x = rna_tf32(v.x); hi.y = rna_tf32(v.y); hi.z = rna_tf32(v.z); hi.w = rna_tf32(v.w);
  lo.x = rna_tf32(v.x - hi.x);                              // v - hi is exact in fp32; |lo| <= 2^-11 |v|
  lo.y = rna_tf32(v.y - hi.y);
  lo.z = rna_tf32(v.z - hi.z);
  lo.w = rna_tf32(v.w - hi.w);
}

// four consecutive K elements starting at column kcol of a row; columns >= kv read as zero
__device__ __forceinline__ float4 tc_load4(const float* p, int kcol, int kv, bool vec, bool stream) {
  if (vec && kcol + 4 <= kv) return stream ? ld_stream4(p) : __ldg(reinterpret_cast<const float4*>(p));
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (kcol < kv) v.x = __ldg(p);
  if (kcol + 1 < kv) v.y = __ldg(p + 1);
  if (kcol + 2 < kv) v.z = __ldg(p + 2);
  if (kcol + 3 < kv) v.w = __ldg(p + 3);
  return v;
}

// byte offset of (row, 16-byte chunk c) inside a [rows x 128 B] K-block tile in the SWIZZLE_128B layout
__device__ __forceinline__ uint32_t sw128(int row, int c) { return (uint32_t)row * 128u + (uint32_t)((c ^ (row & 7)) << 4); }

struct TcSmem {
  uint32_t b_hi, b_lo;       // [KB][NO x 128 B]
  uint32_t a_hi, a_lo;       // [stages][128 x 128 B]
  uint32_t ready, empty;     // [stages] mbarriers
  uint32_t tfull, tempty;    // [2] mbarriers
  uint32_t tmem_ptr;
  uint32_t bias;             // [NO] floats
  uint32_t total;
};

__host__ __device__ inline TcSmem tc_smem_layout(int no, int k, int stages) {
  TcSmem s{};
  const uint32_t kb = (uint32_t)k / kTcKBlock;
  uint32_t off = 0;
  s.b_hi = off; off += kb * (uint32_t)no * 128u;
  s.b_lo = off; off += kb * (uint32_t)no * 128u;
  s.a_hi = off; off += (uint32_t)stages * kTcTileM * 128u;
  s.a_lo = off; off += (uint32_t)stages * kTcTileM * 128u;
  s.ready = off; off += 8u * kTcMaxStages;
  s.empty = off; off += 8u * kTcMaxStages;
  s.tfull = off; off += 16u;
  s.tempty = off; off += 16u;
  s.tmem_ptr = off; off += 16u;
  s.bias = off; off += 4u * 128u;
  s.total = off;
  return s;
}

__global__ void __launch_bounds__(kTcThreads, 1) tc_gemm_rows_kernel(const TcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic shared memory is only guaranteed 16-byte aligned: align the tile area to 1024 bytes by hand
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);                  // generic pointer to the aligned base
  const TcSmem L = tc_smem_layout(a.no, a.k, a.stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KB = a.k / kTcKBlock;
  const int ntiles = (a.rows + kTcTileM - 1) / kTcTileM;
  const uint32_t tmem_cols = a.no * 2 <= 32 ? 32u : (a.no * 2 <= 64 ? 64u : (a.no * 2 <= 128 ? 128u : 256u));

  // ---- one-time setup ------------------------------------------------------------------------------
  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(base + L.ready + 8 * s, 4 * 32);           // one producer group
      mbar_init(base + L.empty + 8 * s, 1);                // tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(base + L.tfull + 8 * i, 1);                // tcgen05.commit
      mbar_init(base + L.tempty + 8 * i, 4 * 32);          // epilogue threads
    }
    fence_barrier_init();
  }
  if (warp == kTcProducerWarps) tmem_alloc(base + L.tmem_ptr, tmem_cols);
  // weights: split once, resident for the whole kernel
  {
    const int chunks = a.no * (a.k / 4);                   // 16-byte chunks of B[NO][K]
    for (int idx = threadIdx.x; idx < chunks; idx += kTcThreads) {
      int n, kc;
      float4 v;
      if (!a.trans_w) {
        kc = idx % (a.k / 4);
        n = idx / (a.k / 4);
        v = tc_load4(a.w + (int64_t)n * a.ldw + 4 * kc, 4 * kc, a.k_valid, a.w_vec != 0, false);
      } else {
        n = idx % a.no;
        kc = idx / a.no;
        const float* p = a.w + (int64_t)(4 * kc) * a.ldw + n;
        v = make_float4(__ldg(p), __ldg(p + a.ldw), __ldg(p + 2 * a.ldw), __ldg(p + 3 * a.ldw));
      }
      float4 hi, lo;
      split_tf32(v, hi, lo);
      const int kb = (4 * kc) / kTcKBlock, c = ((4 * kc) % kTcKBlock) / 4;
      const uint32_t off = (uint32_t)kb * (uint32_t)a.no * 128u + sw128(n, c);
      *reinterpret_cast<float4*>(gen + L.b_hi + off) = hi;
      *reinterpret_cast<float4*>(gen + L.b_lo + off) = lo;
    }
    for (int j = threadIdx.x; j < a.no; j += kTcThreads)
      reinterpret_cast<float*>(gen + L.bias)[j] = a.bias ? __ldg(a.bias + j) : 0.f;
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen + L.tmem_ptr);

  if (warp < kTcProducerWarps) {
    // ===== producers =================================================================================
    const int grp = warp >> 2;                             // K-block iterations with (it & 1) == grp
    const int t = threadIdx.x & 127;
    const int c = t & 7, r0 = t >> 3;                      // 16-byte chunk, first row (then +16 per step)
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int kb = 0; kb < KB; ++kb, ++it) {
        if ((it & 1) != grp) continue;
        const int s = it % a.stages;
        const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int row = tile * kTcTileM + r0 + 16 * i;
          const int kcol = kb * kTcKBlock + 4 * c;
          v[i] = row < a.rows ? tc_load4(a.a + (int64_t)row * a.lda + kcol, kcol, a.k_valid, a.a_vec != 0, true)
                              : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        mbar_wait(base + L.empty + 8 * s, ph ^ 1u);
        uint8_t* hi_t = gen + L.a_hi + (uint32_t)s * kTcTileM * 128u;
        uint8_t* lo_t = gen + L.a_lo + (uint32_t)s * kTcTileM * 128u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float4 hi, lo;
          split_tf32(v[i], hi, lo);
          const uint32_t off = sw128(r0 + 16 * i, c);
          *reinterpret_cast<float4*>(hi_t + off) = hi;
          *reinterpret_cast<float4*>(lo_t + off) = lo;
        }
        fence_proxy_async();
        mbar_arrive(base + L.ready + 8 * s);
      }
    }
  } else if (warp == kTcProducerWarps) {
    // ===== MMA issuer (one thread) ====================================================================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_tf32(a.no);
      int it = 0, ti = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
        const int acc = ti & 1;
        const uint32_t aph = (uint32_t)(ti >> 1) & 1u;
        mbar_wait(base + L.tempty + 8 * acc, aph ^ 1u);    // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * a.no);
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % a.stages;
          const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
          mbar_wait(base + L.ready + 8 * s, ph);
          tc_fence_after();
          const uint32_t ahi = base + L.a_hi + (uint32_t)s * kTcTileM * 128u;
          const uint32_t alo = base + L.a_lo + (uint32_t)s * kTcTileM * 128u;
          const uint32_t bhi = base + L.b_hi + (uint32_t)kb * (uint32_t)a.no * 128u;
          const uint32_t blo = base + L.b_lo + (uint32_t)kb * (uint32_t)a.no * 128u;
#pragma unroll
          for (int k8 = 0; k8 < kTcKBlock / 8; ++k8) {     // one tcgen05.mma covers K = 8 tf32 = 32 bytes
            const uint32_t ko = 32u * k8;
            umma_tf32(d_tmem, umma_desc_k128(ahi + ko), umma_desc_k128(bhi + ko), idesc, (kb | k8) != 0);
            umma_tf32(d_tmem, umma_desc_k128(ahi + ko), umma_desc_k128(blo + ko), idesc, 1u);
            umma_tf32(d_tmem, umma_desc_k128(alo + ko), umma_desc_k128(bhi + ko), idesc, 1u);
          }
          umma_commit(base + L.empty + 8 * s);             // stage free once these MMAs have read it
        }
        umma_commit(base + L.tfull + 8 * acc);             // accumulator complete
      }
    }
  } else {
    // ===== epilogue ===================================================================================
    const int q = warp & 3;                                // TMEM lane quadrant this warp may read
    const float* bias_s = reinterpret_cast<const float*>(gen + L.bias);
    int ti = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
      const int acc = ti & 1;
      const uint32_t aph = (uint32_t)(ti >> 1) & 1u;
      mbar_wait(base + L.tfull + 8 * acc, aph);
      tc_fence_after();
      const int row = tile * kTcTileM + 32 * q + lane;
      float* orow = a.out + (int64_t)row * a.ldo;
      for (int c0 = 0; c0 < a.no; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(acc * a.no + c0), v);
        if (row < a.rows) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            float4 o = make_float4(v[j] + bias_s[c0 + j], v[j + 1] + bias_s[c0 + j + 1], v[j + 2] + bias_s[c0 + j + 2],
                                   v[j + 3] + bias_s[c0 + j + 3]);
            if (a.accumulate) {
              const float4 old = *reinterpret_cast<const float4*>(orow + c0 + j);
              o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
            }
            *reinterpret_cast<float4*>(orow + c0 + j) = o;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(base + L.tempty + 8 * acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTcProducerWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ==================================================================================================
// dW[M=128, d] = dY^T X, db[M] = column sums of dY  (weight gradient of the packed Q|K projection):
// the reduction runs over the NODES, so both operands are "MN-major" in memory (the M / N index is the
// contiguous one).  For 32-bit MN-major operands the only UMMA layout is SWIZZLE_128B_BASE32B (cute::UMMA
// Layout_MN_SW128_32B_Atom, Swizzle<2,5,2>; canonical form ((8,n),(4,k)):((1,LBO),(8,SBO)) in 16-byte units):
// atoms of 4 K-rows x 128 bytes (32 consecutive features of 4 consecutive nodes) whose 32-byte chunks are
// XOR-ed with (K-row & 3) - a row-major block of the source matrix, re-chunked - LBO = byte stride between
// the atoms along M/N, SBO = byte stride between the 4-row K-groups.  One tcgen05.mma (K = 8) spans two
// K-groups; a K-block is 32 nodes = 8 K-groups.
// The bias gradient rides along as 16 extra B columns holding 1.0 (N = d + 16).
// Every CTA accumulates ALL its K-blocks (static interleave) in one TMEM accumulator and writes one partial
// [128, d + 16]; a second kernel adds the partials in CTA order (deterministic).
struct TcTnArgs {
  const float* dy;       // [rows, 128]
  int64_t lddy;
  const float* x;        // [rows, d]
  int64_t ldx;
  float* part;           // [gridDim.x][128][d + 16]
  int32_t rows, d, stages;
};

struct TcTnSmem {
  uint32_t a_hi, a_lo, b_hi, b_lo;   // per stage: A 16 KB, B (d/32 + 1 atoms) x 4 KB
  uint32_t ready, empty, tfull, tmem_ptr, total;
  uint32_t a_stage, b_stage;         // bytes per stage
};

__host__ __device__ inline TcTnSmem tc_tn_layout(int d, int stages) {
  TcTnSmem s{};
  s.a_stage = 4u * 4096u;                                   // 4 M-atoms x (4 K-groups x 1024 B)
  s.b_stage = (uint32_t)(d / 32 + 1) * 4096u;               // N-atoms of x + the ones atom
  uint32_t off = 0;
  s.a_hi = off; off += stages * s.a_stage;
  s.a_lo = off; off += stages * s.a_stage;
  s.b_hi = off; off += stages * s.b_stage;
  s.b_lo = off; off += stages * s.b_stage;
  s.ready = off; off += 8u * kTcMaxStages;
  s.empty = off; off += 8u * kTcMaxStages;
  s.tfull = off; off += 16u;
  s.tmem_ptr = off; off += 16u;
  s.total = off;
  return s;
}

// MN-major SWIZZLE_128B_BASE32B descriptor: LBO = 4096 B between M/N atoms, SBO = 512 B between 4-row K-groups
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (256ull << 16) | (32ull << 32) | (1ull << 46) | (1ull << 61);
}
// byte offset of (node 0..31, 16-byte chunk c 0..7) inside one M/N atom column [8 K-groups][4 rows][128 B]
__device__ __forceinline__ uint32_t sw128_32b(int node, int c) {
  const int c32 = ((c >> 1) ^ (node & 3));                  // 32-byte chunk XOR K-row within the atom
  return (uint32_t)(node >> 2) * 512u + (uint32_t)(node & 3) * 128u + (uint32_t)(c32 << 5) + (uint32_t)((c & 1) << 4);
}

__global__ void __launch_bounds__(kTcThreads, 1) tc_gemm_tn_kernel(const TcTnArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - raw);
  const TcTnSmem L = tc_tn_layout(a.d, a.stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_mma = a.d + 16;                               // x columns + ones columns
  const int nkb = (a.rows + kTcKBlock - 1) / kTcKBlock;     // K-blocks of 32 nodes
  const int xatoms = a.d / 32;
  const uint32_t tmem_cols = n_mma <= 32 ? 32u : (n_mma <= 64 ? 64u : (n_mma <= 128 ? 128u : 256u));

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(base + L.ready + 8 * s, 4 * 32);
      mbar_init(base + L.empty + 8 * s, 1);
    }
    mbar_init(base + L.tfull, 1);
    fence_barrier_init();
  }
  if (warp == kTcProducerWarps) tmem_alloc(base + L.tmem_ptr, tmem_cols);
  // the ones atom of B (hi = 1 in the first 16 columns, lo = 0) is constant: fill it once in every stage
  for (int idx = threadIdx.x; idx < a.stages * 256; idx += kTcThreads) {      // 256 chunks of 16 B per atom
    const int s = idx / 256, r = idx % 256;
    const int node = r / 8, c = r % 8;
    const float one = c < 4 ? 1.f : 0.f;                                      // columns d .. d+15
    const uint32_t off = (uint32_t)s * L.b_stage + (uint32_t)xatoms * 4096u + sw128_32b(node, c);
    *reinterpret_cast<float4*>(gen + L.b_hi + off) = make_float4(one, one, one, one);
    *reinterpret_cast<float4*>(gen + L.b_lo + off) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(gen + L.tmem_ptr);

  if (warp < kTcProducerWarps) {
    const int grp = warp >> 2;
    const int t = threadIdx.x & 127;
    const int c = t & 7;                                    // 16-byte chunk inside a 128-byte atom row
    int it = 0;
    for (int kb = blockIdx.x; kb < nkb; kb += gridDim.x, ++it) {
      if ((it & 1) != grp) continue;
      const int s = it % a.stages;
      const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
      // (atom, node) pairs: t / 8 runs over 16 of them per step; A has 4 atoms x 32 nodes = 128 pairs (8 steps)
      float4 va[8], vb[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int pr = (t >> 3) + 16 * i;                   // 0 .. 127
        const int atom = pr >> 5, node = pr & 31;
        const int row = kb * kTcKBlock + node;
        const bool ok = row < a.rows;
        va[i] = ok ? ld_stream4(a.dy + (int64_t)row * a.lddy + 32 * atom + 4 * c) : make_float4(0.f, 0.f, 0.f, 0.f);
        vb[i] = (ok && atom < xatoms) ? ld_stream4(a.x + (int64_t)row * a.ldx + 32 * atom + 4 * c)
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      mbar_wait(base + L.empty + 8 * s, ph ^ 1u);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int pr = (t >> 3) + 16 * i;
        const int atom = pr >> 5, node = pr & 31;
        const uint32_t off = (uint32_t)atom * 4096u + sw128_32b(node, c);
        float4 hi, lo;
        split_tf32(va[i], hi, lo);
        *reinterpret_cast<float4*>(gen + L.a_hi + (uint32_t)s * L.a_stage + off) = hi;
        *reinterpret_cast<float4*>(gen + L.a_lo + (uint32_t)s * L.a_stage + off) = lo;
        if (atom < xatoms) {
          split_tf32(vb[i], hi, lo);
          *reinterpret_cast<float4*>(gen + L.b_hi + (uint32_t)s * L.b_stage + off) = hi;
          *reinterpret_cast<float4*>(gen + L.b_lo + (uint32_t)s * L.b_stage + off) = lo;
        }
      }
      fence_proxy_async();
      mbar_arrive(base + L.ready + 8 * s);
    }
  } else if (warp == kTcProducerWarps) {
    if (lane == 0) {
      // D = F32, A = B = TF32, both MN-major (bits 15, 16), M = 128, N = d + 16
      const uint32_t idesc = umma_idesc_tf32(n_mma) | (1u << 15) | (1u << 16);
      int it = 0;
      for (int kb = blockIdx.x; kb < nkb; kb += gridDim.x, ++it) {
        const int s = it % a.stages;
        const uint32_t ph = (uint32_t)(it / a.stages) & 1u;
        mbar_wait(base + L.ready + 8 * s, ph);
        tc_fence_after();
        const uint32_t ahi = base + L.a_hi + (uint32_t)s * L.a_stage, alo = base + L.a_lo + (uint32_t)s * L.a_stage;
        const uint32_t bhi = base + L.b_hi + (uint32_t)s * L.b_stage, blo = base + L.b_lo + (uint32_t)s * L.b_stage;
#pragma unroll
        for (int kg = 0; kg < 4; ++kg) {                    // one tcgen05.mma = 8 nodes = two 4-row K-groups
          const uint32_t ko = 1024u * kg;
          umma_tf32(tmem_base, umma_desc_mn128(ahi + ko), umma_desc_mn128(bhi + ko), idesc, (it | kg) != 0);
          umma_tf32(tmem_base, umma_desc_mn128(ahi + ko), umma_desc_mn128(blo + ko), idesc, 1u);
          umma_tf32(tmem_base, umma_desc_mn128(alo + ko), umma_desc_mn128(bhi + ko), idesc, 1u);
        }
        umma_commit(base + L.empty + 8 * s);
      }
      umma_commit(base + L.tfull);
    }
  } else {
    const int q = warp & 3;
    mbar_wait(base + L.tfull, 0u);
    tc_fence_after();
    const int m = 32 * q + lane;
    float* prow = a.part + ((int64_t)blockIdx.x * kTcTileM + m) * n_mma;
    const bool any = (int)blockIdx.x < nkb;                 // a CTA without work leaves TMEM undefined: write zeros
    for (int c0 = 0; c0 < n_mma; c0 += 16) {
      float v[16];
      tmem_ld16(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<float4*>(prow + c0 + j) = any ? make_float4(v[j], v[j + 1], v[j + 2], v[j + 3])
                                                        : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTcProducerWarps) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// dw[m, c] = sum_cta part[cta][m][c] (c < d), db[m] = sum_cta part[cta][m][d]; fixed CTA order
__global__ void __launch_bounds__(256) tc_tn_reduce_kernel(const float* __restrict__ part, int ctas, int d,
                                                         float* __restrict__ dw, float* __restrict__ db) {
  const int n_mma = d + 16;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= kTcTileM * (d + 1)) return;
  const int m = idx / (d + 1), c = idx % (d + 1);
  float acc = 0.f;
  for (int b = 0; b < ctas; ++b) acc += part[((int64_t)b * kTcTileM + m) * n_mma + c];
  if (c < d) dw[(int64_t)m * d + c] = acc;
  else if (db) db[m] = acc;
}

bool tc_gemm_tn_supported(const void* dy, int64_t lddy, const void* x, int64_t ldx, int nout, int d) {
  static const bool disabled = getenv("GRANDB200_NO_TC") != nullptr;
  if (disabled) return false;
  auto al = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  return nout == kTcTileM && d % 32 == 0 && d >= 32 && d <= 128 && al(dy) && al(x) && lddy % 4 == 0 && ldx % 4 == 0;
}

size_t tc_gemm_tn_workspace_bytes(int d) { return (size_t)sm_count() * kTcTileM * (d + 16) * sizeof(float) + 256; }

int tc_gemm_tn_launch(const float* dy, int64_t lddy, const float* x, int64_t ldx, int rows, int d, float* dw, float* db,
                      void* workspace, cudaStream_t stream) {
  TcTnArgs a{};
  a.dy = dy; a.lddy = lddy; a.x = x; a.ldx = ldx; a.rows = rows; a.d = d;
  a.part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 15) & ~(uintptr_t)15);
  a.stages = 3;
  while (a.stages > 2 && tc_tn_layout(d, a.stages).total + 1024 > 220 * 1024) --a.stages;
  const size_t smem = tc_tn_layout(d, a.stages).total + 1024;
  static bool configured = false;
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(tc_gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  const int nkb = (rows + kTcKBlock - 1) / kTcKBlock;
  const int blocks = min(nkb, sm_count());
  {
    LaunchScope ls(K_LINEAR_BWD_W, stream);
    tc_gemm_tn_kernel<<<blocks, kTcThreads, smem, stream>>>(a);
  }
  if (int rc = check_launch("tc_gemm_tn_kernel")) return rc;
  {
    LaunchScope ls(K_LINEAR_BWD_W_REDUCE, stream);
    const int total = kTcTileM * (d + 1);
    tc_tn_reduce_kernel<<<(total + 255) / 256, 256, 0, stream>>>(a.part, blocks, d, dw, db);
  }
  return check_launch("tc_tn_reduce_kernel");
}

// can the tensor-core path take out[rows, no] (+)= a[rows, k] . w + bias ?
bool tc_gemm_supported(const void* a, int64_t lda, const void* w, int64_t ldw, const void* out, int64_t ldo, int no,
                       int k, int trans_w) {
  static const bool disabled = getenv("GRANDB200_NO_TC") != nullptr;
  if (disabled) return false;
  if (no % 16 != 0 || no < 16 || no > 128 || k % kTcKBlock != 0 || k < kTcKBlock || k > 128) return false;
  auto al = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  if (!al(a) || !al(out) || lda % 4 != 0 || ldo % 4 != 0) return false;
  if (!trans_w && (!al(w) || ldw % 4 != 0)) return false;
  return true;
}

// out[rows, no] = a[rows, k] . w[no, k]^T + bias for ANY k (and unaligned a / w rows): slices of <= 128 columns
bool tc_gemm_sliced_supported(const void* out, int64_t ldo, int no) {
  static const bool disabled = getenv("GRANDB200_NO_TC") != nullptr;
  if (disabled) return false;
  return no % 16 == 0 && no >= 16 && no <= 128 && reinterpret_cast<uintptr_t>(out) % 16 == 0 && ldo % 4 == 0;
}

int tc_gemm_launch(const TcArgs& a_in, int kernel_id, cudaStream_t stream);

int tc_gemm_sliced_launch(const float* x, int64_t ldx, int rows, int k, const float* w, int64_t ldw, const float* bias,
                          int no, float* out, int64_t ldo, int kernel_id, cudaStream_t stream) {
  auto al = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  for (int k0 = 0; k0 < k; k0 += 128) {
    const int kv = min(128, k - k0);
    TcArgs t{};
    t.a = x + k0; t.lda = ldx; t.w = w + k0; t.ldw = ldw; t.bias = k0 == 0 ? bias : nullptr;
    t.out = out; t.ldo = ldo; t.rows = rows; t.no = no;
    t.k = (kv + kTcKBlock - 1) / kTcKBlock * kTcKBlock;
    t.k_valid = kv;
    t.a_vec = al(t.a) && ldx % 4 == 0;
    t.w_vec = al(t.w) && ldw % 4 == 0;
    t.accumulate = k0 != 0;
    if (int rc = tc_gemm_launch(t, kernel_id, stream)) return rc;
  }
  return GB_OK;
}

int tc_gemm_launch(const TcArgs& a_in, int kernel_id, cudaStream_t stream) {
  TcArgs a = a_in;
  if (a.k_valid == 0) { a.k_valid = a.k; a.a_vec = 1; a.w_vec = 1; }
  // shared memory: weights (2 x K x NO x 4 B) + stages x 32 KB; keep at least 2 stages
  a.stages = kTcMaxStages;
  while (a.stages > 2 && tc_smem_layout(a.no, a.k, a.stages).total + 1024 > 220 * 1024) --a.stages;
  const size_t smem = tc_smem_layout(a.no, a.k, a.stages).total + 1024;
  if (smem > 227 * 1024) {
    set_error("tc_gemm: shared memory %zu bytes exceeds the 227 KB limit", smem);
    return GB_ERR_UNSUPPORTED;
  }
  static bool configured = false;
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(tc_gemm_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  const int ntiles = (a.rows + kTcTileM - 1) / kTcTileM;
  const int blocks = min(ntiles, sm_count());
  LaunchScope ls(kernel_id, stream);
  tc_gemm_rows_kernel<<<blocks, kTcThreads, smem, stream>>>(a);
  return check_launch("tc_gemm_rows_kernel");
}

}  // namespace gb
