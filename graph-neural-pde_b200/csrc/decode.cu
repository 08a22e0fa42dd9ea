// decode.cu - the per-step evaluation of the early-stopping integrator in ONE launch
// (src/early_stop_solver.py:108-128 `evaluate` + :88-98 `ode_test`; the same three ops close GNN.forward,
// src/GNN.py:62-74): after every accepted solver step the reference runs
//
//     z = relu(y1[:, :width]);  logits = z W2^T + b2;                       (F.relu, F.linear)
//     for mask in (train, val, test):  acc = (logits[mask].argmax(-1) == y[mask]).sum() / mask.sum()
//
// i.e. a relu, a GEMM, three boolean gathers, three arg-max, three compares, six reductions and six .item()
// host synchronisations.  Here: one kernel that reads every state row once and leaves six integers on the device
// (correct / total per mask); the caller reads them back with one copy.
//
// Roofline: HBM - the state is read once (4 d bytes per node), labels 8 bytes, masks 3 bytes; the [C, d] decoder
// weights (<= 48 KB) live in shared memory.  Per node 2 C d flops on the CUDA cores (C = classes <= 64):
// lane = two nodes, 2 C accumulators in registers, the weights arrive by shared-memory broadcast (one LDS.128 per
// eight products), the state rows straight from global memory (L1 keeps the 128-byte lines a lane walks through).
// Integer atomics only (counts): the result does not depend on the order of arrival.
#include "common.cuh"

namespace gb {

constexpr int kDecWarps = 4;
constexpr int kDecThreads = kDecWarps * 32;

struct DecodeArgs {
  const float* z;            // [n, >= d] state rows; columns >= d (augmented half) are ignored
  int64_t ldz;
  const float* w2;           // [classes, d]
  int64_t ldw;
  const float* b2;           // [classes] or null
  const int64_t* y;          // [n] labels or null (then no counts)
  const uint8_t* mask[3];    // train / val / test, one byte per node (torch.bool), each may be null
  int32_t* counts;           // [6]: correct train, val, test; total train, val, test
  float* logits;             // [n, classes] or null
  int64_t ldl;
  int64_t* pred;             // [n] or null
  int32_t n, d, classes, dpad;   // dpad = d rounded up to 4
};

// first maximum, NaN counts as the largest value (torch.max / argmax semantics)
__device__ __forceinline__ bool dec_better(float v, float best) {
  return (v > best) || (v != v && best == best);
}

// F.relu keeps NaN (fmaxf would not)
__device__ __forceinline__ float dec_relu(float v) { return v < 0.f ? 0.f : v; }

// lane = kNpt nodes (rows base + lane, base + 32 + lane): CP x kNpt accumulators in registers; every weight float4 fetched
// from shared memory (one broadcast LDS.128) feeds 4 kNpt FFMAs; the state rows come straight from global memory -
// 16 bytes per lane per step, whole 128-byte lines over 8 steps (L1), one load per 4 CP kNpt FFMAs - so shared memory holds
// only the weights and the resident warps are limited by registers alone.
constexpr int kNpt = 2;

template <int CP, bool VEC>
__global__ void __launch_bounds__(kDecThreads) decode_eval_kernel(const DecodeArgs a) {
  extern __shared__ __align__(16) float dec_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d4 = a.dpad >> 2;
  float* w_s = dec_smem;                                      // [CP][dpad], zero padded
  float* b_s = w_s + CP * a.dpad;                             // [CP]
  __shared__ int cnt_s[6];
  if (threadIdx.x < 6) cnt_s[threadIdx.x] = 0;
  for (int i = threadIdx.x; i < CP * a.dpad; i += kDecThreads) {
    const int c = i / a.dpad, k = i % a.dpad;
    w_s[i] = (c < a.classes && k < a.d) ? __ldg(a.w2 + (int64_t)c * a.ldw + k) : 0.f;
  }
  for (int c = threadIdx.x; c < CP; c += kDecThreads) b_s[c] = (c < a.classes && a.b2) ? __ldg(a.b2 + c) : 0.f;
  __syncthreads();

  const int ntiles = (a.n + 32 * kNpt - 1) / (32 * kNpt);
  int correct[3] = {0, 0, 0}, total[3] = {0, 0, 0};
  for (int t = blockIdx.x * kDecWarps + warp; t < ntiles; t += gridDim.x * kDecWarps) {
    int row[kNpt];
    bool live[kNpt];
    const float* zr[kNpt];
#pragma unroll
    for (int p = 0; p < kNpt; ++p) {
      row[p] = t * 32 * kNpt + 32 * p + lane;
      live[p] = row[p] < a.n;
      zr[p] = a.z + (int64_t)(live[p] ? row[p] : 0) * a.ldz;
    }
    float acc[kNpt][CP];
#pragma unroll
    for (int p = 0; p < kNpt; ++p)
#pragma unroll
      for (int c = 0; c < CP; ++c) acc[p][c] = b_s[c];
#pragma unroll 2
    for (int k4 = 0; k4 < d4; ++k4) {
      float4 zv[kNpt];
#pragma unroll
      for (int p = 0; p < kNpt; ++p) {
        if (VEC) {
          zv[p] = __ldg(reinterpret_cast<const float4*>(zr[p]) + k4);
        } else {                                              // ragged width / unaligned rows
          const int k = 4 * k4;
          zv[p].x = k < a.d ? __ldg(zr[p] + k) : 0.f;
          zv[p].y = k + 1 < a.d ? __ldg(zr[p] + k + 1) : 0.f;
          zv[p].z = k + 2 < a.d ? __ldg(zr[p] + k + 2) : 0.f;
          zv[p].w = k + 3 < a.d ? __ldg(zr[p] + k + 3) : 0.f;
        }
        zv[p].x = dec_relu(zv[p].x); zv[p].y = dec_relu(zv[p].y); zv[p].z = dec_relu(zv[p].z); zv[p].w = dec_relu(zv[p].w);
      }
#pragma unroll
      for (int c = 0; c < CP; ++c) {
        const float4 w = *reinterpret_cast<const float4*>(w_s + c * a.dpad + 4 * k4);        // broadcast
#pragma unroll
        for (int p = 0; p < kNpt; ++p) {
          acc[p][c] = fmaf(zv[p].x, w.x, acc[p][c]);
          acc[p][c] = fmaf(zv[p].y, w.y, acc[p][c]);
          acc[p][c] = fmaf(zv[p].z, w.z, acc[p][c]);
          acc[p][c] = fmaf(zv[p].w, w.w, acc[p][c]);
        }
      }
    }
#pragma unroll
    for (int p = 0; p < kNpt; ++p) {
      int best = 0;
      float bv = acc[p][0];
#pragma unroll
      for (int c = 1; c < CP; ++c)
        if (c < a.classes && dec_better(acc[p][c], bv)) { bv = acc[p][c]; best = c; }
      if (live[p]) {
        if (a.logits) {
#pragma unroll
          for (int c = 0; c < CP; ++c)
            if (c < a.classes) a.logits[(int64_t)row[p] * a.ldl + c] = acc[p][c];
        }
        if (a.pred) a.pred[row[p]] = best;
        if (a.y) {
          const bool hit = a.y[row[p]] == (int64_t)best;
#pragma unroll
          for (int m = 0; m < 3; ++m) {
            if (a.mask[m] && a.mask[m][row[p]]) {
              ++total[m];
              correct[m] += hit ? 1 : 0;
            }
          }
        }
      }
    }
  }
  if (a.y && a.counts) {
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      const int c = __reduce_add_sync(0xffffffffu, correct[m]);
      const int tt = __reduce_add_sync(0xffffffffu, total[m]);
      if (lane == 0) {
        if (c) atomicAdd(&cnt_s[m], c);
        if (tt) atomicAdd(&cnt_s[3 + m], tt);
      }
    }
    __syncthreads();
    if (threadIdx.x < 6 && cnt_s[threadIdx.x]) atomicAdd(a.counts + threadIdx.x, cnt_s[threadIdx.x]);
  }
}

template <int CP>
static int decode_launch(const DecodeArgs& a, cudaStream_t stream) {
  const size_t smem = ((size_t)CP * a.dpad + CP) * sizeof(float);
  if (smem > 226 * 1024) {
    set_error("gb_decode_eval: %d classes x %d features need %zu bytes of shared memory (limit 226 KB)", a.classes, a.d, smem);
    return GB_ERR_UNSUPPORTED;
  }
  const bool vec = (a.d % 4 == 0) && (a.ldz % 4 == 0) && (reinterpret_cast<uintptr_t>(a.z) % 16 == 0);
  static bool configured = false;
  if (!configured) {
    GB_CUDA(cudaFuncSetAttribute(decode_eval_kernel<CP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));   // + 24 B static
    GB_CUDA(cudaFuncSetAttribute(decode_eval_kernel<CP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    configured = true;
  }
  const int ntiles = (a.n + 32 * kNpt - 1) / (32 * kNpt);
  const int blocks = max(1, min((ntiles + kDecWarps - 1) / kDecWarps, sm_count() * 8));
  LaunchScope ls(K_DECODE_EVAL, stream);
  if (vec) decode_eval_kernel<CP, true><<<blocks, kDecThreads, smem, stream>>>(a);
  else decode_eval_kernel<CP, false><<<blocks, kDecThreads, smem, stream>>>(a);
  return check_launch("decode_eval_kernel");
}

}  // namespace gb

using namespace gb;

extern "C" {

int gb_decode_eval_max_classes(void) { return 64; }

int gb_decode_eval(const float* z, int32_t ldz, int32_t n, int32_t d, const float* w2, int32_t ldw, const float* b2,
                   int32_t classes, const int64_t* y, const uint8_t* train_mask, const uint8_t* val_mask,
                   const uint8_t* test_mask, int32_t* counts, float* logits, int32_t ldl, int64_t* pred, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(z && w2 && n > 0 && d > 0 && classes > 0, "gb_decode_eval: bad arguments");
  GB_REQUIRE(ldz >= d && ldw >= d, "gb_decode_eval: leading dimension too small");
  GB_REQUIRE(!logits || ldl >= classes, "gb_decode_eval: logits leading dimension too small");
  GB_REQUIRE(!counts || y, "gb_decode_eval: counts need labels");
  if (classes > 64) {
    set_error("gb_decode_eval: %d classes (limit 64: one accumulator register per class)", classes);
    return GB_ERR_UNSUPPORTED;
  }
  DecodeArgs a{};
  a.z = z; a.ldz = ldz; a.w2 = w2; a.ldw = ldw; a.b2 = b2; a.y = y;
  a.mask[0] = train_mask; a.mask[1] = val_mask; a.mask[2] = test_mask;
  a.counts = counts; a.logits = logits; a.ldl = ldl; a.pred = pred;
  a.n = n; a.d = d; a.classes = classes; a.dpad = (d + 3) & ~3;
  if (counts) GB_CUDA(cudaMemsetAsync(counts, 0, 6 * sizeof(int32_t), stream));
  if (classes <= 8) return decode_launch<8>(a, stream);
  if (classes <= 16) return decode_launch<16>(a, stream);
  if (classes <= 32) return decode_launch<32>(a, stream);
  if (classes <= 48) return decode_launch<48>(a, stream);
  return decode_launch<64>(a, stream);
}

}  // extern "C"
