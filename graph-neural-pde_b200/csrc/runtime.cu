// runtime.cu - error reporting, launch accounting and per-kernel event timing of libgrandb200.
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>
#include <mutex>
#include <vector>
#include <nvtx3/nvToolsExt.h>
#include "common.cuh"

namespace gb {

static thread_local std::string g_error;
static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_prof_on{0};

struct EventRec {
  int id;
  bool side;                                   // recorded on the side stream: the interval overlaps the main kernel
  cudaEvent_t start, stop;
};
static std::mutex g_prof_mu;
static std::vector<EventRec*> g_pending;       // recorded, not yet collected
static std::vector<EventRec*> g_free;          // event pool
static double g_total_ms[K_NUM_KERNELS];
static int64_t g_count[K_NUM_KERNELS];
static int64_t g_side_count[K_NUM_KERNELS];    // launches recorded on the side stream

static const char* kNames[K_NUM_KERNELS] = {
    "spmm_fwd",           "spmm_fwd_long",     "spmm_bwd",          "spmm_bwd_long",     "reduce_partials",
    "linear_fwd",         "linear_bwd_x",      "linear_bwd_w",      "linear_bwd_w_reduce", "edge_softmax_fwd",
    "edge_softmax_fwd_long", "sddmm",          "edge_softmax_bwd",  "edge_score_bwd_q",  "edge_score_bwd_k",
    "plan",               "edge_util",         "fused_att_spmm_fwd", "fused_att_bwd_row", "fused_att_bwd_col",
    "linear_tc",          "stage_axpy",        "nl_fwd",            "nl_fwd_long",       "nl_adj_rows",
    "nl_adj_long",        "gat_softmax",       "decode_eval"};

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_error = buf;
}

int cuda_fail(cudaError_t err, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)err, cudaGetErrorString(err), file, line, what);
  return GB_ERR_CUDA;
}

int check_launch(const char* name) {
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) {
    set_error("launch of %s failed: %s", name, cudaGetErrorString(err));
    return GB_ERR_CUDA;
  }
  return GB_OK;
}

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}

int persistent_sms() {
  static int reserve = -1;
  if (reserve < 0) {
    const char* e = getenv("GRANDB200_SM_RESERVE");
    reserve = e != nullptr ? atoi(e) : 0;
    if (reserve < 0) reserve = 0;
  }
  const int n = sm_count() - reserve;
  return n < 1 ? 1 : n;
}

// ---- side stream: the CTA-per-long-segment kernels run next to the main kernel of the same pass ------------
// A handful of hub rows / columns (one CTA each, up to ~1 ms for a 20 000-edge hub) would otherwise run alone
// on a few SMs after the main kernel.  They touch disjoint output rows, so they are launched FIRST on a side
// stream that is ordered after everything already enqueued on the caller's stream (fork) and joined back
// before the next dependent launch.  Works under stream capture (the side stream joins the capture through the
// fork event and leaves it at the join).  GRANDB200_NO_SIDE_STREAM=1 keeps everything on the caller's stream.
namespace {
struct SideStream {
  cudaStream_t s = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool failed = false;
};
SideStream g_side[16];
std::mutex g_side_mu;
}  // namespace

cudaStream_t side_fork(cudaStream_t main) {
  static const bool off = [] {
    const char* e = getenv("GRANDB200_NO_SIDE_STREAM");
    return e && e[0] && e[0] != '0';
  }();
  if (off) return main;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return main;
  std::lock_guard<std::mutex> lk(g_side_mu);
  SideStream& ss = g_side[dev];
  if (ss.failed) return main;
  if (!ss.s) {
    int lo = 0, hi = 0;                               // highest priority: the few hub CTAs are placed before the
    cudaDeviceGetStreamPriorityRange(&lo, &hi);       // main kernel's CTAs fill every SM
    if (cudaStreamCreateWithPriority(&ss.s, cudaStreamNonBlocking, hi) != cudaSuccess ||
        cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming) != cudaSuccess) {
      ss.failed = true;
      ss.s = nullptr;
      cudaGetLastError();
      return main;
    }
  }
  if (cudaEventRecord(ss.fork, main) != cudaSuccess || cudaStreamWaitEvent(ss.s, ss.fork, 0) != cudaSuccess) {
    cudaGetLastError();
    return main;
  }
  return ss.s;
}

static bool is_side_stream(cudaStream_t s) {
  if (s == nullptr) return false;
  std::lock_guard<std::mutex> lk(g_side_mu);
  for (const SideStream& ss : g_side)
    if (ss.s == s) return true;
  return false;
}

int side_join(cudaStream_t main, cudaStream_t side) {
  if (side == main) return GB_OK;
  int dev = 0;
  GB_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(g_side_mu);
  SideStream& ss = g_side[dev];
  GB_CUDA(cudaEventRecord(ss.join, side));
  GB_CUDA(cudaStreamWaitEvent(main, ss.join, 0));
  return GB_OK;
}

// GRANDB200_NVTX=1: every launch sits in an NVTX range named after its kernel id (nsys / ncu --nvtx timelines)
static bool nvtx_on() {
  static const bool on = [] {
    const char* e = getenv("GRANDB200_NVTX");
    return e != nullptr && e[0] == '1';
  }();
  return on;
}

LaunchScope::LaunchScope(int kernel_id, cudaStream_t s) : id(kernel_id), stream(s), rec(nullptr) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  if (nvtx_on()) nvtxRangePushA(kernel_id >= 0 && kernel_id < K_NUM_KERNELS ? kNames[kernel_id] : "gb_kernel");
  if (!g_prof_on.load(std::memory_order_relaxed)) return;
  EventRec* r = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_free.empty()) {
      r = g_free.back();
      g_free.pop_back();
    }
  }
  if (!r) {
    r = new EventRec();
    if (cudaEventCreate(&r->start) != cudaSuccess || cudaEventCreate(&r->stop) != cudaSuccess) {
      delete r;
      return;
    }
  }
  r->id = kernel_id;
  r->side = is_side_stream(stream);
  cudaEventRecord(r->start, stream);
  rec = r;
}

LaunchScope::~LaunchScope() {
  if (nvtx_on()) nvtxRangePop();
  if (!rec) return;
  EventRec* r = static_cast<EventRec*>(rec);
  cudaEventRecord(r->stop, stream);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_pending.push_back(r);
}

}  // namespace gb

using namespace gb;

extern "C" {

int gb_abi_version(void) { return GB_ABI_VERSION; }

const char* gb_last_error(void) { return g_error.c_str(); }

int gb_device_info(int32_t* sm_count_host, int32_t* cc_major_host, int32_t* cc_minor_host) {
  int dev = 0;
  GB_CUDA(cudaGetDevice(&dev));
  int sms = 0, maj = 0, min = 0;
  GB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GB_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
  GB_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count_host) *sm_count_host = sms;
  if (cc_major_host) *cc_major_host = maj;
  if (cc_minor_host) *cc_minor_host = min;
  return GB_OK;
}

int64_t gb_launch_count(void) { return g_launches.load(); }

int gb_prof_enable(int on) {
  g_prof_on.store(on ? 1 : 0);
  return GB_OK;
}

int gb_prof_reset(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (EventRec* r : g_pending) g_free.push_back(r);
  g_pending.clear();
  for (int i = 0; i < K_NUM_KERNELS; ++i) {
    g_total_ms[i] = 0.0;
    g_count[i] = 0;
    g_side_count[i] = 0;
  }
  return GB_OK;
}

int gb_prof_collect(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (EventRec* r : g_pending) {
    GB_CUDA(cudaEventSynchronize(r->stop));
    float ms = 0.f;
    GB_CUDA(cudaEventElapsedTime(&ms, r->start, r->stop));
    g_total_ms[r->id] += ms;
    g_count[r->id] += 1;
    if (r->side) g_side_count[r->id] += 1;
    g_free.push_back(r);
  }
  g_pending.clear();
  return GB_OK;
}

int gb_prof_get(int kernel_id, double* total_ms_host, int64_t* launches_host) {
  GB_REQUIRE(kernel_id >= 0 && kernel_id < K_NUM_KERNELS, "gb_prof_get: bad kernel id %d", kernel_id);
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (total_ms_host) *total_ms_host = g_total_ms[kernel_id];
  if (launches_host) *launches_host = g_count[kernel_id];
  return GB_OK;
}

int64_t gb_prof_overlapped(int kernel_id) {
  if (kernel_id < 0 || kernel_id >= K_NUM_KERNELS) return 0;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  return g_side_count[kernel_id];
}

int gb_prof_num_kernels(void) { return K_NUM_KERNELS; }

const char* gb_prof_name(int kernel_id) {
  if (kernel_id < 0 || kernel_id >= K_NUM_KERNELS) return "";
  return kNames[kernel_id];
}

}  // extern "C"
