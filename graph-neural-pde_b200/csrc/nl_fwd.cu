// nl_fwd.cu - forward instantiations of the single-gather GRAND-nl kernels (see nl_kernels.cuh) and gb_nl_fwd.
#include <stdlib.h>
#include "nl2_kernels.cuh"

namespace gb {
// GRANDB200_NL_IMPL=1 selects the first-generation row kernels (register-resident gathers, nl_kernels.cuh) for A/B
// timing; default: the shared-memory-staged second generation (nl2_kernels.cuh) wherever it covers the shape.
int nl_impl() {
  static int impl = 0;
  if (impl == 0) {
    const char* s = getenv("GRANDB200_NL_IMPL");
    impl = (s != nullptr && s[0] == '1') ? 1 : 2;
  }
  return impl;
}

int nl_dispatch_fwd(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
  if (nl_impl() == 2) {
    const int rc = nl2_dispatch<false>(a, heads, n_long, stream);
    if (rc != GB_ERR_UNSUPPORTED) return rc;
  }
  return nl_dispatch<false>(a, heads, n_long, stream);
}

// argument checks shared by gb_nl_fwd / gb_nl_adjoint; fills the common part of NlArgs
int nl_fill_args(NlArgs& a, const gb_graph* g, int row_lo, int row_hi, const float* x, int ldx, const float* q,
                 int ldq, const float* wk, int ldwk, int heads, int dk, int d, const float* x0, int ldx0,
                 const float* alpha_train, int alpha_sigmoid, const float* beta_train, float* out, int ldo,
                 const gb_stage* stage, void* workspace, const char* who) {
  GB_REQUIRE(g && g->n > 0 && g->rowptr && (g->e == 0 || g->colidx), "%s: bad graph", who);
  GB_REQUIRE(row_lo >= 0 && row_lo <= row_hi && row_hi <= g->n, "%s: bad row range [%d, %d)", who, row_lo, row_hi);
  GB_REQUIRE(x && q && wk && alpha_train && workspace, "%s: null pointer", who);
  GB_REQUIRE(gb_nl_supported(heads, dk, d), "%s: unsupported shape heads=%d dk=%d d=%d", who, heads, dk, d);
  GB_REQUIRE(ldx >= d && ldq >= heads * dk && ldwk >= d, "%s: leading dimension too small", who);
  GB_REQUIRE(vec_width(x, ldx, d) == 4 && vec_width(wk, ldwk, d) == 4, "%s: x / Wk rows must be 16-byte aligned", who);
  GB_REQUIRE(g->long_threshold > 0 || g->e <= 1024, "%s: plan without long-segment lists", who);
  a = NlArgs{};
  a.rowptr = g->rowptr; a.colidx = g->colidx; a.row_lo = row_lo; a.row_hi = row_hi;
  a.dv = d / 4; a.adim = heads * dk; a.dk = dk; a.q = q; a.ldq = ldq; a.wk = wk; a.ldwk = ldwk;
  a.counter = static_cast<int32_t*>(workspace);
  a.long_list = g->long_rows;
  // rows above this are worked by a whole CTA.  Few listed rows (small graphs: 9 on the ogbn-arxiv shape, the longest
  // 1367 edges = 171 batches on ONE warp, a third of the launch) -> all of them; many (C4 / C5) -> only the real hubs
  const int cta_rows = g->n_long_rows <= 4 * sm_count() ? g->long_threshold : kNlLongRow;
  a.long_threshold = g->long_threshold > 0 ? (g->long_threshold > cta_rows ? g->long_threshold : cta_rows) : 0x7fffffff;
  SpmmArgs& sp = a.sp;
  sp.src = x; sp.lds = ldx; sp.alpha_train = alpha_train; sp.alpha_sigmoid = alpha_sigmoid; sp.n = g->n; sp.dv = d / 4;
  const bool source = x0 != nullptr && beta_train != nullptr;
  sp.x0 = source ? x0 : nullptr; sp.ldx0 = ldx0; sp.beta_train = source ? beta_train : nullptr;
  sp.out = out; sp.ldo = ldo;
  if (out) GB_REQUIRE(ldo >= d && vec_width(out, ldo, d) == 4, "%s: out rows must be 16-byte aligned", who);
  if (source) GB_REQUIRE(ldx0 >= d && vec_width(x0, ldx0, d) == 4, "%s: x0 rows must be 16-byte aligned", who);
  if (stage) {
    GB_REQUIRE(stage->y && stage->ynext && stage->n_prev >= 0 && stage->n_prev <= 6 && stage->ld >= d, "%s: bad stage", who);
    int v = min_int(vec_width(stage->y, stage->ld, d), vec_width(stage->ynext, stage->ld, d));
    for (int j = 0; j < stage->n_prev; ++j) {
      GB_REQUIRE(stage->k[j] != nullptr, "%s: null stage k[%d]", who, j);
      v = min_int(v, vec_width(stage->k[j], stage->ld, d));
    }
    GB_REQUIRE(v == 4, "%s: stage buffers must be 16-byte aligned", who);
    sp.has_stage = 1; sp.stage = *stage;
  }
  return GB_OK;
}
}  // namespace gb

using namespace gb;

extern "C" {

int gb_nl_supported(int32_t heads, int32_t dk, int32_t d) {
  const bool h_ok = heads == 1 || heads == 2 || heads == 4 || heads == 8;
  // attention_dim <= 128: the limit of the column pass of the VJP (one float4 of dk per lane); wider layers take the
  // composite kernels for the forward evaluation too, so that a function never evaluates on one route and fails on the
  // other when it is differentiated
  return h_ok && dk > 0 && dk % 4 == 0 && d > 0 && d % 4 == 0 && d <= 128 && heads * dk <= 128;
}

size_t gb_nl_workspace_bytes(void) { return 256; }

int gb_nl_fwd(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
              int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk, int32_t d, const float* x0,
              int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, float* out,
              int32_t ldo, const gb_stage* stage, void* workspace, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  NlArgs a;
  if (int rc = nl_fill_args(a, g, row_lo, row_hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train,
                            alpha_sigmoid, beta_train, out, ldo, stage, workspace, "gb_nl_fwd"))
    return rc;
  GB_REQUIRE(out != nullptr, "gb_nl_fwd: null out");
  return nl_dispatch_fwd(a, heads, g->long_threshold > 0 ? g->n_long_rows : 0, stream);
}

}  // extern "C"
