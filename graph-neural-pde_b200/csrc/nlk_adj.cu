// nlk_adj.cu - adjoint (evaluation + VJP row pass) instantiations of the k-gather GRAND-nl row kernels and
// gb_nlk_adjoint_phase (row pass here, column pass = the one of gb_nl_adjoint_phase, fused_att.cu).
#include "nlk_kernels.cuh"

namespace gb {
int nlk_fill_args(NlArgs& a, const gb_graph* g, int row_lo, int row_hi, const float* x, int ldx, const float* q,
                  int ldq, const float* k, int ldk, int heads, int dk, int d, const float* x0, int ldx0,
                  const float* alpha_train, int alpha_sigmoid, const float* beta_train, float* out, int ldo,
                  const gb_stage* stage, void* workspace, const char* who);   // nlk_fwd.cu
int nl_cols_pass(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
                 int32_t ldq, int32_t heads, int32_t dk, int32_t d, const float* x0, int32_t ldx0,
                 const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, const float* gout,
                 int32_t ldg, float* ds, float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk, float* dx,
                 int32_t lddx, float* partials, float* dparams, cudaStream_t stream);   // fused_att.cu
}  // namespace gb

using namespace gb;

extern "C" {

int gb_nlk_adjoint_phase(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
                         int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk, int32_t d,
                         const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                         const float* beta_train, const float* gout, int32_t ldg, float* out, int32_t ldo, float* ds,
                         float* dabar_scratch, float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk,
                         float* dx, int32_t lddx, float* partials, float* dparams, void* workspace, int32_t phase,
                         void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  GB_REQUIRE(phase == GB_NL_PHASE_BOTH || phase == GB_NL_PHASE_ROWS || phase == GB_NL_PHASE_COLS,
             "gb_nlk_adjoint_phase: bad phase");
  NlArgs a;
  if (int rc = nlk_fill_args(a, g, row_lo, row_hi, x, ldx, q, ldq, k, ldk, heads, dk, d, x0, ldx0, alpha_train,
                             alpha_sigmoid, beta_train, out, ldo, nullptr, workspace, "gb_nlk_adjoint_phase"))
    return rc;
  GB_REQUIRE(gout && (g->e == 0 || (ds && dabar_scratch && abar)) && dq && dk_out && dx && partials, "gb_nlk_adjoint_phase: null pointer");
  GB_REQUIRE(g->colptr && (g->e == 0 || (g->rowidx && g->tperm)), "gb_nlk_adjoint_phase: graph without CSC arrays");
  GB_REQUIRE(ldg >= d && lddx >= d && vec_width(gout, ldg, d) == 4 && vec_width(dx, lddx, d) == 4,
             "gb_nlk_adjoint_phase: gout / dx rows must be 16-byte aligned");
  const int adim = heads * dk;
  GB_REQUIRE(lddq >= adim && lddk >= adim && vec_width(dq, lddq, adim) == 4 && vec_width(dk_out, lddk, adim) == 4,
             "gb_nlk_adjoint_phase: dq / dk rows must be 16-byte aligned");
  GB_REQUIRE(reinterpret_cast<uintptr_t>(partials) % 8 == 0, "gb_nlk_adjoint_phase: partials must be 8-byte aligned");
  const bool use_long = g->long_threshold > 0;
  GB_REQUIRE(!use_long || g->long_threshold == 256, "gb_nlk_adjoint_phase: plan long_threshold must be 256");
  GB_REQUIRE(use_long || g->e <= 256, "gb_nlk_adjoint_phase: plan without long-segment lists");
  a.write_f = out != nullptr;
  a.g = gout; a.ldg = ldg; a.ds = ds; a.dabar = dabar_scratch; a.abar = abar; a.dq = dq; a.lddq = lddq;
  if (phase != GB_NL_PHASE_COLS) {
    if (int rc = nlk_dispatch<true>(a, heads, use_long ? g->n_long_rows : 0, stream)) return rc;
  }
  if (phase == GB_NL_PHASE_ROWS) return GB_OK;
  return nl_cols_pass(g, row_lo, row_hi, x, ldx, q, ldq, heads, dk, d, x0, ldx0, alpha_train, alpha_sigmoid, beta_train,
                      gout, ldg, ds, abar, dq, lddq, dk_out, lddk, dx, lddx, partials, dparams, stream);
}

}  // extern "C"
