// nlk_fwd.cu - forward instantiations of the k-gather GRAND-nl row kernels (nlk_kernels.cuh), gb_nlk_fwd and the
// argument checks shared with gb_nlk_adjoint_phase.
#include "nlk_kernels.cuh"

namespace gb {
int nl_fill_args(NlArgs& a, const gb_graph* g, int row_lo, int row_hi, const float* x, int ldx, const float* q,
                 int ldq, const float* wk, int ldwk, int heads, int dk, int d, const float* x0, int ldx0,
                 const float* alpha_train, int alpha_sigmoid, const float* beta_train, float* out, int ldo,
                 const gb_stage* stage, void* workspace, const char* who);   // nl_fwd.cu

int nlk_dispatch_fwd(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
  return nlk_dispatch<false>(a, heads, n_long, stream);
}

// NlArgs of the k-gather kernels: the `wk` / `ldwk` fields carry k / ldk
int nlk_fill_args(NlArgs& a, const gb_graph* g, int row_lo, int row_hi, const float* x, int ldx, const float* q,
                  int ldq, const float* k, int ldk, int heads, int dk, int d, const float* x0, int ldx0,
                  const float* alpha_train, int alpha_sigmoid, const float* beta_train, float* out, int ldo,
                  const gb_stage* stage, void* workspace, const char* who) {
  GB_REQUIRE(gb_nlk_supported(heads, dk, d), "%s: unsupported shape heads=%d dk=%d d=%d", who, heads, dk, d);
  GB_REQUIRE(k != nullptr && q != nullptr, "%s: null q / k", who);
  const int adim = heads * dk;
  GB_REQUIRE(ldk >= adim && vec_width(k, ldk, adim) == 4 && ldq >= adim && vec_width(q, ldq, adim) == 4,
             "%s: q / k rows must be 16-byte aligned", who);
  // the common checks and fields (x stands in for the Wk argument of the single-gather kernels)
  if (int rc = nl_fill_args(a, g, row_lo, row_hi, x, ldx, q, ldq, x, ldx, heads, dk, d, x0, ldx0, alpha_train,
                            alpha_sigmoid, beta_train, out, ldo, stage, workspace, who))
    return rc;
  a.wk = k;
  a.ldwk = ldk;
  return GB_OK;
}
}  // namespace gb

using namespace gb;

extern "C" {

int gb_nlk_supported(int32_t heads, int32_t dk, int32_t d) {
  const bool h_ok = heads == 1 || heads == 2 || heads == 4 || heads == 8;
  return h_ok && dk > 0 && dk % 4 == 0 && d > 0 && d % 4 == 0 && d <= 128 && heads * dk <= 128;
}

int gb_nlk_fwd(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
               int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk, int32_t d, const float* x0,
               int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, float* out,
               int32_t ldo, const gb_stage* stage, void* workspace, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  NlArgs a;
  if (int rc = nlk_fill_args(a, g, row_lo, row_hi, x, ldx, q, ldq, k, ldk, heads, dk, d, x0, ldx0, alpha_train,
                             alpha_sigmoid, beta_train, out, ldo, stage, workspace, "gb_nlk_fwd"))
    return rc;
  GB_REQUIRE(out != nullptr, "gb_nlk_fwd: null out");
  return nlk_dispatch_fwd(a, heads, g->long_threshold > 0 ? g->n_long_rows : 0, stream);
}

}  // extern "C"
