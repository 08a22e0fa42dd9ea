// nl_adj.cu - adjoint (evaluation + VJP row pass) instantiations of the single-gather GRAND-nl kernels.
#include "nl2_kernels.cuh"

namespace gb {
int nl_impl();   // nl_fwd.cu

int nl_dispatch_adj(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
  if (nl_impl() == 2) {
    const int rc = nl2_dispatch<true>(a, heads, n_long, stream);
    if (rc != GB_ERR_UNSUPPORTED) return rc;
  }
  return nl_dispatch<true>(a, heads, n_long, stream);
}
}  // namespace gb
