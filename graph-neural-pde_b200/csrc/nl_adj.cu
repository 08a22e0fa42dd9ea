// nl_adj.cu - adjoint (evaluation + VJP row pass) instantiations of the single-gather GRAND-nl kernels.
#include "nl_kernels.cuh"

namespace gb {
int nl_dispatch_adj(const NlArgs& a, int heads, int n_long, cudaStream_t stream) {
  return nl_dispatch<true>(a, heads, n_long, stream);
}
}  // namespace gb
