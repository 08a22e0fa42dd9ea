"""CPU oracle for the GRAND/BLEND per-step graph-diffusion operator.

TEST INFRASTRUCTURE ONLY.  Nothing under `oracle/` is part of the product: only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py` may import
it, and there only as the checker (or as the timed CPU baseline), never as the thing shipped.  The
product package (`graph-neural-pde_b200/`) never imports it and fails loudly when its CUDA
library is missing.

What it restates (all file:line citations are relative to `/root/reference`):

* `graph.py`   - self-loop insertion, random-walk / GCN normalisation, COO -> CSR/CSC ordering
                 (`src/utils.py:16-42,177-233`, `src/base_classes.py:70-90`; un-batched intent,
                 PyG 1.7 `add_remaining_self_loops` semantics because the fork's helper is broken,
                 SURVEY.md F3).
* `funcs.py`   - segment softmax (`src/utils.py:116-127`), the attention layer
                 (`src/function_transformer_attention.py:159-267`), `LaplacianODEFunc.forward`
                 (`src/function_laplacian_diffusion.py:39-77`) and
                 `ODEFuncTransformerAtt.forward` (`src/function_transformer_attention.py:44-59`).
* `integrate.py` - the torchdiffeq 0.2.x solvers the reference calls (`src/base_classes.py:45-48`,
                 `src/block_constant.py:34-51`): euler / midpoint / rk4 (3/8 rule) / dopri5 /
                 adaptive_heun and the adjoint driver.  torchdiffeq is a third-party dependency
                 that is neither vendored in the reference nor installable here (README pin 0.2.1),
                 so this part restates its published algorithm and is **parity unpinned**.

Pinning status (see `oracle/make_golden.py`, `tests/test_oracle_golden.py`):
the graph/softmax/ODE-function parts are pinned against outputs of the reference's OWN modules
imported unmodified in the build container (with the import-only stubs in `oracle/ref_stubs/`) and
stored in `tests/golden/*.npz`, plus the reference's known-answer tests.
The one deliberate difference - `scaled_dot` scores are per-edge q.k/sqrt(dk) (upstream /
north_star semantics) instead of the fork's E x E sum (`function_transformer_attention.py:249`) -
is documented by a `fork_literal=True` switch that reproduces the fork bit-for-bit on tiny graphs.
"""
