/* grandb200.h - C ABI of libgrandb200.so: the B200 (sm_100a) graph-diffusion operator.
 *
 * This is the drop-in boundary for ONE hot path of alimt1992/graph-neural-pde: one evaluation of
 *     f(t, x) = alpha * (A(x) x - x) + beta * x0
 * (forward and vector-Jacobian product) as computed by the reference's
 *     LaplacianODEFunc.forward          src/function_laplacian_diffusion.py:39-77
 *     ODEFuncTransformerAtt.forward     src/function_transformer_attention.py:20-59
 *     SpGraphTransAttentionLayer.forward src/function_transformer_attention.py:159-267
 *     utils.softmax                     src/utils.py:116-127
 *     utils.get_rw_adj / gcn_norm_fill_val   src/utils.py:177-233
 * and their autograd backward (the reference has no hand-written backward; torchdiffeq's
 * odeint_adjoint calls autograd.grad on the same forward, src/base_classes.py:45-48).
 *
 * Conventions
 *  - Plain C: pointers + sizes, no torch types.  Every pointer is a DEVICE pointer unless its name
 *    ends in _host.  The library owns no memory between calls: the caller (torch's caching
 *    allocator in the Python host, any allocator elsewhere) provides outputs and workspaces.
 *  - `stream` is a cudaStream_t passed as void*.  All work is enqueued on it; no call synchronises
 *    the device except gb_plan_finalize_host (plan build, once per graph) and gb_prof_collect.
 *  - Return value: 0 = GB_OK, otherwise a gb_status; gb_last_error() gives the text
 *    (thread-local).  Nothing ever falls back to a CPU path.
 *  - Feature matrices are row-major fp32 with an explicit leading dimension (in floats).
 *  - Edge orientation: row = edge_index[0] = node summed INTO, col = edge_index[1] = node whose
 *    features are gathered (the orientation of sparse_coo_tensor([row, col]) @ x at
 *    function_laplacian_diffusion.py:41-57).
 *  - "CSR slot" s in [0,E): position after the stable sort of the edges by (row, col).
 *    "CSC slot" p in [0,E): position after the stable sort of the CSR slots by (col, row).
 *    Per-edge arrays handed to kernels are in CSR-slot order unless stated otherwise.
 */
#ifndef GRANDB200_H
#define GRANDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GB_ABI_VERSION 2

typedef enum {
  GB_OK = 0,
  GB_ERR_INVALID = 1,      /* bad argument (null pointer, size, alignment) */
  GB_ERR_CUDA = 2,         /* a CUDA runtime call or launch failed */
  GB_ERR_UNSUPPORTED = 3,  /* valid request the library does not implement */
  GB_ERR_WORKSPACE = 4     /* workspace too small */
} gb_status;

/* attention_type of SpGraphTransAttentionLayer (function_transformer_attention.py:246-259) */
typedef enum {
  GB_ATT_SCALED_DOT = 0,   /* per-edge q.k/sqrt(dk) (upstream semantics) */
  GB_ATT_COSINE = 1,       /* cosine similarity, eps 1e-5 */
  GB_ATT_PEARSON = 2,      /* mean-centred cosine similarity */
  GB_ATT_EXP_KERNEL = 3    /* output_var^2 * exp(-|q-k|^2 / (2 lengthscale^2)) */
} gb_att_type;

/* Immutable device-side description of one graph ("plan").  Built once per
 * ODEblock.reset_graph_data (src/base_classes.py:70-90) and reused by every evaluation. */
typedef struct {
  int32_t n;               /* nodes */
  int32_t e;               /* edges (duplicates kept as separate entries) */
  const int32_t* rowptr;   /* [n+1] CSR */
  const int32_t* colidx;   /* [e]   column of CSR slot s */
  const int32_t* eperm;    /* [e]   CSR slot -> caller's edge id (COO position) */
  const int32_t* colptr;   /* [n+1] CSC */
  const int32_t* rowidx;   /* [e]   row of CSC slot p */
  const int32_t* tperm;    /* [e]   CSC slot -> CSR slot */
  const int32_t* long_rows;/* [n_long_rows] rows with more than long_threshold edges (may be NULL) */
  const int32_t* long_cols;/* [n_long_cols] columns with more than long_threshold edges */
  int32_t n_long_rows;
  int32_t n_long_cols;
  int32_t long_threshold;
  int32_t reserved;
} gb_graph;

/* Optional Runge-Kutta stage fusion for gb_spmm_fwd: besides f the kernel also writes
 *   ynext = y + sum_{j<n_prev} c[j]*k[j] + c[n_prev]*f
 * (the stage combination torchdiffeq does with separate axpy kernels). */
typedef struct {
  const float* y;          /* [n, ld] state at the start of the step */
  const float* k[6];       /* previous stage derivatives */
  float c[7];              /* coefficients already multiplied by dt */
  int32_t n_prev;
  int32_t write_f;         /* 0: f itself is not needed (last stage), skip its store */
  float* ynext;            /* [n, ld] */
  int32_t ld;              /* leading dimension of y, k[*], ynext */
  int32_t reserved;
} gb_stage;

/* ---- library ------------------------------------------------------------------------------ */
int gb_abi_version(void);
const char* gb_last_error(void);
/* sm count / compute capability of the current device */
int gb_device_info(int32_t* sm_count_host, int32_t* cc_major_host, int32_t* cc_minor_host);
/* number of kernels this library launched since load (monotone; used for bench "gpu_launches") */
int64_t gb_launch_count(void);

/* per-kernel CUDA-event timing (off by default). ids are gb_kernel_id values. */
int gb_prof_enable(int on);
int gb_prof_reset(void);
int gb_prof_collect(void);                      /* synchronises the recorded events */
int gb_prof_get(int kernel_id, double* total_ms_host, int64_t* launches_host);
/* launches of the kernel that ran on the library's side stream next to the main kernel of their pass (hub rows /
 * columns): their event intervals include the wait for free SMs and overlap the main kernel's */
int64_t gb_prof_overlapped(int kernel_id);
int gb_prof_num_kernels(void);
const char* gb_prof_name(int kernel_id);

/* ---- graph plan (replaces utils.get_rw_adj / gcn_norm_fill_val scatter_add + the COO->dense
 *      conversion of sparse_multiply; src/utils.py:177-233, function_laplacian_diffusion.py:39-58) */
size_t gb_plan_workspace_bytes(int32_t n, int64_t e);
/* COO (int64, caller order) -> CSR + CSC arrays.  Bit-exact contract: oracle/graph.py coo_to_csr /
 * csr_to_csc.  All outputs are caller-allocated int32 device arrays. */
int gb_plan_build(const int64_t* row, const int64_t* col, int64_t e, int32_t n,
                  int32_t* rowptr, int32_t* colidx, int32_t* eperm,
                  int32_t* colptr, int32_t* rowidx, int32_t* tperm,
                  void* workspace, size_t workspace_bytes, void* stream);
/* Lists the segments longer than `threshold`. ptr = rowptr or colptr. Writes ids to `out` (capacity
 * `cap`) in increasing order and the count to *count_host (synchronises the stream). */
int gb_plan_long_segments(const int32_t* ptr, int32_t n, int32_t threshold, int32_t* out, int32_t cap,
                          int32_t* count_host, void* workspace, size_t workspace_bytes, void* stream);

/* deg[i] = sum of w over edges with idx == i, summed SEQUENTIALLY IN EDGE ORDER (the order of
 * torch's CPU scatter_add, so the fp32 bits match the oracle); utils.py:191,230 */
size_t gb_degree_workspace_bytes(int32_t n, int64_t e);
int gb_degree(const int64_t* idx, const float* w, int64_t e, int32_t n, float* deg,
              void* workspace, size_t workspace_bytes, void* stream);
/* utils.get_rw_adj: w_out = w * (1/deg)[idx]   (reciprocal, then multiply; utils.py:231-232) */
int gb_rw_normalize(const int64_t* idx, const float* w, const float* deg, int64_t e, float* w_out, void* stream);
/* utils.gcn_norm_fill_val: dis = deg^-1/2 (inf -> 0); w_out = (dis[row]*w)*dis[col]; utils.py:192-194 */
int gb_gcn_normalize(const int64_t* row, const int64_t* col, const float* w, const float* deg, int64_t e,
                     float* w_out, void* stream);

/* per-edge plumbing between the caller's COO order and CSR-slot order */
int gb_gather_f32(const float* src, const int32_t* perm, int64_t e, float* dst, void* stream);          /* dst[s] = src[perm[s]] */
int gb_head_mean_gather(const float* att_coo, int32_t heads, const int32_t* eperm, int64_t e,
                        float* abar_csr, void* stream);   /* mean over heads (function_laplacian_diffusion.py:46) */
int gb_scatter_rows_f32(const float* src_csr, int32_t width, const int32_t* eperm, int64_t e,
                        float* dst_coo, void* stream);    /* dst[eperm[s], :] = src[s, :] */

/* ---- K1: CSR SpMM + fused epilogue (LaplacianODEFunc.forward, function_laplacian_diffusion.py:60-77)
 *   out[i,:] = alpha * (sum_{s in row i} vals[s] * x[colidx[s],:] - x[i,:]) + beta * x0[i,:]
 *   alpha = sigmoid(*alpha_train) if alpha_sigmoid else *alpha_train ; beta = *beta_train.
 *   x0 == NULL or beta_train == NULL  -> no source term (opt['add_source'] False).
 *   stage == NULL -> plain evaluation. */
int gb_spmm_fwd(const gb_graph* g, const float* vals, const float* x, int32_t ldx,
                const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                const float* beta_train, int32_t d, float* out, int32_t ldo,
                const gb_stage* stage, void* stream);

/* Row-partitioned form: only rows [row_lo, row_hi) are evaluated (1-D row partition); x is
 * addressed by global column id, x0 / out / stage buffers by global row id. */
int gb_spmm_fwd_part(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* vals, const float* x,
                     int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                     const float* beta_train, int32_t d, float* out, int32_t ldo, const gb_stage* stage,
                     void* stream);

/* ---- K5a: transpose (CSC) SpMM + parameter-gradient reductions (VJP of K1 w.r.t. x, alpha_train,
 *      beta_train).  gout = dL/df.
 *   t[j,:]  = sum_{p in col j} vals[tperm[p]] * gout[rowidx[p],:] - gout[j,:]
 *   dx[j,:] = alpha * t[j,:]                      (accumulate != 0: dx += ...)
 *   dparams[0] = dalpha_train = (alpha_sigmoid ? alpha(1-alpha) : 1) * sum_j <x[j,:], t[j,:]>
 *   dparams[1] = dbeta_train  = sum_j <gout[j,:], x0[j,:]>          (0 if x0 == NULL)
 *   Reductions are two-pass and order-fixed (no float atomics). `partials` needs
 *   gb_spmm_bwd_partials(g, d) floats. */
size_t gb_spmm_bwd_partials(const gb_graph* g, int32_t d);   /* in floats */
int gb_spmm_bwd(const gb_graph* g, const float* vals, const float* gout, int32_t ldg,
                const float* x, int32_t ldx, const float* x0, int32_t ldx0,
                const float* alpha_train, int32_t alpha_sigmoid, int32_t d,
                float* dx, int32_t lddx, int32_t accumulate,
                float* partials, float* dparams, void* stream);

/* Row-partitioned form: the plan holds only this rank's rows; every column j gets this rank's share
 * sum_e a_e gout[row_e,:], and only columns in [own_lo, own_hi) also get the "- gout[j,:]" self term and
 * contribute <gout_j, x0_j> to dbeta (own_lo == own_hi == 0: all columns owned).  Summing dx / dparams over
 * the ranks (reduce-scatter / all-reduce) gives the single-GPU result. */
int gb_spmm_bwd_part(const gb_graph* g, int32_t own_lo, int32_t own_hi, const float* vals, const float* gout,
                     int32_t ldg, const float* x, int32_t ldx, const float* x0, int32_t ldx0,
                     const float* alpha_train, int32_t alpha_sigmoid, int32_t d, float* dx, int32_t lddx,
                     int32_t accumulate, float* partials, float* dparams, void* stream);

/* ---- gb_spmm_bwd without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 *      for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 *      harnesses that forbid it).  Same arguments, arithmetic and results as gb_spmm_bwd minus `partials`: the
 *      reduction scratch is a file-scope device array of GB_SPMM_BWD_STATIC_MAX_FLOATS floats (4 MiB, held by every
 *      GPU that loads the library until the process exits).  Graphs / widths whose gb_spmm_bwd_partials(g, d)
 *      exceeds that limit are refused with GB_ERR_WORKSPACE (use gb_spmm_bwd).  Calls no cudaMalloc / cudaFree.
 *      All calls on a device share the one array: calls that overlap on the GPU (different streams) race on it. */
#define GB_SPMM_BWD_STATIC_MAX_FLOATS (1 << 20)
int gb_spmm_bwd_static(const gb_graph* g, const float* vals, const float* gout, int32_t ldg,
                       const float* x, int32_t ldx, const float* x0, int32_t ldx0,
                       const float* alpha_train, int32_t alpha_sigmoid, int32_t d,
                       float* dx, int32_t lddx, int32_t accumulate, float* dparams, void* stream);

/* ---- K3: Q/K projection (nn.Linear x2, function_transformer_attention.py:224-225)
 *   out[i, j] = sum_c x[i,c] * w[j,c] + b[j],  j < nout   (w is torch's Linear.weight layout [nout, d];
 *   pass Q and K weights concatenated along dim 0 to get [q | k] rows in one pass) */
/* Tensor-core path (csrc/linear_tc.cu): tcgen05.mma kind::tf32 with a 3-term hi/lo split (fp32-accurate), TMEM
 * accumulators, persistent warp-specialised CTAs.  Taken by gb_linear_fwd / gb_linear_bwd_x when the reduction
 * length is a multiple of 32 and <= 128, the output width a multiple of 16 and <= 128, rows 16-byte aligned
 * (every GRAND-nl configuration of BASELINE.json).  gb_linear_fwd also takes it for ANY reduction length >= 32 when
 * n >= 12288 and the output qualifies (the encoder m1 / mx / mp of src/GNN.py:30-38, e.g. [N, 1433] x [1433, 64]):
 * slices of 128 input columns accumulate into `out`, a ragged last slice is zero-filled, rows that are not 16-byte
 * aligned are loaded element-wise.  Other shapes run the CUDA-core SGEMM.  Returns 1 / 0 (the un-sliced form). */
int gb_linear_tc_eligible(const float* a, int32_t lda, const float* w, int32_t ldw, const float* out, int32_t ldo,
                          int32_t nout, int32_t k, int32_t trans_w);
int gb_linear_fwd(const float* x, int32_t ldx, int32_t n, int32_t d, const float* w, const float* b,
                  int32_t nout, float* out, int32_t ldo, void* stream);
/* dx[i,c] (+)= sum_j dy[i,j] * w[j,c] */
int gb_linear_bwd_x(const float* dy, int32_t lddy, int32_t n, int32_t nout, const float* w, int32_t d,
                    float* dx, int32_t lddx, int32_t accumulate, void* stream);
/* dw[j,c] = sum_i dy[i,j]*x[i,c] ; db[j] = sum_i dy[i,j]   (split over i, fixed-order second pass) */
size_t gb_linear_bwd_w_workspace_bytes(int32_t n, int32_t d, int32_t nout);
int gb_linear_bwd_w(const float* dy, int32_t lddy, const float* x, int32_t ldx, int32_t n, int32_t d,
                    int32_t nout, float* dw, float* db, void* workspace, size_t workspace_bytes, void* stream);

/* ---- K2+K4: fused per-edge score + segment softmax + head mean
 *      (function_transformer_attention.py:240-266 + utils.py:116-127)
 *   s[e,h]   = score(q[row_e, h, :], k[col_e, h, :])      (x edge_w[e] if edge_w != NULL)
 *   att[e,h] = exp(s - max_g) / (sum_g exp(s - max_g) + 1e-16), groups g = rows (norm_idx 0) or
 *              columns (norm_idx 1)
 *   abar[e]  = mean_h att[e,h]
 *   att / abar / prods are in CSR-slot order; prods (raw scores) may be NULL.
 *   kparams: [output_var, lengthscale] device floats for GB_ATT_EXP_KERNEL, else NULL. */
int gb_edge_softmax_fwd(const gb_graph* g, int32_t norm_idx, const float* q, int32_t ldq,
                        const float* k, int32_t ldk, int32_t heads, int32_t dk, int32_t att_type,
                        const float* edge_w, const float* kparams,
                        float* att, float* abar, float* prods, void* stream);

/* ---- K5b: attention VJP.
 *  (1) SDDMM:        dabar[s] = alpha * <gout[row_s,:], x[col_s,:]>                        (CSR pass)
 *  (2) softmax bwd:  da[s,h] = dabar[s]/H (+ datt[s,h]);  ds[s,h] = att[s,h] * (da[s,h] - sum_{g} att[.,h]*da[.,h])
 *      (dabar = dL/d mean_h att [E], datt = dL/d att [E,H]; either may be NULL)                (segment pass)
 *  (3) dq[i,h,:] = sum_{s in row i} ds[s,h] * dscore/dq ;  dk[j,h,:] = sum_{p in col j} ds[.,h] * dscore/dk
 *  All reductions are segment reductions over the pre-sorted CSR/CSC orders: deterministic. */
int gb_sddmm(const gb_graph* g, const float* gout, int32_t ldg, const float* x, int32_t ldx, int32_t d,
             const float* alpha_train, int32_t alpha_sigmoid, float* dabar, void* stream);
int gb_edge_softmax_bwd(const gb_graph* g, int32_t norm_idx, const float* att, const float* dabar,
                        const float* datt, int32_t heads, const float* edge_w, float* ds, void* stream);
int gb_edge_score_bwd(const gb_graph* g, const float* ds, const float* q, int32_t ldq, const float* k,
                      int32_t ldk, int32_t heads, int32_t dk, int32_t att_type, const float* kparams,
                      const float* prods, float* dq, int32_t lddq, float* dk_out, int32_t lddk,
                      float* dkparams_partials, float* dkparams, void* stream);

/* ---- GAT scores + segment softmax + head mean (function_GAT_attention.py:117-141 replaces the per-edge gather of
 *      both endpoints' [dk,H] projections + `sum(a * edge_h)`; utils.py:116-127 the softmax).  The score factorises
 *      into per-node scalars the caller projects once per evaluation (s_src[i,h] = <a[:dk], (xW)[i,h,:]>,
 *      s_dst[j,h] = <a[dk:], (xW)[j,h,:]>, both [n, lds >= heads]):
 *   prods[s,h] = leaky_relu(s_src[row_s,h] + s_dst[col_s,h], slope)
 *   att[s,h]   = exp(prods - max_g) / (sum_g exp(prods - max_g) + 1e-16), groups = rows (norm_idx 0) / columns (1)
 *   abar[s]    = mean_h att[s,h]
 *   CSR-slot order; prods may be NULL; heads <= 8.  The backward of att is gb_edge_softmax_bwd. */
int gb_gat_softmax_fwd(const gb_graph* g, int32_t norm_idx, const float* s_src, const float* s_dst, int32_t lds,
                       int32_t heads, float slope, float* att, float* abar, float* prods, void* stream);

/* ---- utils.softmax(src, index, num_nodes) on ready-made scores (utils.py:116-127), for callers that hold their own
 *      [E,H] scores and a group id per edge rather than a graph plan: the host sorts the group ids once
 *      (ptr [n+1] = segment boundaries, perm [E] = position in sorted order -> caller's slot, NULL if already sorted);
 *   att[s,h] = exp(src[s,h] - max_g) / (sum_g exp(src[., h] - max_g) + 1e-16);  abar (head mean) may be NULL. heads <= 8
 *   long_list [n_long]: the groups with more than long_threshold members (one CTA each instead of one warp); may be
 *   NULL / 0 - then every group is taken by a warp whatever its size. */
int gb_segment_softmax_fwd(const int32_t* ptr, const int32_t* perm, int32_t n, int32_t heads, const float* scores,
                           float* att, float* abar, const int32_t* long_list, int32_t n_long, int32_t long_threshold,
                           void* stream);

/* ---- per-step evaluation of the early-stopping integrator in one launch (csrc/decode.cu).  Replaces
 *      EarlyStopDopri5/EarlyStopRK4.evaluate + ode_test (src/early_stop_solver.py:108-128, :88-98 and :210-226):
 *          logits = relu(z[:, :d]) w2^T + b2;   pred = argmax(logits, -1)     (first maximum; NaN counts as largest)
 *          counts[m]     = #{i : mask_m[i] and pred[i] == y[i]},  counts[3 + m] = #{i : mask_m[i]},  m = train, val, test
 *      z [n, ldz >= d] device fp32 (columns >= d - the augmented half - are not read); w2 [classes, ldw] = m2.weight,
 *      b2 [classes] = m2.bias (may be NULL); y [n] int64 labels; masks: one byte per node (torch.bool), each may be NULL;
 *      counts [6] int32 on the device, zeroed by the call (NULL: no counting); logits [n, ldl] and pred [n] int64 are
 *      optional outputs (NULL to skip).  classes <= gb_decode_eval_max_classes() (64), else GB_ERR_UNSUPPORTED.
 *      The accuracies are counts[m] / counts[3 + m] - divided by the caller after ONE 24-byte read-back (the
 *      reference synchronises six times per step with .item()). */
int gb_decode_eval_max_classes(void);
int gb_decode_eval(const float* z, int32_t ldz, int32_t n, int32_t d, const float* w2, int32_t ldw, const float* b2,
                   int32_t classes, const int64_t* y, const uint8_t* train_mask, const uint8_t* val_mask,
                   const uint8_t* test_mask, int32_t* counts, float* logits, int32_t ldl, int64_t* pred, void* stream);

/* ---- fused GRAND-nl evaluation (warp-per-row kernels, csrc/fused_att.cu) -------------------------
 * gb_att_fused_supported: 1 when the shapes/alignment fit the fused kernels (heads a power of two <= 8,
 * dk/4 a power of two, attention_dim <= 128, 16-byte aligned rows, d % 4 == 0, d <= 512; pass x == NULL to
 * ask about the attention-only mode).  The plan's long_threshold must be 256.
 *
 * gb_att_fwd_fused: score + segment softmax + head mean (+ when x != NULL, attention_norm_idx 0: the SpMM
 * and the diffusion epilogue of gb_spmm_fwd with the freshly computed abar) in ONE kernel.
 *   store != 0 : att [E,H] / abar [E] are written for every edge (a backward pass will follow)
 *   store == 0 : they are only used as scratch for segments longer than 256 edges (may be NULL if the
 *                plan has none); x == NULL requires store semantics (att and abar are the outputs). */
int gb_att_fused_supported(const float* q, int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk,
                           const float* x, int32_t ldx, int32_t d);
int gb_att_fwd_fused(const gb_graph* g, int32_t norm_idx, const float* q, int32_t ldq, const float* k, int32_t ldk,
                     int32_t heads, int32_t dk, int32_t att_type, const float* edge_w, const float* kparams,
                     float* att, float* abar, float* prods, int32_t store,
                     const float* x, int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train,
                     int32_t alpha_sigmoid, const float* beta_train, int32_t d, float* out, int32_t ldo,
                     const gb_stage* stage, void* stream);
/* VJP of the fused evaluation (scaled_dot, attention_norm_idx 0) given gout = dL/df:
 *   rows   : dabar = alpha <g_i, x[col]>, softmax backward -> ds [E,H], dq [N,A]
 *   columns: dx = alpha (A^T g - g) (direct part only; add dq Wq + dk Wk with gb_linear_bwd_x),
 *            dk [N,A], dparams = {dalpha_train, dbeta_train} (fixed-order reductions)
 * ds [E,H] and dabar_scratch [E] (needed only when the plan has long rows) are caller scratch;
 * partials needs gb_att_bwd_fused_partials(g) floats, 8-byte aligned. */
size_t gb_att_bwd_fused_partials(const gb_graph* g);
int gb_att_bwd_fused(const gb_graph* g, const float* q, int32_t ldq, const float* k, int32_t ldk, int32_t heads,
                     int32_t dk, const float* att, const float* abar, const float* gout, int32_t ldg,
                     const float* x, int32_t ldx, const float* x0, int32_t ldx0, const float* alpha_train,
                     int32_t alpha_sigmoid, int32_t d, float* ds, float* dabar_scratch, float* dq, int32_t lddq,
                     float* dk_out, int32_t lddk, float* dx, int32_t lddx, float* partials, float* dparams,
                     void* stream);

/* ---- Runge-Kutta stage combination (replaces torchdiffeq's chain of axpy kernels in the fixed-grid solvers,
 *      rk4_alt_step_func / _runge_kutta_step; third-party, call sites src/base_classes.py:45-48):
 *   out[i] = y[i] + sum_{j < nk} c_host[j] * k_host[j][i],  nk <= 7, one pass over `count` floats.
 *   k_host: HOST array of nk device pointers; c_host: HOST array of nk coefficients (already times dt). */
int gb_stage_combine(const float* y, const float* const* k_host, const float* c_host, int32_t nk, int64_t count,
                     float* out, void* stream);
/* The same combination stored to n_out (1..8) destinations at once: the caller's buffer plus, on a row-partitioned
 * graph, this rank's slab inside the gathered state matrix of every peer (peer pointers over NVLink: CUDA IPC or
 * symmetric memory) - a push-style all-gather fused into the kernel that produces the next stage's state.  The
 * caller orders the peers' reads after these writes (barrier / signal). */
int gb_stage_combine_multi(const float* y, const float* const* k_host, const float* c_host, int32_t nk, int64_t count,
                           float* const* outs_host, int32_t n_out, void* stream);

/* Sending half of a sparse halo exchange as one kernel over peer memory (NO COUNTERPART IN THE REFERENCE, which has no
 * multi-GPU graph path): request r (row idx[r] of src, or row r when idx is NULL) of segment s =
 * [seg_start_host[s], seg_start_host[s + 1]) is stored to row r - seg_start_host[s] of dst_host[s] (a peer's matrix:
 * symmetric memory / CUDA IPC pointer).  nseg <= 8; idx is a DEVICE array, seg_start_host [nseg + 1] and dst_host
 * [nseg] are HOST arrays; d % 4 == 0, 16-byte aligned rows.  The caller orders the peers' reads after the writes. */
int gb_rows_push(const float* src, int32_t ld, const int64_t* idx, int32_t nseg, const int64_t* seg_start_host,
                 float* const* dst_host, int32_t dst_ld, int32_t d, void* stream);

/* ---- single-gather GRAND-nl evaluation and VJP (csrc/nl_kernels.cuh; scaled_dot, attention_norm_idx 0) ----
 * One row gather per edge: scores are formed as <Wk_h^T q_ih / sqrt(dk), x_j> (the per-(row, head) constant
 * <q_ih, bk_h> cancels in the row softmax), the softmax is online, so neither k = x Wk^T + bk nor the
 * attention [E,H] is ever materialised in the forward pass.  Replaces, per evaluation,
 * ODEFuncTransformerAtt.forward (function_transformer_attention.py:44-59) after the Q projection.
 *
 * Rows [row_lo, row_hi) are evaluated (1-D row partition; pass 0, g->n on one GPU).  Matrices indexed by the
 * row (q, x0, out, gout, dq) are addressed with the GLOBAL row id; x is addressed by global column id.
 * workspace: gb_nl_workspace_bytes() bytes of device memory (scheduler counter). */
int gb_nl_supported(int32_t heads, int32_t dk, int32_t d);   /* heads in {1,2,4,8}, dk % 4 == 0, d % 4 == 0, d <= 128, heads * dk <= 128 */
size_t gb_nl_workspace_bytes(void);
int gb_nl_fwd(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx,
              const float* q, int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk, int32_t d,
              const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
              const float* beta_train, float* out, int32_t ldo, const gb_stage* stage,
              void* workspace, void* stream);
/* Evaluation + VJP in one row pass and one column pass, given gout = dL/df (rows [row_lo,row_hi)):
 *   rows    : [out = f when out != NULL], dq [rows, A], ds [E,H], abar [E]   (dabar_scratch [E] is caller scratch;
 *             the three per-edge buffers may be NULL when the graph has no edges: f = -alpha x + beta x0, dq = dk = 0)
 *   columns : dx[j,:] = alpha (sum_{e: col_e = j} abar_e gout[row_e,:] - [j owned] gout[j,:])   for ALL columns j
 *             dk[j,:] = sum_e ds[e,:] q[row_e,:] / sqrt(dk)                                      for ALL columns j
 *             dparams = {dalpha_train, dbeta_train} (this rank's share when partitioned)
 * The caller finishes with dense GEMMs: dx += dq Wq + dk Wk, dWq = dq^T x, dWk = dk^T x, db = column sums.
 * partials: gb_att_bwd_fused_partials(g) floats, 8-byte aligned. */
int gb_nl_adjoint(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx,
                  const float* q, int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk, int32_t d,
                  const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                  const float* beta_train, const float* gout, int32_t ldg, float* out, int32_t ldo,
                  float* ds, float* dabar_scratch, float* abar, float* dq, int32_t lddq,
                  float* dk_out, int32_t lddk, float* dx, int32_t lddx, float* partials, float* dparams,
                  void* workspace, void* stream);

/* The same call in two halves, so that a host that exchanges states between devices can start the exchange of the
 * NEXT stage's state (which depends only on f) while the column pass of this evaluation runs:
 *   GB_NL_PHASE_ROWS : the row pass only  -> out (f), dq, ds, abar (and the dabar scratch) are complete on return
 *   GB_NL_PHASE_COLS : the column pass only (same arguments; reads q, abar, ds, gout, x) -> dx, dk, dparams
 *   GB_NL_PHASE_BOTH : gb_nl_adjoint. */
enum { GB_NL_PHASE_BOTH = 0, GB_NL_PHASE_ROWS = 1, GB_NL_PHASE_COLS = 2 };
int gb_nl_adjoint_phase(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx,
                        const float* q, int32_t ldq, const float* wk, int32_t ldwk, int32_t heads, int32_t dk,
                        int32_t d, const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                        const float* beta_train, const float* gout, int32_t ldg, float* out, int32_t ldo,
                        float* ds, float* dabar_scratch, float* abar, float* dq, int32_t lddq,
                        float* dk_out, int32_t lddk, float* dx, int32_t lddx, float* partials, float* dparams,
                        void* workspace, int32_t phase, void* stream);

/* ---- K-gather form of the fused GRAND-nl kernels (csrc/nlk_kernels.cuh): same function, results and buffers as
 *      gb_nl_fwd / gb_nl_adjoint_phase (ODEFuncTransformerAtt.forward, function_transformer_attention.py:44-59, and
 *      its VJP), but the caller projects k = x Wk^T + bk as well (one GEMM for [q | k]) and passes it instead of Wk:
 *      the kernels gather x[col] AND k[col] per edge and do no per-row Wk work.  Faster whenever d > attention_dim
 *      (measured, profiles/r02*); one GPU / un-partitioned rows or any row range whose columns have k rows. */
int gb_nlk_supported(int32_t heads, int32_t dk, int32_t d);
int gb_nlk_fwd(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx, const float* q,
               int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk, int32_t d, const float* x0,
               int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid, const float* beta_train, float* out,
               int32_t ldo, const gb_stage* stage, void* workspace, void* stream);
int gb_nlk_adjoint_phase(const gb_graph* g, int32_t row_lo, int32_t row_hi, const float* x, int32_t ldx,
                         const float* q, int32_t ldq, const float* k, int32_t ldk, int32_t heads, int32_t dk,
                         int32_t d, const float* x0, int32_t ldx0, const float* alpha_train, int32_t alpha_sigmoid,
                         const float* beta_train, const float* gout, int32_t ldg, float* out, int32_t ldo, float* ds,
                         float* dabar_scratch, float* abar, float* dq, int32_t lddq, float* dk_out, int32_t lddk,
                         float* dx, int32_t lddx, float* partials, float* dparams, void* workspace, int32_t phase,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRANDB200_H */
