"""Oracle: graph construction for the diffusion operator (TEST INFRASTRUCTURE ONLY).

Restates, un-batched ([2,E] / [E]) and with the intent pinned by the reference's own tests
(`test/test_utils.py:27-30,62-77`, `test/test_function_laplacian_diffusion.py:68-85`):

* `add_remaining_self_loops`  - `src/utils.py:16-42` is broken in the fork (sorts the two index
  columns independently); upstream called PyG 1.7's helper, whose published
  behaviour is restated here: drop existing self-loops from the edge list, then append one loop
  (i,i) per node i, carrying the existing loop's weight if the node had one, else `fill_value`.
* `get_rw_adj`                - `src/utils.py:215-233`: deg = scatter_add(w, index); w * deg^-1.
* `gcn_norm_fill_val`         - `src/utils.py:177-194`: deg^-1/2[row] * w * deg^-1/2[col], inf -> 0.
* `reset_graph_data`          - `src/base_classes.py:70-90`: normalise, then (if
  self_loop_weight > 0) add remaining self loops again, which is a no-op under PyG semantics.
* `coo_to_csr` / `csr_to_csc` - the ordering contract of the CUDA plan builder: stable sort by
  (row, col) / (col, row); duplicates are kept as separate entries.

All integer work is numpy int64; weights are float32 and every float op is done in the order
the reference does it (sequential scatter_add in edge order, reciprocal THEN multiply) so that the
CUDA plan builder can be compared bit-for-bit.
"""
import numpy as np


def _seq_scatter_add_f32(w, idx, n):
    """float32 scatter_add with the summation order of torch's CPU kernel (sequential over the
    edge list) - np.add.at is sequential and unbuffered, so the bits match."""
    deg = np.zeros(n, dtype=np.float32)
    np.add.at(deg, idx, w.astype(np.float32, copy=False))
    return deg


def add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes):
    """PyG 1.7 semantics (see module docstring). edge_index: int64 [2,E]; edge_weight: f32 [E] or None."""
    edge_index = np.asarray(edge_index, dtype=np.int64)
    row, col = edge_index[0], edge_index[1]
    mask = row != col
    loops = np.arange(num_nodes, dtype=np.int64)
    new_index = np.concatenate([edge_index[:, mask], np.stack([loops, loops])], axis=1)
    if edge_weight is None:
        return new_index, None
    edge_weight = np.asarray(edge_weight, dtype=np.float32)
    loop_w = np.full(num_nodes, fill_value, dtype=np.float32)
    inv = ~mask
    if inv.any():
        # later duplicates overwrite earlier ones, as torch index assignment does
        loop_w[row[inv]] = edge_weight[inv]
    return new_index, np.concatenate([edge_weight[mask], loop_w])


def get_rw_adj(edge_index, edge_weight=None, norm_dim=1, fill_value=0.0, num_nodes=None):
    """`src/utils.py:215-233`. Returns (edge_index [2,E'], edge_weight f32 [E'])."""
    edge_index = np.asarray(edge_index, dtype=np.int64)
    if num_nodes is None:
        num_nodes = int(edge_index.max()) + 1
    if edge_weight is None:
        edge_weight = np.ones(edge_index.shape[1], dtype=np.float32)
    edge_weight = np.asarray(edge_weight, dtype=np.float32)
    if not fill_value == 0:
        edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes)
    row, col = edge_index[0], edge_index[1]
    indices = row if norm_dim == 0 else col
    deg = _seq_scatter_add_f32(edge_weight, indices, num_nodes)
    with np.errstate(divide='ignore'):
        deg_inv = (np.float32(1.0) / deg).astype(np.float32)      # deg.pow_(-1)
    return edge_index, (edge_weight * deg_inv[indices]).astype(np.float32)


def gcn_norm_fill_val(edge_index, edge_weight=None, fill_value=0.0, num_nodes=None):
    """`src/utils.py:177-194` (== PyG gcn_norm with a custom fill value)."""
    edge_index = np.asarray(edge_index, dtype=np.int64)
    if num_nodes is None:
        num_nodes = int(edge_index.max()) + 1
    if edge_weight is None:
        edge_weight = np.ones(edge_index.shape[1], dtype=np.float32)
    edge_weight = np.asarray(edge_weight, dtype=np.float32)
    if not int(fill_value) == 0:
        edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes)
    row, col = edge_index[0], edge_index[1]
    deg = _seq_scatter_add_f32(edge_weight, col, num_nodes)
    with np.errstate(divide='ignore'):
        dis = (np.float32(1.0) / np.sqrt(deg, dtype=np.float32)).astype(np.float32)   # deg.pow_(-0.5)
    dis[np.isinf(dis)] = 0
    return edge_index, ((dis[row] * edge_weight).astype(np.float32) * dis[col]).astype(np.float32)


def reset_graph_data(edge_index, edge_attr, num_nodes, data_norm='rw', self_loop_weight=1.0):
    """`src/base_classes.py:70-90`: what ODEblock pushes onto the ODE function."""
    if data_norm == 'rw':
        ei, ew = get_rw_adj(edge_index, edge_attr, norm_dim=1, fill_value=self_loop_weight, num_nodes=num_nodes)
    else:
        ei, ew = gcn_norm_fill_val(edge_index, edge_attr, fill_value=self_loop_weight, num_nodes=num_nodes)
    if self_loop_weight > 0:
        ei, ew = add_remaining_self_loops(ei, ew, self_loop_weight, num_nodes)
    return ei, ew


def coo_to_csr(edge_index, num_nodes):
    """Ordering contract of the plan builder: stable sort of the edges by (row, col).
    Returns rowptr int32 [N+1], colidx int32 [E], eperm int32 [E] (CSR slot -> original edge id)."""
    edge_index = np.asarray(edge_index, dtype=np.int64)
    row, col = edge_index[0], edge_index[1]
    key = row * np.int64(num_nodes) + col
    eperm = np.argsort(key, kind='stable')
    counts = np.bincount(row, minlength=num_nodes)
    rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(counts, out=rowptr[1:])
    return rowptr.astype(np.int32), col[eperm].astype(np.int32), eperm.astype(np.int32)


def csr_to_csc(rowptr, colidx, num_nodes):
    """Transpose ordering: stable sort of the CSR slots by (col, row).
    Returns colptr int32 [N+1], rowidx int32 [E], tperm int32 [E] (CSC slot -> CSR slot)."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    colidx = np.asarray(colidx, dtype=np.int64)
    rows = np.repeat(np.arange(num_nodes, dtype=np.int64), np.diff(rowptr))
    key = colidx * np.int64(num_nodes) + rows
    tperm = np.argsort(key, kind='stable')
    counts = np.bincount(colidx, minlength=num_nodes)
    colptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(counts, out=colptr[1:])
    return colptr.astype(np.int32), rows[tperm].astype(np.int32), tperm.astype(np.int32)


def to_dense_adj(edge_index, edge_weight, num_nodes, dtype=np.float64):
    """`src/utils.py:100-112`: duplicates add (sparse_coo_tensor(...).to_dense())."""
    a = np.zeros((num_nodes, num_nodes), dtype=dtype)
    np.add.at(a, (edge_index[0], edge_index[1]), np.asarray(edge_weight, dtype=dtype))
    return a


# ----------------------------------------------------------------------------------------------
# Synthetic graphs of the BASELINE.json shapes.  Host-side, numpy, seeded.
# ----------------------------------------------------------------------------------------------
def shaped_graph(num_nodes, num_edges, seed):
    """Exactly `num_nodes` nodes and `num_edges` directed edges (num_edges even): num_edges/2
    distinct unordered pairs u != v, u uniform, v ~ (rank+1)^-0.5 under a fixed random
    permutation, symmetrised.  Returns int64 [2, num_edges]."""
    assert num_edges % 2 == 0
    rng = np.random.default_rng(seed)
    half = num_edges // 2
    perm = rng.permutation(num_nodes)
    # inverse-CDF sampling of p(r) ~ (r+1)^-0.5 : r = floor(N * U^2)
    got = np.empty(0, dtype=np.int64)
    while got.size < half:
        need = int((half - got.size) * 1.2) + 16
        u = rng.integers(0, num_nodes, size=need, dtype=np.int64)
        v = perm[np.minimum((num_nodes * rng.random(need) ** 2).astype(np.int64), num_nodes - 1)]
        keep = u != v
        lo = np.minimum(u[keep], v[keep])
        hi = np.maximum(u[keep], v[keep])
        key = np.concatenate([got, lo * np.int64(num_nodes) + hi])
        _, first = np.unique(key, return_index=True)
        got = key[np.sort(first)]
    got = got[:half]
    lo, hi = got // num_nodes, got % num_nodes
    return np.stack([np.concatenate([lo, hi]), np.concatenate([hi, lo])]).astype(np.int64)


def rmat_graph(scale, edge_factor, seed, a=0.57, b=0.19, c=0.19):
    """Graph500 R-MAT, vertex ids permuted, duplicates and self loops kept. int64 [2, E]."""
    rng = np.random.default_rng(seed)
    n = 1 << scale
    e = n * edge_factor
    row = np.zeros(e, dtype=np.int64)
    col = np.zeros(e, dtype=np.int64)
    ab, c_norm, a_norm = a + b, c / (1.0 - (a + b)), a / (a + b)
    for _ in range(scale):
        i_bit = rng.random(e) > ab
        j_bit = rng.random(e) > np.where(i_bit, c_norm, a_norm)
        row = (row << 1) | i_bit
        col = (col << 1) | j_bit
    perm = rng.permutation(n)
    return np.stack([perm[row], perm[col]]).astype(np.int64)
