"""torch_geometric.utils.convert.to_scipy_sparse_matrix as documented by PyG (edge_index [2,E], edge_attr [E] ->
scipy COO matrix with entries A[row, col] = attr, duplicates kept) - used by src/CGNN.py:283.
TEST INFRASTRUCTURE ONLY (oracle/make_golden.py)."""
import scipy.sparse


def to_scipy_sparse_matrix(edge_index, edge_attr=None, num_nodes=None):
    row, col = edge_index.cpu().numpy()
    n = int(num_nodes) if num_nodes is not None else int(max(row.max(), col.max())) + 1
    val = edge_attr.detach().cpu().numpy() if edge_attr is not None else None
    if val is None:
        import numpy as np
        val = np.ones(row.size, dtype='float32')
    return scipy.sparse.coo_matrix((val, (row, col)), shape=(n, n))
