"""Known-answer tests of the reference for graph normalisation, run against oracle/graph.py
(test/test_utils.py:27-30,62-77; test/test_function_laplacian_diffusion.py:68-85)."""
import numpy as np
import pytest
from sklearn.preprocessing import normalize

from oracle import graph

EDGE = np.array([[0, 2, 2, 1], [1, 0, 1, 2]])          # test/test_utils.py:37
EDGE_SYM = np.array([[0, 1, 2, 1], [1, 0, 1, 2]])      # test/test_function_laplacian_diffusion.py:33


def dense(ei, w, n):
    return graph.to_dense_adj(ei, w, n)


@pytest.mark.parametrize('self_loop', [0, 0.3, 1, 3.2])
@pytest.mark.parametrize('norm_dim', [0, 1])
def test_get_rw_adj_matches_sklearn(self_loop, norm_dim):
    base = dense(EDGE, np.ones(4), 3)
    ei, ew = graph.get_rw_adj(EDGE, norm_dim=norm_dim, fill_value=self_loop, num_nodes=3)
    want = normalize(base + np.identity(3) * self_loop, norm='l1', axis=0 if norm_dim == 1 else 1)
    assert np.allclose(want, dense(ei, ew, 3))


def test_rw_and_sym_adjacency_toy():
    n = 3
    ei, ew = graph.get_rw_adj(EDGE_SYM, None, norm_dim=1, fill_value=1, num_nodes=n)
    aug = dense(EDGE_SYM, np.ones(4), n) + np.identity(n)
    assert np.allclose(normalize(aug, norm='l1', axis=0), dense(ei, ew, n))
    ei, ew = graph.gcn_norm_fill_val(EDGE_SYM, None, 1, n)
    sq = np.sqrt(aug.sum(axis=1))
    assert np.allclose(aug / sq[:, None] / sq[None, :], dense(ei, ew, n))


def test_add_remaining_self_loops_pyg_semantics():
    ei = np.array([[0, 1, 1, 2], [1, 1, 0, 0]])
    w = np.array([0.5, 7.0, 0.25, 2.0], dtype=np.float32)
    ni, nw = graph.add_remaining_self_loops(ei, w, 3.0, 4)
    assert ni.tolist() == [[0, 1, 2, 0, 1, 2, 3], [1, 0, 0, 0, 1, 2, 3]]
    assert nw.tolist() == [0.5, 0.25, 2.0, 3.0, 7.0, 3.0, 3.0]
    # second application is a no-op up to ordering (base_classes.py:83-85)
    ni2, nw2 = graph.add_remaining_self_loops(ni, nw, 3.0, 4)
    assert np.array_equal(ni2, ni) and np.array_equal(nw2, nw)


def test_reset_graph_data_columns_sum_to_one():
    ei = graph.shaped_graph(200, 1000, seed=3)
    assert ei.shape == (2, 1000) and (ei[0] != ei[1]).all()
    e2, w = graph.reset_graph_data(ei, None, 200, 'rw', 1.0)
    assert e2.shape[1] == 1200
    assert np.allclose(dense(e2, w, 200).sum(axis=0), 1.0, atol=1e-6)


def test_csr_csc_ordering_contract():
    rng = np.random.default_rng(0)
    n, e = 17, 120
    ei = rng.integers(0, n, size=(2, e))
    rowptr, colidx, eperm = graph.coo_to_csr(ei, n)
    rows = np.repeat(np.arange(n), np.diff(rowptr))
    assert np.array_equal(rows, ei[0][eperm]) and np.array_equal(colidx, ei[1][eperm])
    key = rows.astype(np.int64) * n + colidx
    assert (np.diff(key) >= 0).all()
    same = np.diff(key) == 0
    assert (np.diff(eperm)[same] > 0).all()              # stable among duplicates
    colptr, rowidx, tperm = graph.csr_to_csc(rowptr, colidx, n)
    assert np.array_equal(rowidx, rows[tperm])
    cols = np.repeat(np.arange(n), np.diff(colptr))
    assert np.array_equal(cols, colidx[tperm])
    assert sorted(tperm.tolist()) == list(range(e))


def test_rmat_shape():
    ei = graph.rmat_graph(8, 4, seed=1)
    assert ei.shape == (2, 1024) and ei.max() < 256
