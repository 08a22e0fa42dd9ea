"""GPU: plan construction and graph normalisation are BIT-EXACT vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import graph as og
from helpers import rand_graph, dev

pytestmark = pytest.mark.gpu


def check_plan(ei, n, long_threshold=2048):
    import grand_b200
    plan = grand_b200.GraphPlan(dev(ei), n, long_threshold=long_threshold)
    rowptr, colidx, eperm = og.coo_to_csr(ei, n)
    colptr, rowidx, tperm = og.csr_to_csc(rowptr, colidx, n)
    for name, want in (('rowptr', rowptr), ('colidx', colidx), ('eperm', eperm), ('colptr', colptr),
                       ('rowidx', rowidx), ('tperm', tperm)):
        got = getattr(plan, name).cpu().numpy()
        assert got.dtype == np.int32 and np.array_equal(got, want), name
    deg_r, deg_c = np.diff(rowptr), np.diff(colptr)
    assert np.array_equal(plan.long_rows.cpu().numpy(), np.nonzero(deg_r > long_threshold)[0])
    assert np.array_equal(plan.long_cols.cpu().numpy(), np.nonzero(deg_c > long_threshold)[0])
    return plan


@pytest.mark.parametrize('n,e,seed', [(3, 4, 0), (17, 120, 1), (1000, 20000, 2), (2708, 10556, 3), (50000, 400000, 4)])
def test_plan_matches_oracle(n, e, seed):
    check_plan(rand_graph(n, e, seed), n)


def test_plan_edge_cases():
    check_plan(np.array([[0, 2, 2, 1], [1, 0, 1, 2]]), 3)                 # reference toy graph
    check_plan(np.array([[5], [5]]), 9)                                    # single self loop, empty rows around it
    check_plan(np.stack([np.zeros(64, dtype=np.int64), np.zeros(64, dtype=np.int64)]), 4)   # 64 duplicates
    check_plan(np.zeros((2, 0), dtype=np.int64), 5)                        # no edges
    check_plan(rand_graph(300, 2000, 7, hub=(11, 5000)), 300, long_threshold=512)    # hub row + hub column


def test_plan_large_ids():
    n = 3_000_000                                                           # key needs > 32 bits
    rng = np.random.default_rng(5)
    ei = rng.integers(0, n, size=(2, 200_000))
    check_plan(ei, n)


@pytest.mark.parametrize('weighted', [False, True])
@pytest.mark.parametrize('fill', [0.0, 0.3, 1.0, 3.2])
@pytest.mark.parametrize('norm_dim', [0, 1])
def test_rw_normalisation_bit_exact(weighted, fill, norm_dim):
    import grand_b200
    n = 500
    ei = rand_graph(n, 6000, 11, self_loops=False)
    rng = np.random.default_rng(12)
    w = rng.random(ei.shape[1]).astype(np.float32) if weighted else None
    want_ei, want_w = og.get_rw_adj(ei, w, norm_dim=norm_dim, fill_value=fill, num_nodes=n)
    got_ei, got_w = grand_b200.get_rw_adj(dev(ei), None if w is None else dev(w), norm_dim=norm_dim,
                                          fill_value=fill, num_nodes=n)
    assert np.array_equal(got_ei.cpu().numpy(), want_ei)
    assert np.array_equal(got_w.cpu().numpy().view(np.uint32), want_w.view(np.uint32))      # same bits


@pytest.mark.parametrize('weighted', [False, True])
def test_gcn_normalisation_bit_exact(weighted):
    import grand_b200
    n = 400
    ei = rand_graph(n, 5000, 21, self_loops=False)
    w = np.random.default_rng(2).random(ei.shape[1]).astype(np.float32) if weighted else None
    want_ei, want_w = og.gcn_norm_fill_val(ei, w, 1.0, n)
    got_ei, got_w = grand_b200.gcn_norm_fill_val(dev(ei), None if w is None else dev(w), fill_value=1.0, num_nodes=n)
    assert np.array_equal(got_ei.cpu().numpy(), want_ei)
    assert np.array_equal(got_w.cpu().numpy().view(np.uint32), want_w.view(np.uint32))


def test_reference_known_answers_toy():
    """test/test_utils.py:62-77 and test/test_function_laplacian_diffusion.py:68-85 through the CUDA path"""
    import grand_b200
    from sklearn.preprocessing import normalize
    edge = np.array([[0, 2, 2, 1], [1, 0, 1, 2]])
    base = og.to_dense_adj(edge, np.ones(4), 3)
    for sl in (0, 0.3, 1, 3.2):
        for nd in (0, 1):
            ei, w = grand_b200.get_rw_adj(dev(edge)[None], norm_dim=nd, fill_value=sl, num_nodes=3)   # fork layout
            assert ei.shape[0] == 1 and w.shape[0] == 1
            dense = og.to_dense_adj(ei[0].cpu().numpy(), w[0].cpu().numpy(), 3)
            assert np.allclose(normalize(base + np.identity(3) * sl, norm='l1', axis=0 if nd == 1 else 1), dense)


def test_block_graph_reset_matches_oracle():
    import grand_b200
    n = 300
    ei = og.shaped_graph(n, 2000, 5)
    want_ei, want_w = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    opt = {'hidden_dim': 8, 'block': 'constant', 'max_nfe': 100, 'no_alpha_sigmoid': False, 'add_source': True,
           'augment': False, 'adjoint': False, 'tol_scale': 1.0, 'data_norm': 'rw', 'self_loop_weight': 1.0,
           'method': 'rk4', 'step_size': 1.0, 'max_iters': 100}
    block = grand_b200.ConstantODEblock(grand_b200.LaplacianODEFunc, [], opt, torch.device('cuda'),
                                        t=torch.tensor([0., 1.])).cuda()
    data = grand_b200.GraphData()
    data.new_graph(dev(ei), n)
    block.reset_graph_data(data, torch.float32)
    assert np.array_equal(block.odefunc.edge_index.cpu().numpy(), want_ei)
    assert np.array_equal(block.odefunc.edge_weight.cpu().numpy().view(np.uint32), want_w.view(np.uint32))
    first = block.odefunc.edge_index
    block.reset_graph_data(data, torch.float32)
    assert block.odefunc.edge_index is first                                 # cached while the input is unchanged
