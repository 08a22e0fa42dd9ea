"""GPU: LaplacianODEFunc (fused CSR SpMM + epilogue, transpose SpMM VJP) vs the reference-generated
golden vectors and vs the fp64 oracle.  Tolerances: per-evaluation outputs rtol 1e-5, gradients 1e-4
(BASELINE.json north_star)."""
import glob
import os
import numpy as np
import pytest
import torch

from oracle import funcs, graph as og
from helpers import rand_graph, assert_close, dev

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden')
F_RTOL, G_RTOL = 1e-5, 1e-4


def make_func(d, block='constant', add_source=True, no_alpha_sigmoid=False, alpha=0.3, beta=-0.2, max_nfe=1000):
    import grand_b200
    opt = {'hidden_dim': d, 'block': block, 'max_nfe': max_nfe, 'no_alpha_sigmoid': no_alpha_sigmoid,
           'add_source': add_source, 'multi_modal': False}
    f = grand_b200.LaplacianODEFunc(d, d, opt, torch.device('cuda')).cuda()
    with torch.no_grad():
        f.alpha_train.fill_(alpha)
        f.beta_train.fill_(beta)
    return f


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_laplacian_*.npz'))))
@pytest.mark.parametrize('batched', [True, False])
def test_golden_reference_vectors(path, batched):
    z = np.load(path)
    n, d = z['x'].shape
    f = make_func(d, block=str(z['block']), no_alpha_sigmoid=bool(z['no_alpha_sigmoid']),
                  alpha=float(z['alpha_train']), beta=float(z['beta_train']))
    wrap = (lambda t: t[None]) if batched else (lambda t: t)      # fork layout [1,...] and upstream layout
    f.edge_index = wrap(dev(z['edge_index']))
    f.edge_weight = wrap(dev(z['edge_weight']))
    f.attention_weights = wrap(dev(z['attention']))
    f.x0 = wrap(dev(z['x0']))
    x = wrap(dev(z['x'])).requires_grad_(True)
    out = f(torch.tensor(0.5, device='cuda'), x)
    assert out.shape == x.shape and f.nfe == int(z['nfe'])
    gx, ga, gb = torch.autograd.grad(out, (x, f.alpha_train, f.beta_train), wrap(dev(z['gout'])))
    assert_close(out.reshape(n, d), z['f'], F_RTOL, 'f')
    assert_close(gx.reshape(n, d), z['gx'], G_RTOL, 'gx')
    assert_close(ga, z['galpha'], G_RTOL, 'galpha')
    assert_close(gb, z['gbeta'], G_RTOL, 'gbeta')


def oracle_eval(ei, w, x, x0, gout, alpha, beta, add_source, no_sig):
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    xo = t64(x).requires_grad_(True)
    a = t64(alpha).requires_grad_(True)
    b = t64(beta).requires_grad_(True)
    f = funcs.laplacian_f(xo, torch.tensor(ei), t64(w), a, b, t64(x0), add_source, no_sig)
    gs = torch.autograd.grad(f, (xo, a, b), t64(gout), allow_unused=True)
    return f.detach(), gs


@pytest.mark.parametrize('n,e,d', [(50, 300, 6), (2708, 10556, 64), (1500, 9000, 128), (1200, 8000, 96),
                                   (900, 5000, 162), (700, 4000, 7), (400, 3000, 520)])
@pytest.mark.parametrize('add_source,no_sig', [(True, False), (False, True)])
def test_random_graphs_vs_fp64_oracle(n, e, d, add_source, no_sig):
    ei, w = og.reset_graph_data(og.shaped_graph(n, e, n), None, n, 'rw', 1.0)
    rng = np.random.default_rng(d)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    f = make_func(d, add_source=add_source, no_alpha_sigmoid=no_sig, alpha=0.3, beta=0.1)
    f.edge_index, f.edge_weight, f.x0 = dev(ei), dev(w), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = f(0.0, xg)
    grads = torch.autograd.grad(out, (xg, f.alpha_train, f.beta_train), dev(gout), allow_unused=True)
    ref, rg = oracle_eval(ei, w, x, x0, gout, 0.3, 0.1, add_source, no_sig)
    assert_close(out, ref, F_RTOL, 'f')
    assert_close(grads[0], rg[0], G_RTOL, 'gx')
    assert_close(grads[1], rg[1], G_RTOL, 'galpha')
    if add_source:
        assert_close(grads[2], rg[2], G_RTOL, 'gbeta')
    else:
        assert grads[2] is None


def test_hub_rows_and_duplicates_long_path():
    """rows / columns longer than the plan's long threshold go through the CTA-per-row kernels"""
    n, d = 600, 64
    ei = rand_graph(n, 4000, 3, hub=(17, 9000))
    w = np.random.default_rng(0).random(ei.shape[1]).astype(np.float32) / 50
    rng = np.random.default_rng(1)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    f = make_func(d, alpha=-0.5, beta=0.4)
    f.edge_index, f.edge_weight, f.x0 = dev(ei), dev(w), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = f(0.0, xg)
    import grand_b200
    plan = grand_b200.get_plan(f.edge_index, n)
    assert plan.long_rows.numel() >= 1 and plan.long_cols.numel() >= 1
    grads = torch.autograd.grad(out, (xg, f.alpha_train, f.beta_train), dev(gout))
    ref, rg = oracle_eval(ei, w, x, x0, gout, -0.5, 0.4, True, False)
    assert_close(out, ref, F_RTOL, 'f')
    for g, r, name in zip(grads, rg, ('gx', 'galpha', 'gbeta')):
        assert_close(g, r, G_RTOL, name)


@pytest.mark.parametrize('n,e,d,hub', [(3000, 24000, 128, None), (2500, 20000, 64, (17, 5000)), (1500, 9000, 96, (3, 2500)),
                                       (1200, 7000, 40, None), (900, 6000, 16, (5, 3000))])
def test_dynamically_scheduled_spmm_vs_oracle_and_vs_the_static_kernels(n, e, d, hub, monkeypatch):
    """csrc/spmm2_kernels.cuh (the large-graph path: persistent CTAs, TMA-staged gathers, CTA-wide rows > 2048 edges)
    forced onto small graphs: f and all gradients vs the fp64 oracle, run-to-run identical, and within rounding of the
    statically tiled kernels"""
    import grand_b200
    ei = rand_graph(n, e, n + d, hub=hub)
    ei = np.concatenate([ei, np.array([[7, 7, 7], [9, 9, 9]])], axis=1)           # duplicate edges add
    w = (np.random.default_rng(2).random(ei.shape[1]).astype(np.float32) + 0.1) / 40
    rng = np.random.default_rng(d)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    x[11] = 0                                                                     # and an isolated-looking row
    res = {}
    for impl in ('dynamic', 'dynamic-again', 'static'):
        monkeypatch.setenv('GRANDB200_SPMM2_MIN_ROWS', '1')
        monkeypatch.setenv('GRANDB200_SPMM_IMPL', '1' if impl == 'static' else '2')
        f = make_func(d, alpha=0.3, beta=0.1)
        f.edge_index, f.edge_weight, f.x0 = dev(ei), dev(w), dev(x0)
        xg = dev(x).requires_grad_(True)
        with grand_b200._lib.Profiler() as prof:
            out = f(0.0, xg)
            grads = torch.autograd.grad(out, (xg, f.alpha_train, f.beta_train), dev(gout))
            kernels = prof.results()
        assert ('spmm_fwd_long' in kernels) == (impl == 'static' and hub is not None), sorted(kernels)
        res[impl] = (out.detach(),) + tuple(g.detach() for g in grads)
    ref, rg = oracle_eval(ei, w, x, x0, gout, 0.3, 0.1, True, False)
    for got, want, name in zip(res['dynamic'], (ref,) + tuple(rg), ('f', 'gx', 'galpha', 'gbeta')):
        assert_close(got, want, F_RTOL if name == 'f' else G_RTOL, name)
    for a_, b_ in zip(res['dynamic'], res['dynamic-again']):
        assert torch.equal(a_, b_)
    for a_, b_, name in zip(res['dynamic'], res['static'], ('f', 'gx', 'galpha', 'gbeta')):
        assert_close(a_, b_, 2e-6 if name in ('f', 'gx') else 1e-5, name + ' vs the static kernels')


def test_isolated_nodes_and_empty_graph():
    n, d = 40, 8
    f = make_func(d, alpha=0.2, beta=0.0, add_source=False)
    x = torch.randn(n, d, device='cuda')
    f.edge_index = torch.zeros(2, 0, dtype=torch.long, device='cuda')
    f.edge_weight = torch.zeros(0, device='cuda')
    out = f(0.0, x)
    assert_close(out, -torch.sigmoid(torch.tensor(0.2)) * x.cpu(), 1e-6, 'empty graph: f = -alpha x')


def test_deterministic_and_nfe_guard():
    import grand_b200
    n, d = 3000, 64
    ei, w = og.reset_graph_data(og.shaped_graph(n, 20000, 9), None, n, 'rw', 1.0)
    f = make_func(d, max_nfe=3)
    f.edge_index, f.edge_weight = dev(ei), dev(w)
    x = torch.randn(n, d, device='cuda', requires_grad=True)
    f.x0 = x.detach().clone()
    g = torch.randn(n, d, device='cuda')
    runs = []
    for _ in range(2):
        out = f(0.0, x)
        runs.append((out.detach().clone(),) + tuple(t.clone() for t in torch.autograd.grad(out, (x, f.alpha_train, f.beta_train), g)))
    for a, b in zip(*runs):
        assert torch.equal(a, b)                                            # no atomics: bit-reproducible
    f.nfe = 4
    with pytest.raises(grand_b200.MaxNFEException):
        f(0.0, x)
    assert f.nfe == 4                                                        # raised before the increment


def test_attention_block_gradient_reaches_attention_weights():
    """non-adjoint training: attention_weights set by AttODEblock stay in the autograd graph"""
    n, d, h = 200, 16, 4
    ei = rand_graph(n, 1500, 4)
    e = ei.shape[1]
    f = make_func(d, block='attention', alpha=0.1, beta=0.3)
    rng = np.random.default_rng(3)
    att = rng.random((e, h)).astype(np.float32)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    f.edge_index, f.x0 = dev(ei)[None], dev(x0)[None]
    att_g = dev(att)[None].requires_grad_(True)
    f.attention_weights = att_g
    out = f(0.0, dev(x)[None])
    gatt, = torch.autograd.grad(out, att_g, dev(gout)[None])
    t64 = lambda a: torch.tensor(a, dtype=torch.float64)
    att_o = t64(att).requires_grad_(True)
    ref = funcs.laplacian_f(t64(x), torch.tensor(ei), att_o.mean(1), t64(0.1), t64(0.3), t64(x0))
    ratt, = torch.autograd.grad(ref, att_o, t64(gout))
    assert_close(out[0], ref, F_RTOL, 'f')
    assert_close(gatt[0], ratt, G_RTOL, 'g_attention')
