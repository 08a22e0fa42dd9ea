"""GPU: SpGraphTransAttentionLayer / ODEFuncTransformerAtt (projection GEMM + fused score/softmax/
head-mean kernel + fused SpMM) vs reference-generated goldens, reference known answers and the fp64
oracle."""
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

BASE_OPT = {'leaky_relu_slope': 0.2, 'heads': 2, 'attention_dim': 8, 'attention_norm_idx': 0, 'add_source': True,
            'hidden_dim': 8, 'block': 'attention', 'function': 'transformer', 'max_nfe': 1000,
            'no_alpha_sigmoid': False, 'reweight_attention': False, 'beltrami': False,
            'attention_type': 'scaled_dot', 'square_plus': False, 'mix_features': False, 'multi_modal': False}


def set_layer(layer, z):
    with torch.no_grad():
        layer.Q.weight.copy_(dev(z['wq']))
        layer.Q.bias.copy_(dev(z['bq']))
        layer.K.weight.copy_(dev(z['wk']))
        layer.K.bias.copy_(dev(z['bk']))


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_transformer_*.npz'))))
def test_golden_attention_and_f(path):
    """forward vs the reference's own outputs.  scaled_dot goldens hold the fork-literal E x E score
    (SURVEY.md F4), so for that type the comparison target is the per-edge oracle instead."""
    import grand_b200
    z = np.load(path)
    att_type = str(z['attention_type'])
    n, d = z['x'].shape
    opt = dict(BASE_OPT, attention_type=att_type, attention_norm_idx=int(z['attention_norm_idx']), hidden_dim=d)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    set_layer(func.multihead_att_layer, z)
    with torch.no_grad():
        func.alpha_train.fill_(float(z['alpha_train']))
        func.beta_train.fill_(float(z['beta_train']))
        if att_type == 'exp_kernel':
            func.multihead_att_layer.output_var.fill_(float(z['output_var']))
            func.multihead_att_layer.lengthscale.fill_(float(z['lengthscale']))
    ei = dev(z['edge_index'])[None]
    func.edge_index, func.x0 = ei, dev(z['x0'])[None]
    x = dev(z['x'])[None]
    att, (v, prods) = func.multihead_att_layer(x, ei)
    out = func(torch.tensor(0.0, device='cuda'), x)
    assert att.shape == (1, ei.shape[2], 2) and out.shape == x.shape
    if att_type == 'scaled_dot':
        T = lambda k: torch.tensor(z[k], dtype=torch.float64)
        kw = dict(attention_norm_idx=int(z['attention_norm_idx']))
        ref_att, ref_prods = funcs.attention_layer(T('x'), torch.tensor(z['edge_index']), T('wq'), T('bq'), T('wk'),
                                                   T('bk'), 2, **kw)
        ref_f = funcs.transformer_f(T('x'), torch.tensor(z['edge_index']), T('wq'), T('bq'), T('wk'), T('bk'), 2,
                                    torch.tensor(float(z['alpha_train']), dtype=torch.float64),
                                    torch.tensor(float(z['beta_train']), dtype=torch.float64), T('x0'), **kw)
    else:
        ref_att, ref_prods, ref_f = z['attention'], z['prods'], z['f']
    assert_close(prods[0], ref_prods, F_RTOL, 'prods')
    assert_close(att[0], ref_att, F_RTOL, 'attention')
    assert_close(out[0], ref_f, F_RTOL, 'f')


def test_k3_known_answer_exact_half():
    """test/test_transformer_attention.py:98-106: attention == 0.5 EXACTLY (torch.eq)"""
    import grand_b200
    z = np.load(os.path.join(GOLD, 'ref_k3.npz'))
    opt = dict(BASE_OPT, heads=2, attention_dim=32, hidden_dim=6)
    layer = grand_b200.SpGraphTransAttentionLayer(2, 2, opt, torch.device('cuda'), concat=True).cuda()
    assert torch.all(torch.isclose(layer.Q.weight, layer.K.weight))         # reference init: constant 1e-5
    set_layer(layer, z)
    att, _ = layer(dev(z['x'])[None], dev(z['edge_index'])[None])
    assert torch.all(torch.eq(att, 0.5 * torch.ones((1, 6, 2), device='cuda')))


def test_softmax_invariants_toy_and_random():
    """test/test_transformer_attention.py:57-76: shape [B,E,H]; groups sum to 1; 0 < a <= 1"""
    import grand_b200
    for norm_idx in (0, 1):
        opt = dict(BASE_OPT, heads=2, attention_dim=32, hidden_dim=6, attention_norm_idx=norm_idx)
        edge = torch.tensor([[[0, 2, 2, 1], [1, 0, 1, 2]]], device='cuda')
        x = torch.tensor([[[1., 2.], [3., 2.], [4., 5.]]], device='cuda')
        layer = grand_b200.SpGraphTransAttentionLayer(2, 2, opt, torch.device('cuda')).cuda()
        att, _ = layer(x, edge)
        assert att.shape == (1, 4, 2)
        sums = torch.zeros(3, 2, device='cuda').index_add_(0, edge[0, norm_idx], att[0])
        present = torch.bincount(edge[0, norm_idx], minlength=3) > 0
        assert torch.allclose(sums[present], torch.ones_like(sums[present]), atol=1e-3)
        assert bool((att > 0).all()) and bool((att <= 1).all())


def oracle_transformer(ei, x, x0, gout, wq, bq, wk, bk, heads, alpha, beta, norm_idx):
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    leaves = [t64(v).requires_grad_(True) for v in (x, wq, bq, wk, bk)]
    a, b = t64(alpha).requires_grad_(True), t64(beta).requires_grad_(True)
    f = funcs.transformer_f(leaves[0], torch.tensor(ei), leaves[1], leaves[2], leaves[3], leaves[4], heads, a, b,
                            t64(x0), attention_norm_idx=norm_idx)
    grads = torch.autograd.grad(f, leaves + [a, b], t64(gout))
    return f.detach(), grads


@pytest.mark.parametrize('n,e,d,heads,adim', [(60, 400, 8, 2, 8), (2000, 12000, 64, 4, 64), (1500, 9000, 128, 4, 64),
                                               (900, 6000, 96, 4, 64), (500, 3000, 80, 8, 128), (400, 2500, 16, 4, 8),
                                               (300, 2000, 6, 2, 6), (700, 4000, 162, 2, 32),
                                               (60, 400, 16, 8, 8), (60, 400, 4, 4, 128)])
@pytest.mark.parametrize('norm_idx', [0, 1])
def test_scaled_dot_forward_and_all_gradients(n, e, d, heads, adim, norm_idx):
    import grand_b200
    ei, _ = og.reset_graph_data(og.shaped_graph(n, e, n + 1), None, n, 'rw', 1.0)
    rng = np.random.default_rng(d + heads)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    wq, wk = (rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d) for _ in range(2))
    bq, bk = (0.1 * rng.standard_normal(adim).astype(np.float32) for _ in range(2))
    opt = dict(BASE_OPT, heads=heads, attention_dim=adim, hidden_dim=d, attention_norm_idx=norm_idx)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    set_layer(func.multihead_att_layer, dict(wq=wq, bq=bq, wk=wk, bk=bk))
    with torch.no_grad():
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = func(0.0, xg)
    L = func.multihead_att_layer
    params = [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]
    grads = torch.autograd.grad(out, params, dev(gout))
    assert L.V.weight.grad is None                                           # V is dead code, as in the reference
    ref, rg = oracle_transformer(ei, x, x0, gout, wq, bq, wk, bk, heads, 0.3, 0.1, norm_idx)
    assert_close(out, ref, F_RTOL, 'f')
    bias_scale = max(float(rg[2].abs().max()), float(rg[4].abs().max()))      # one bias gradient is exactly 0 in maths
    for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
        assert_close(g, r, G_RTOL, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


def test_hub_segments_long_path_both_norms():
    import grand_b200
    n, d, heads, adim = 500, 32, 4, 32
    ei = rand_graph(n, 3000, 8, hub=(5, 7000))
    rng = np.random.default_rng(2)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    wq, wk = (rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d) for _ in range(2))
    bq, bk = (0.1 * rng.standard_normal(adim).astype(np.float32) for _ in range(2))
    for norm_idx in (0, 1):
        opt = dict(BASE_OPT, heads=heads, attention_dim=adim, hidden_dim=d, attention_norm_idx=norm_idx)
        func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
        set_layer(func.multihead_att_layer, dict(wq=wq, bq=bq, wk=wk, bk=bk))
        func.edge_index, func.x0 = dev(ei), dev(x0)
        xg = dev(x).requires_grad_(True)
        out = func(0.0, xg)
        L = func.multihead_att_layer
        grads = torch.autograd.grad(out, [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train,
                                          func.beta_train], dev(gout))
        ref, rg = oracle_transformer(ei, x, x0, gout, wq, bq, wk, bk, heads, 0.0, 0.0, norm_idx)
        assert_close(out, ref, F_RTOL, 'f')
        bias_scale = max(float(rg[2].abs().max()), float(rg[4].abs().max()))
        for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
            assert_close(g, r, G_RTOL, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


@pytest.mark.parametrize('impl', ['3', '2', '1'])
@pytest.mark.parametrize('e', [0, 40])
def test_isolated_nodes_and_empty_graph_through_the_row_kernels(impl, e, monkeypatch):
    """rows without any edge (no self loops: the function is also called on raw edge lists, test/test_*.py) and
    columns nobody points to; e = 0: f = -alpha x + beta x0 and every attention gradient is zero"""
    import grand_b200
    monkeypatch.setenv('GRANDB200_NL_IMPL', impl)
    n, d, heads, adim = 97, 32, 4, 32
    rng = np.random.default_rng(e)
    ei = np.stack([rng.integers(0, n // 3, e), rng.integers(n // 2, n, e)]).astype(np.int64)   # two thirds of the rows empty
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    wq, wk = (rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d) for _ in range(2))
    bq, bk = (0.1 * rng.standard_normal(adim).astype(np.float32) for _ in range(2))
    opt = dict(BASE_OPT, heads=heads, attention_dim=adim, hidden_dim=d)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    set_layer(func.multihead_att_layer, dict(wq=wq, bq=bq, wk=wk, bk=bk))
    with torch.no_grad():
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = func(0.0, xg)
    L = func.multihead_att_layer
    grads = torch.autograd.grad(out, [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train],
                                dev(gout))
    ref, rg = oracle_transformer(ei, x, x0, gout, wq, bq, wk, bk, heads, 0.3, 0.1, 0)
    assert_close(out, ref, F_RTOL, 'f')
    scale = max(float(r.abs().max()) for r in rg[1:5]) if e else 1.0
    for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
        assert_close(g, r, G_RTOL, name, floor=scale if name in ('gwq', 'gbq', 'gwk', 'gbk') else 0.0)
    f2, dx2, p2 = func.eval_vjp(0.0, dev(x), dev(gout), [L.Q.weight, func.alpha_train])      # the fused adjoint evaluation
    assert_close(f2, ref, F_RTOL, 'f (eval_vjp)')
    assert_close(dx2, rg[0], G_RTOL, 'dx (eval_vjp)')


def test_head_mean_identity_through_kernels():
    """test/test_transformer_attention.py:118-143: mean_h(A_h x) == (mean_h A_h) x"""
    import grand_b200
    n, d, h = 300, 16, 4
    ei = rand_graph(n, 2000, 6)
    att = np.random.default_rng(0).random((ei.shape[1], h)).astype(np.float32)
    x = np.random.default_rng(1).standard_normal((n, d)).astype(np.float32)
    opt = {'hidden_dim': d, 'block': 'attention', 'max_nfe': 100, 'no_alpha_sigmoid': True, 'add_source': False}
    f = grand_b200.LaplacianODEFunc(d, d, opt, torch.device('cuda')).cuda()
    with torch.no_grad():
        f.alpha_train.fill_(1.0)
    f.edge_index, f.attention_weights = dev(ei), dev(att)
    got = f(0.0, dev(x)) + dev(x)                                            # alpha = 1: f = Ax - x
    t64 = lambda a: torch.tensor(a, dtype=torch.float64)
    per_head = torch.stack([funcs.spmm(torch.tensor(ei), t64(att[:, i]), t64(x)) for i in range(h)]).mean(0)
    assert_close(got, per_head, 1e-5, 'head mean')


def test_unsupported_options_raise():
    import grand_b200
    x = torch.randn(5, 8, device='cuda')
    ei = torch.tensor([[0, 1, 2], [1, 2, 0]], device='cuda')
    for key, val in (('mix_features', True),):                # square_plus: supported since round 2 (test_gpu_variants.py)
        func = grand_b200.ODEFuncTransformerAtt(8, 8, dict(BASE_OPT, **{key: val}), torch.device('cuda')).cuda()
        func.edge_index, func.x0 = ei, x
        with pytest.raises(NotImplementedError):
            func(0.0, x)
    with pytest.raises(NotImplementedError, match='attention_dim'):           # wider than a warp's 32 x 4 floats
        grand_b200.ODEFuncTransformerAtt(128, 128, dict(BASE_OPT, heads=8, attention_dim=256, hidden_dim=128), torch.device('cuda'))


@pytest.mark.parametrize('d,heads,adim', [(64, 4, 64), (128, 4, 64), (96, 4, 64), (80, 8, 128), (32, 1, 16)])
def test_fused_path_matches_composite_path_and_is_deterministic(d, heads, adim, monkeypatch):
    """the warp-per-row fused kernels and the general composite kernels are two implementations of the
    same evaluation: compare them with each other (tight) in addition to the oracle comparisons above"""
    import grand_b200
    from grand_b200 import function_transformer_attention as fta
    n, e = 3000, 30000
    ei = rand_graph(n, e, 12, hub=(7, 1200))                                # includes > 256-edge segments
    rng = np.random.default_rng(5)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    opt = dict(BASE_OPT, heads=heads, attention_dim=adim, hidden_dim=d)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    g = torch.Generator().manual_seed(1)
    L = func.multihead_att_layer
    with torch.no_grad():
        for lin in (L.Q, L.K):
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / d ** 0.5).cuda())
        func.alpha_train.fill_(0.2)
        func.beta_train.fill_(-0.1)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]

    def run():
        xg = dev(x).requires_grad_(True)
        out = func(0.0, xg)
        return [out.detach()] + list(torch.autograd.grad(out, [xg] + params, dev(gout)))

    plan = grand_b200.get_plan(func.edge_index, n)
    assert fta.fused_transformer_supported(plan, dev(x), heads, adim, 0, 0, None)
    assert fta.nl_supported(plan, dev(x), heads, adim, 0, 0, None)
    names = ('f', 'gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')

    def compare(a_list, b_list, tol):
        bias_scale = max(float(b_list[3].abs().max()), float(b_list[5].abs().max()))
        for a, b, name in zip(a_list, b_list, names):
            assert_close(a, b, tol, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)

    # (1) single-gather kernels (nl_kernels.cuh): deterministic, no-grad variant identical, fused eval+VJP identical
    nl_a, nl_b = run(), run()
    for a, b in zip(nl_a, nl_b):
        assert torch.equal(a, b)
    with torch.no_grad():
        assert torch.equal(func(0.0, dev(x)), nl_a[0])
    f2, vx, vps = func.eval_vjp(0.0, dev(x), dev(gout), params)
    assert_close(f2, nl_a[0], 1e-6, 'f from the adjoint row kernel vs the forward kernel')
    for a, b in zip([vx] + vps, nl_a[1:]):
        assert torch.equal(a, b)
    # (2) k-gather warp-per-row kernels (fused_att.cu)
    monkeypatch.setattr(fta, 'nl_supported', lambda *a, **k: False)
    assert func.eval_vjp(0.0, dev(x), dev(gout), params) is None
    fused_a, fused_b = run(), run()
    for a, b in zip(fused_a, fused_b):
        assert torch.equal(a, b)                                             # bit-reproducible
    with torch.no_grad():
        nograd = func(0.0, dev(x))                                           # store == 0 variant of the kernel
    assert torch.equal(nograd, fused_a[0])
    # (3) general composite kernels
    monkeypatch.setattr(fta, 'fused_transformer_supported', lambda *a, **k: False)
    comp = run()
    compare(fused_a, comp, 2e-5)
    compare(nl_a, comp, 2e-5)
