"""GPU parity of the widened rows of SURVEY.md 8f: training paths of attention_type cosine_sim / pearson / exp_kernel
(incl. the beltrami split kernel), the GAT function, utils.softmax, MixedODEblock / HardAttODEblock, the two-phase
adjoint, the 2^16-node R-MAT parity gate and the GNN / GNNEarly mirrors.  First run on a B200 in round 2
(profiles/r02a_pytest_all.log: all green); part of the plain `pytest -m gpu` run."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import funcs, graph as og
from helpers import assert_close, dev, device

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden')
F_RTOL, G_RTOL = 1e-5, 1e-4

BASE_OPT = {'leaky_relu_slope': 0.2, 'heads': 2, 'attention_dim': 8, 'attention_norm_idx': 0, 'add_source': True,
            'hidden_dim': 8, 'block': 'attention', 'function': 'transformer', 'max_nfe': 1000,
            'no_alpha_sigmoid': False, 'reweight_attention': False, 'beltrami': False,
            'attention_type': 'scaled_dot', 'square_plus': False, 'mix_features': False, 'multi_modal': False}


def _func(opt, z):
    import grand_b200
    d = int(opt['hidden_dim'])
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, device()).to(device())
    L = func.multihead_att_layer
    with torch.no_grad():
        for lin, w, b in ((L.Q, 'wq', 'bq'), (L.K, 'wk', 'bk')):
            lin.weight.copy_(dev(z[w]))
            lin.bias.copy_(dev(z[b]))
        func.alpha_train.fill_(float(z['alpha_train']))
        func.beta_train.fill_(float(z['beta_train']))
    return func


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_transformer_cosine_sim_*.npz')) +
                                        glob.glob(os.path.join(GOLD, 'ref_transformer_pearson_*.npz'))))
def test_cosine_and_pearson_gradients_match_the_reference(path):
    """training path of attention_type cosine_sim / pearson (projection -> unit_heads -> scaled_dot edge kernels
    -> SpMM) against the gradients the reference itself produced for the same inputs"""
    z = np.load(path)
    n, d = z['x'].shape
    opt = dict(BASE_OPT, attention_type=str(z['attention_type']), attention_norm_idx=int(z['attention_norm_idx']),
               hidden_dim=d, heads=int(z['heads']), attention_dim=int(z['wq'].shape[0]))
    func = _func(opt, z)
    func.edge_index, func.x0 = dev(z['edge_index']), dev(z['x0'])
    xg = dev(z['x']).requires_grad_(True)
    out = func(0.0, xg)
    L = func.multihead_att_layer
    params = [xg, func.alpha_train, func.beta_train, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias]
    grads = torch.autograd.grad(out, params, dev(z['gout']))
    assert_close(out, z['f'], F_RTOL, 'f')
    for g, key in zip(grads, ('gx', 'galpha', 'gbeta', 'gwq', 'gbq', 'gwk', 'gbk')):
        assert_close(g, z[key], G_RTOL, key)
    with torch.no_grad():                                        # inference keeps the dedicated forward kernels
        assert_close(func(0.0, dev(z['x'])), z['f'], F_RTOL, 'f (no grad)')


@pytest.mark.parametrize('att_type', ['cosine_sim', 'pearson'])
def test_cosine_and_pearson_train_under_the_adjoint_block(att_type):
    """AttODEblock + odeint_adjoint with a non-scaled_dot function: eval_vjp declines, the solver falls back to
    autograd over forward(), which now has a backward for these types; compared with the fp64 oracle"""
    import grand_b200
    from oracle import integrate as oi
    n, e, d, heads, adim = 300, 1800, 16, 2, 8
    ei, _ = og.reset_graph_data(og.shaped_graph(n, e, 5), None, n, 'rw', 1.0)
    rng = np.random.default_rng(11)
    x, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(2))
    z = dict(wq=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bq=0.1 * rng.standard_normal(adim).astype(np.float32),
             wk=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bk=0.1 * rng.standard_normal(adim).astype(np.float32),
             alpha_train=0.3, beta_train=0.1)
    opt = dict(BASE_OPT, attention_type=att_type, hidden_dim=d, heads=heads, attention_dim=adim)
    func = _func(opt, z)
    func.edge_index, func.x0 = dev(ei), dev(x)
    from grand_b200 import integrators
    xg = dev(x).requires_grad_(True)
    t = torch.tensor([0.0, 2.0])
    zT = integrators.odeint_adjoint(func, xg, t, method='rk4', options=dict(step_size=1.0))[1]
    L = func.multihead_att_layer
    params = [xg, L.Q.weight, L.K.weight, func.alpha_train, func.beta_train]
    grads = torch.autograd.grad(zT, params, dev(gout))

    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)

    class Ref(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.wq, self.bq, self.wk, self.bk = (torch.nn.Parameter(t64(z[k])) for k in ('wq', 'bq', 'wk', 'bk'))
            self.a, self.b = torch.nn.Parameter(t64(0.3)), torch.nn.Parameter(t64(0.1))

        def forward(self, tt, y):
            return funcs.transformer_f(y, torch.tensor(ei), self.wq, self.bq, self.wk, self.bk, heads, self.a, self.b,
                                       t64(x), attention_type=att_type)

    ref = Ref()
    xo = t64(x).requires_grad_(True)
    zr = oi.odeint_adjoint(ref, xo, [0.0, 2.0], 'rk4', 1.0)[1]
    rg = torch.autograd.grad(zr, [xo, ref.wq, ref.wk, ref.a, ref.b], t64(gout))
    assert_close(zT, zr, 1e-4, 'z(T)')
    for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gwk', 'galpha', 'gbeta')):
        assert_close(g, r, 1e-4, name)


def test_rmat_scale16_parity():
    """SURVEY.md 8d parity gate: a 2^16-node R-MAT graph (edge factor 16, duplicates coalesced by the plan's
    ordering contract: kept and added), GRAND-l and GRAND-nl evaluation + all gradients vs the fp64 oracle"""
    import grand_b200
    scale, d, heads, adim = 16, 64, 4, 64
    n = 1 << scale
    ei, w = og.reset_graph_data(og.rmat_graph(scale, 16, 3), None, n, 'rw', 1.0)
    rng = np.random.default_rng(2)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    eit = torch.tensor(ei)
    # GRAND-l
    lopt = dict(BASE_OPT, function='laplacian', block='constant', hidden_dim=d)
    lf = grand_b200.LaplacianODEFunc(d, d, lopt, device()).to(device())
    with torch.no_grad():
        lf.alpha_train.fill_(0.3)
        lf.beta_train.fill_(0.1)
    lf.edge_index, lf.edge_weight, lf.x0 = dev(ei), dev(w), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = lf(0.0, xg)
    gx, ga, gb = torch.autograd.grad(out, [xg, lf.alpha_train, lf.beta_train], dev(gout))
    xo, a, b = t64(x).requires_grad_(True), t64(0.3).requires_grad_(True), t64(0.1).requires_grad_(True)
    ref = funcs.laplacian_f(xo, eit, t64(w), a, b, t64(x0))
    rgx, rga, rgb = torch.autograd.grad(ref, [xo, a, b], t64(gout))
    assert_close(out, ref, F_RTOL, 'GRAND-l f')
    for g, r, name in ((gx, rgx, 'gx'), (ga, rga, 'galpha'), (gb, rgb, 'gbeta')):
        assert_close(g, r, G_RTOL, 'GRAND-l ' + name)
    # GRAND-nl
    z = dict(wq=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bq=0.1 * rng.standard_normal(adim).astype(np.float32),
             wk=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bk=0.1 * rng.standard_normal(adim).astype(np.float32),
             alpha_train=0.3, beta_train=0.1)
    func = _func(dict(BASE_OPT, hidden_dim=d, heads=heads, attention_dim=adim), z)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    xg = dev(x).requires_grad_(True)
    out = func(0.0, xg)
    L = func.multihead_att_layer
    grads = torch.autograd.grad(out, [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train],
                                dev(gout))
    leaves = [t64(v).requires_grad_(True) for v in (x, z['wq'], z['bq'], z['wk'], z['bk'])]
    a, b = t64(0.3).requires_grad_(True), t64(0.1).requires_grad_(True)
    ref = funcs.transformer_f(leaves[0], eit, leaves[1], leaves[2], leaves[3], leaves[4], heads, a, b, t64(x0))
    rg = torch.autograd.grad(ref, leaves + [a, b], t64(gout))
    assert_close(out, ref, F_RTOL, 'GRAND-nl f')
    bias_scale = max(float(rg[2].abs().max()), float(rg[4].abs().max()))
    for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
        assert_close(g, r, G_RTOL, 'GRAND-nl ' + name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_transformer_exp_kernel_*.npz'))))
def test_exp_kernel_gradients_match_the_reference(path):
    """training path of attention_type exp_kernel: projection -> edge softmax (exp_kernel forward kernel) -> SpMM,
    backward through ops.exp_kernel_score_vjp over the multi-head SpMM kernels; incl. output_var / lengthscale"""
    z = np.load(path)
    n, d = z['x'].shape
    opt = dict(BASE_OPT, attention_type='exp_kernel', attention_norm_idx=int(z['attention_norm_idx']), hidden_dim=d,
               heads=int(z['heads']), attention_dim=int(z['wq'].shape[0]))
    func = _func(opt, z)
    L = func.multihead_att_layer
    with torch.no_grad():
        L.output_var.fill_(float(z['output_var']))
        L.lengthscale.fill_(float(z['lengthscale']))
    func.edge_index, func.x0 = dev(z['edge_index']), dev(z['x0'])
    xg = dev(z['x']).requires_grad_(True)
    out = func(0.0, xg)
    params = [xg, func.alpha_train, func.beta_train, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, L.output_var,
              L.lengthscale]
    grads = torch.autograd.grad(out, params, dev(z['gout']))
    assert_close(out, z['f'], F_RTOL, 'f')
    keys = ('gx', 'galpha', 'gbeta', 'gwq', 'gbq', 'gwk', 'gbk', 'goutput_var', 'glengthscale')
    for g, key in zip(grads, keys):
        assert_close(g.reshape(-1), np.asarray(z[key]).reshape(-1), G_RTOL, key)


@pytest.mark.parametrize('att_type,norm_idx,reweight', [('scaled_dot', 0, False), ('scaled_dot', 1, True),
                                                         ('exp_kernel', 0, True), ('cosine_sim', 1, False)])
def test_square_plus_and_reweighted_exp_kernel_train(att_type, norm_idx, reweight):
    """opt['square_plus'] (utils.squareplus instead of the segment softmax) and the exp_kernel backward under
    reweight_attention: f and every gradient against the fp64 oracle on a graph with hub segments"""
    import grand_b200
    from helpers import rand_graph
    n, d, heads, adim = 500, 32, 4, 32
    ei = rand_graph(n, 4000, 31, hub=(9, 600))
    rng = np.random.default_rng(6)
    x, x0, gout = (0.5 * rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    ew = (0.5 + rng.random(ei.shape[1])).astype(np.float32)
    z = dict(wq=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bq=0.1 * rng.standard_normal(adim).astype(np.float32),
             wk=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bk=0.1 * rng.standard_normal(adim).astype(np.float32),
             alpha_train=0.3, beta_train=0.1)
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    for square_plus in (True, False):
        if not square_plus and not (att_type == 'exp_kernel' and reweight):
            continue                                               # the plain softmax cases are covered elsewhere
        opt = dict(BASE_OPT, attention_type=att_type, attention_norm_idx=norm_idx, hidden_dim=d, heads=heads,
                   attention_dim=adim, square_plus=square_plus, reweight_attention=reweight)
        import grand_b200
        func = grand_b200.ODEFuncTransformerAtt(d, d, opt, device()).to(device())
        L = func.multihead_att_layer
        L.edge_weights = dev(ew) if reweight else None            # per edge, in the caller's edge order
        with torch.no_grad():
            for lin, w, b in ((L.Q, 'wq', 'bq'), (L.K, 'wk', 'bk')):
                lin.weight.copy_(dev(z[w]))
                lin.bias.copy_(dev(z[b]))
            func.alpha_train.fill_(0.3)
            func.beta_train.fill_(0.1)
            if att_type == 'exp_kernel':
                L.output_var.fill_(1.2)
                L.lengthscale.fill_(1.5)
        func.edge_index, func.x0 = dev(ei), dev(x0)
        xg = dev(x).requires_grad_(True)
        params = [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias] + ([L.output_var, L.lengthscale] if att_type == 'exp_kernel' else [])
        out = func(0.0, xg)
        grads = torch.autograd.grad(out, params, dev(gout))
        leaves = [t64(v).requires_grad_(True) for v in (x, z['wq'], z['bq'], z['wk'], z['bk'])]
        kw = dict(attention_type=att_type, attention_norm_idx=norm_idx, square_plus=square_plus,
                  edge_weight=t64(ew) if reweight else None, reweight_attention=reweight)
        extra = []
        if att_type == 'exp_kernel':
            extra = [t64(1.2).requires_grad_(True), t64(1.5).requires_grad_(True)]
            kw.update(output_var=extra[0], lengthscale=extra[1])
        ref = funcs.transformer_f(leaves[0], torch.tensor(ei), leaves[1], leaves[2], leaves[3], leaves[4], heads,
                                  t64(0.3), t64(0.1), t64(x0), **kw)
        rg = torch.autograd.grad(ref, leaves + extra, t64(gout))
        tag = f'{att_type} norm{norm_idx} sq{int(square_plus)}'
        assert_close(out, ref, F_RTOL, 'f ' + tag)
        scale = max(float(r.abs().max()) for r in rg[1:5])
        for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'gov', 'gls')):
            assert_close(g.reshape(-1), r.reshape(-1), G_RTOL, f'{name} {tag}', floor=scale if name in ('gbq', 'gbk') else 0.0)
        with torch.no_grad():
            assert_close(func(0.0, dev(x)), ref, F_RTOL, 'f (no grad) ' + tag)


def test_exp_kernel_gradients_on_a_graph_with_hub_segments():
    import grand_b200
    from helpers import rand_graph
    n, d, heads, adim = 600, 32, 4, 32
    ei = rand_graph(n, 5000, 21, hub=(5, 700))                    # > 256-edge row and column: long-segment kernels
    rng = np.random.default_rng(4)
    x, x0, gout = (0.5 * rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    z = dict(wq=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bq=0.1 * rng.standard_normal(adim).astype(np.float32),
             wk=rng.standard_normal((adim, d)).astype(np.float32) / np.sqrt(d), bk=0.1 * rng.standard_normal(adim).astype(np.float32),
             alpha_train=0.3, beta_train=0.1)
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    for norm_idx in (0, 1):
        opt = dict(BASE_OPT, attention_type='exp_kernel', attention_norm_idx=norm_idx, hidden_dim=d, heads=heads,
                   attention_dim=adim)
        func = _func(opt, z)
        L = func.multihead_att_layer
        with torch.no_grad():
            L.output_var.fill_(1.2)
            L.lengthscale.fill_(1.5)
        func.edge_index, func.x0 = dev(ei), dev(x0)
        xg = dev(x).requires_grad_(True)
        out = func(0.0, xg)
        grads = torch.autograd.grad(out, [xg, L.Q.weight, L.K.weight, L.output_var, L.lengthscale], dev(gout))
        leaves = [t64(v).requires_grad_(True) for v in (x, z['wq'], z['bq'], z['wk'], z['bk'])]
        ov, ls = t64(1.2).requires_grad_(True), t64(1.5).requires_grad_(True)
        ref = funcs.transformer_f(leaves[0], torch.tensor(ei), leaves[1], leaves[2], leaves[3], leaves[4], heads,
                                  t64(0.3), t64(0.1), t64(x0), attention_type='exp_kernel', attention_norm_idx=norm_idx,
                                  output_var=ov, lengthscale=ls)
        rg = torch.autograd.grad(ref, [leaves[0], leaves[1], leaves[3], ov, ls], t64(gout))
        assert_close(out, ref, F_RTOL, f'f norm{norm_idx}')
        for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gwk', 'goutput_var', 'glengthscale')):
            assert_close(g.reshape(-1), r.reshape(-1), G_RTOL, f'{name} norm{norm_idx}')


@pytest.mark.parametrize('norm_idx,labels', [(0, 0), (1, 4)])
def test_beltrami_split_exp_kernel_matches_the_oracle(norm_idx, labels):
    """beltrami + exp_kernel (function_transformer_attention.py:164-216): feature kernel x positional kernel, run as
    ONE exp_kernel evaluation on per-head vectors [qx / lx, qp / lp] of width 2 d_k; attention, f and all gradients
    (x, the four projections, both output variances and lengthscales) against the literal two-kernel product"""
    import grand_b200
    from helpers import rand_graph
    n, feat, pos, heads, adim = 300, 16, 8, 2, 8
    d = feat + pos + labels
    ei = rand_graph(n, 2400, 33, hub=(3, 400))
    rng = np.random.default_rng(8)
    x, x0, gout = (0.5 * rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    opt = dict(BASE_OPT, attention_type='exp_kernel', beltrami=True, attention_norm_idx=norm_idx, hidden_dim=d,
               feat_hidden_dim=feat, pos_enc_hidden_dim=pos, heads=heads, attention_dim=adim)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, device()).to(device())
    L = func.multihead_att_layer
    names = ['Qx.weight', 'Qx.bias', 'Kx.weight', 'Kx.bias', 'Qp.weight', 'Qp.bias', 'Kp.weight', 'Kp.bias',
             'output_var_x', 'lengthscale_x', 'output_var_p', 'lengthscale_p']
    params = dict(L.named_parameters())
    with torch.no_grad():
        for k in names[:8]:
            params[k].copy_(dev((rng.standard_normal(tuple(params[k].shape)) / 3).astype(np.float32)))
        for k, v in zip(names[8:], (1.2, 1.5, 0.8, 0.7)):
            params[k].fill_(v)
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    assert params['Qx.weight'].shape == (adim, d - pos) and params['Qp.weight'].shape == (adim, pos)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    t64 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)
    w64 = {k: t64(params[k].detach().cpu().numpy()).requires_grad_(True) for k in names}
    x64 = t64(x).requires_grad_(True)
    ei_t = torch.tensor(ei)
    att_ref, prods_ref = funcs.beltrami_attention_layer(x64, ei_t, w64, heads, feat, pos, attention_norm_idx=norm_idx)
    with torch.no_grad():                                        # the layer alone, caller's edge order
        att, (_v, prods) = L(dev(x), dev(ei))
    assert_close(prods, prods_ref, F_RTOL, 'prods')
    assert_close(att, att_ref, F_RTOL, 'attention')
    xg = dev(x).requires_grad_(True)
    out = func(0.0, xg)
    grads = torch.autograd.grad(out, [xg, func.alpha_train, func.beta_train] + [params[k] for k in names], dev(gout))
    a64, b64 = t64(0.3).requires_grad_(True), t64(0.1).requires_grad_(True)
    ref = funcs.laplacian_f(x64, ei_t, att_ref.mean(dim=1), a64, b64, t64(x0))
    rg = torch.autograd.grad(ref, [x64, a64, b64] + [w64[k] for k in names], t64(gout))
    assert_close(out, ref, F_RTOL, 'f')
    for g, r, name in zip(grads, rg, ['gx', 'galpha', 'gbeta'] + names):
        assert_close(g.reshape(-1), r.reshape(-1), G_RTOL, name, floor=1e-4 * float(rg[0].abs().max()))


@pytest.mark.parametrize('norm_idx,heads', [(0, 2), (1, 4), (0, 8), (1, 1), (0, 3)])
def test_gat_function_matches_the_oracle(norm_idx, heads):
    """ODEFuncAtt (function_GAT_attention.py): per-node score halves x (W_h a) -> gb_gat_softmax_fwd (leaky relu +
    segment softmax + head mean over H gathered floats per edge) -> fused SpMM, against the literal restatement
    (gather both endpoints' projections, a * edge_h); scores, attention, f and every gradient; hub row and column
    longer than the plan's long-segment threshold"""
    import grand_b200
    from helpers import rand_graph
    n, d, adim = 500, 24, (12 if heads == 3 else 16)
    ei = rand_graph(n, 4000, 17, hub=(7, 600))
    rng = np.random.default_rng(2)
    x, x0, gout = (0.7 * rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    opt = dict(BASE_OPT, function='GAT', attention_norm_idx=norm_idx, hidden_dim=d, heads=heads, attention_dim=adim)
    func = grand_b200.ODEFuncAtt(d, d, opt, device()).to(device())
    L = func.multihead_att_layer
    assert sorted(k for k, _ in L.named_parameters()) == ['W', 'Wout', 'a'] and L.a.shape == (1, 2 * adim // heads, 1, 1)
    with torch.no_grad():
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    func.edge_index, func.x0 = dev(ei), dev(x0)
    t64 = lambda v: torch.tensor(np.asarray(v), dtype=torch.float64)
    w64 = t64(L.W.detach().cpu().numpy()).requires_grad_(True)
    a64 = t64(L.a.detach().cpu().numpy().reshape(-1)).requires_grad_(True)
    x64 = t64(x).requires_grad_(True)
    al64, be64 = t64(0.3).requires_grad_(True), t64(0.1).requires_grad_(True)
    ei_t = torch.tensor(ei)
    att_ref, prods_ref = funcs.gat_attention_layer(x64, ei_t, w64, a64, heads, 0.2, norm_idx)
    with torch.no_grad():
        att, wx = L(dev(x), dev(ei))
    assert_close(att, att_ref, F_RTOL, 'attention')
    assert_close(wx, x64 @ w64, F_RTOL, 'wx')
    xg = dev(x).requires_grad_(True)
    out = func(0.0, xg)
    grads = torch.autograd.grad(out, [xg, func.alpha_train, func.beta_train, L.W, L.a], dev(gout))
    ref = funcs.gat_f(x64, ei_t, w64, a64, heads, 0.2, al64, be64, t64(x0), attention_norm_idx=norm_idx)
    rg = torch.autograd.grad(ref, [x64, al64, be64, w64, a64], t64(gout))
    assert_close(out, ref, F_RTOL, 'f')
    for g, r, name in zip(grads, rg, ('gx', 'galpha', 'gbeta', 'gW', 'ga')):
        assert_close(g.reshape(-1), r.reshape(-1), G_RTOL, name)
    assert func.nfe == 1


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_gat_*.npz'))))
def test_gat_function_matches_the_reference(path):
    """ODEFuncAtt against the numbers the reference's own function_GAT_attention.py produced (tests/golden/ref_gat_*:
    forward values; the gradients are checked against the gradcheck-pinned oracle in the test above)"""
    import grand_b200
    z = np.load(path)
    n, d = z['x'].shape
    opt = dict(BASE_OPT, function='GAT', attention_norm_idx=int(z['attention_norm_idx']), hidden_dim=d,
               heads=int(z['heads']), attention_dim=int(z['w'].shape[1]), leaky_relu_slope=float(z['slope']))
    func = grand_b200.ODEFuncAtt(d, d, opt, device()).to(device())
    L = func.multihead_att_layer
    with torch.no_grad():
        L.W.copy_(dev(z['w']))
        L.a.copy_(dev(z['a']))
        func.alpha_train.fill_(float(z['alpha_train']))
        func.beta_train.fill_(float(z['beta_train']))
    func.edge_index, func.x0 = dev(z['edge_index']), dev(z['x0'])
    with torch.no_grad():
        att, wx = L(dev(z['x']), dev(z['edge_index']))
    assert_close(att, z['attention'], F_RTOL, 'attention')
    assert_close(wx, z['wx'], F_RTOL, 'wx')
    with torch.no_grad():
        out = func(0.0, dev(z['x']))
    assert_close(out, z['f'], F_RTOL, 'f')
    assert func.nfe == int(z['nfe'])


def test_gat_function_under_the_constant_block_with_the_adjoint_solver():
    """function 'GAT' through the registry and ConstantODEblock, rk4 + odeint_adjoint (autograd fallback of the
    adjoint dynamics: no fused eval_vjp for this function) against the oracle integration"""
    import grand_b200
    from oracle import integrate as oint
    n, e, d, heads, adim, T = 150, 900, 16, 4, 16, 2.0
    opt = dict(BLOCK_OPT, function='GAT', block='constant', hidden_dim=d, heads=heads, attention_dim=adim, adjoint=True)
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device(), t=torch.tensor([0., T])).to(device())
    block.train()
    ei = og.shaped_graph(n, e, 4)
    data = grand_b200.GraphData()
    data.new_graph(dev(ei), n)
    func = block.odefunc
    L = func.multihead_att_layer
    with torch.no_grad():
        func.alpha_train.fill_(-0.2)
        func.beta_train.fill_(0.4)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(n, d, generator=g)
    gout = torch.randn(n, d, generator=g)
    xg = dev(x.numpy()).requires_grad_(True)
    block.set_x0(xg)
    z = block(xg, data)
    grads = torch.autograd.grad(z, [xg, func.alpha_train, func.beta_train, L.W, L.a], dev(gout.numpy()))
    t64 = lambda v: torch.tensor(np.asarray(v), dtype=torch.float64)
    ei_n, _ = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    leaves = [t64(x.numpy()), t64(-0.2), t64(0.4), t64(L.W.detach().cpu().numpy()),
              t64(L.a.detach().cpu().numpy().reshape(-1))]
    leaves = [v.requires_grad_(True) for v in leaves]
    f = lambda t, y: funcs.gat_f(y, torch.tensor(ei_n), leaves[3], leaves[4], heads, 0.2, leaves[1], leaves[2],
                                 t64(x.numpy()))
    zo = oint.odeint_adjoint(f, leaves[0], [0., T], 'rk4', 1.0, adjoint_params=tuple(leaves[1:]))[1]
    ro = torch.autograd.grad(zo, leaves, t64(gout.numpy()))
    assert_close(z, zo, 1e-4, 'z(T)')
    for got, r, name in zip(grads, ro, ('gx', 'galpha', 'gbeta', 'gW', 'ga')):
        assert_close(got.reshape(-1), r.reshape(-1), 1e-4, name)


def test_utils_softmax_matches_the_reference_and_its_invariants():
    """stand-alone utils.softmax (utils.py:116-127) on ready-made scores with an UNSORTED group vector: the golden the
    reference itself produced (tests/golden/ref_softmax.npz), the reference tests' invariants (every group sums to
    one, values in (0, 1]; test/test_transformer_attention.py:57-76), batched layout, empty groups, gradient vs the
    oracle, run-to-run identity"""
    from grand_b200 import utils
    z = np.load(os.path.join(GOLD, 'ref_softmax.npz'))
    out = utils.softmax(dev(z['src']), dev(z['index']), int(z['n']))
    assert_close(out, z['out'], 1e-6, 'softmax vs reference golden')
    out_b = utils.softmax(dev(z['src'])[None], dev(z['index'])[None], int(z['n']))
    assert out_b.shape == (1,) + tuple(z['src'].shape) and torch.equal(out_b[0], out)
    rng = np.random.default_rng(5)
    n, e, h = 700, 9000, 4
    index = rng.integers(0, n, e)
    index[:3000] = 13                                                  # one long group
    index[index == 200] = 201                                          # and an empty one
    src = (3 * rng.standard_normal((e, h))).astype(np.float32)
    sg = dev(src).requires_grad_(True)
    att = utils.softmax(sg, dev(index), n)
    sums = torch.zeros(n, h, device=att.device).index_add_(0, dev(index).long(), att.detach())
    present = torch.zeros(n, dtype=torch.bool, device=att.device)
    present[dev(index).long()] = True
    assert torch.allclose(sums[present], torch.ones_like(sums[present]), atol=1e-5) and not bool(present[200])
    assert float(att.detach().min()) > 0 and float(att.detach().max()) <= 1
    gout = rng.standard_normal((e, h)).astype(np.float32)
    g, = torch.autograd.grad(att, sg, dev(gout))
    s64 = torch.tensor(src, dtype=torch.float64, requires_grad=True)
    ref = funcs.segment_softmax(s64, torch.tensor(index), n)
    rg, = torch.autograd.grad(ref, s64, torch.tensor(gout, dtype=torch.float64))
    assert_close(att, ref, F_RTOL, 'softmax')
    assert_close(g, rg, G_RTOL, 'd softmax', floor=1e-4 * float(rg.abs().max()))
    assert torch.equal(utils.softmax(dev(src), dev(index), n), att.detach())


# ---- MixedODEblock / HardAttODEblock (host compositions over the verified kernels) ----------------------------
BLOCK_OPT = {'hidden_dim': 16, 'block': 'mixed', 'function': 'laplacian', 'max_nfe': 10000, 'no_alpha_sigmoid': False,
             'add_source': True, 'augment': False, 'adjoint': False, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0,
             'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': 'rk4', 'step_size': 1.0, 'max_iters': 100,
             'adjoint_method': 'rk4', 'adjoint_step_size': 1.0, 'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 16,
             'attention_norm_idx': 0, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
             'square_plus': False, 'mix_features': False, 'multi_modal': False, 'kinetic_energy': None,
             'jacobian_norm2': None, 'total_deriv': None, 'directional_penalty': None, 'att_samp_pct': 0.6,
             'use_flux': False}


def _block(opt, n, e, T, seed):
    import grand_b200
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device(),
                                      t=torch.tensor([0., T])).to(device())
    ei = og.shaped_graph(n, e, seed)
    data = grand_b200.GraphData()
    data.new_graph(dev(ei), n)
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        L = block.multihead_att_layer
        for lin in (L.Q, L.K):
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / lin.weight.shape[1] ** 0.5).to(device()))
            lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=g)).to(device()))
        block.odefunc.alpha_train.fill_(0.4)
        block.odefunc.beta_train.fill_(0.2)
    return block, data, ei


def _t64(a):
    return torch.tensor(np.asarray(a.detach().cpu() if torch.is_tensor(a) else a), dtype=torch.float64)


def test_mixed_block_state_and_gradients_incl_the_gate():
    from oracle import integrate as oint
    n, e, d, T = 250, 1500, 16, 2.0
    block, data, ei = _block(dict(BLOCK_OPT), n, e, T, 4)
    block.train()
    with torch.no_grad():
        block.gamma.fill_(0.3)
    x = torch.randn(n, d, device=device(), requires_grad=True)
    gout = torch.randn(n, d, device=device())
    block.set_x0(x)
    z = block(x, data)
    B = block.multihead_att_layer
    params = [x, B.Q.weight, B.K.weight, block.gamma, block.odefunc.alpha_train]
    grads = torch.autograd.grad(z, params, gout)
    ei_n, w = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    eit = torch.tensor(ei_n)
    leaves = [_t64(p).requires_grad_(True) for p in params]
    att, _ = funcs.attention_layer(leaves[0], eit, leaves[1], _t64(B.Q.bias), leaves[2], _t64(B.K.bias), 4)
    gate = torch.sigmoid(leaves[3])
    mixed = att.mean(1) * (1 - gate) + _t64(w) * gate
    x0 = _t64(x)
    f = lambda t, y: funcs.laplacian_f(y, eit, mixed, leaves[4], _t64(0.2), x0)
    zo = oint.odeint(f, leaves[0], [0., T], 'rk4', 1.0)[1]
    ro = torch.autograd.grad(zo, leaves, _t64(gout))
    assert_close(z, zo, 1e-4, 'z(T)')
    for g, r, name in zip(grads, ro, ('gx', 'gwq', 'gwk', 'ggamma', 'galpha')):
        assert_close(g.reshape(-1), r.reshape(-1), 1e-4, name)


@pytest.mark.parametrize('function', ['laplacian', 'transformer', 'GAT'])
def test_hard_attention_block_samples_edges_while_training(function, capsys):
    from oracle import integrate as oint
    n, e, d, T = 300, 2400, 16, 2.0
    opt = dict(BLOCK_OPT, block='hard_attention', function=function)
    import grand_b200
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device(),
                                      t=torch.tensor([0., T])).to(device())
    ei = og.shaped_graph(n, e, 9)
    data = grand_b200.GraphData()
    data.new_graph(dev(ei), n)
    L = block.multihead_att_layer if function == 'laplacian' else block.odefunc.multihead_att_layer
    g = torch.Generator().manual_seed(2)
    with torch.no_grad():
        for lin in (() if function == 'GAT' else (L.Q, L.K)):             # GAT keeps its xavier init (W, a)
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / 4.0).to(device()))
        block.odefunc.alpha_train.fill_(0.4)
        block.odefunc.beta_train.fill_(0.2)
    x = torch.randn(n, d, device=device())
    block.set_x0(x)
    block.train()
    z = block(x, data)
    func = block.odefunc
    e_full = int(block.data_edge_index.shape[-1])
    kept = func.edge_index[0] if func.edge_index.dim() == 3 else func.edge_index
    assert abs(int(kept.shape[1]) - 0.6 * e_full) <= 2
    assert 'retaining' in capsys.readouterr().out
    w = func.attention_weights.reshape(-1)
    sums = torch.zeros(n, device=device()).index_add_(0, kept[0], w)
    present = torch.zeros(n, dtype=torch.bool, device=device())
    present[kept[0]] = True
    assert_close(sums[present], torch.ones(int(present.sum())), 1e-5, 'kept rows sum to one')
    eit = kept.cpu()
    x0 = _t64(x)
    if function == 'laplacian':
        f = lambda t, y: funcs.laplacian_f(y, eit, _t64(w), _t64(0.4), _t64(0.2), x0)
    elif function == 'GAT':
        f = lambda t, y: funcs.gat_f(y, eit, _t64(L.W), _t64(L.a).reshape(-1), 4, 0.2, _t64(0.4), _t64(0.2), x0)
    else:
        f = lambda t, y: funcs.transformer_f(y, eit, _t64(L.Q.weight), _t64(L.Q.bias), _t64(L.K.weight), _t64(L.K.bias), 4,
                                             _t64(0.4), _t64(0.2), x0)
    zo = oint.odeint(f, _t64(x), [0., T], 'rk4', 1.0)[1]
    assert_close(z, zo, 1e-4, 'z(T) on the sampled graph')
    block.eval()                                                  # eval mode: the full graph, head-mean attention
    with torch.no_grad():
        z_eval = block(x, data)
    assert func.edge_index.shape[-1] == e_full and z_eval.shape == x.shape


def test_two_phase_adjoint_is_bit_identical_to_the_single_call(monkeypatch):
    """gb_nl_adjoint_phase (row pass, then column pass) through the phased adjoint solve on a 1-rank partition
    (LoopbackHalo) against the ordinary partitioned solve (gb_nl_adjoint in one call): same kernels, same buffers,
    same order of arithmetic -> torch.equal"""
    import grand_b200
    n, d = 400, 32
    opt = dict(BLOCK_OPT, block='attention', function='transformer', adjoint=True, hidden_dim=d)
    ei = dev(og.shaped_graph(n, 2400, 9))
    g = torch.Generator().manual_seed(0)
    x, gz = dev(torch.randn(n, d, generator=g).numpy()), dev(torch.randn(n, d, generator=g).numpy())
    results = []
    for overlap in (False, True):
        monkeypatch.setenv('GRANDB200_OVERLAP', '1' if overlap else '0')
        block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device(),
                                          t=torch.tensor([0., 2.])).to(device())
        block.train()
        L = block.odefunc.multihead_att_layer
        gw = torch.Generator().manual_seed(5)
        with torch.no_grad():
            for lin in (L.Q, L.K):
                lin.weight.copy_((torch.randn(lin.weight.shape, generator=gw) / 6).to(device()))
                lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=gw)).to(device()))
            block.odefunc.alpha_train.fill_(0.3)
            block.odefunc.beta_train.fill_(0.1)
        part = grand_b200.RowPartition(n, 1, 0)
        block.set_partition(part, grand_b200.LoopbackHalo(part))
        data = grand_b200.GraphData()
        data.new_graph(ei, n)
        xg = x.clone().requires_grad_(True)
        block.set_x0(xg)
        z = block(xg, data)
        params = [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, block.odefunc.alpha_train, block.odefunc.beta_train]
        results.append([z.detach()] + list(torch.autograd.grad((z * gz).sum(), params)))
        assert block.odefunc.nfe == 16
    for a, b in zip(*results):
        assert torch.equal(a, b)


# ---- model wrappers (GNN / GNNEarly): the CPU tests of tests/test_gnn_host_cpu.py on the device -------------------
@pytest.mark.parametrize('batched', [False, True])
@pytest.mark.parametrize('adjoint', [False, True])
def test_gnn_logits_and_gradients(batched, adjoint):
    import test_gnn_host_cpu as tg
    tg.test_gnn_logits_and_gradients_match_the_oracle_composition(None, batched, adjoint)


def test_gnn_early_and_option_branches():
    import test_gnn_host_cpu as tg
    tg.test_gnn_early_runs_the_early_stop_integrator_at_test_time(None)
    tg.test_base_gnn_option_branches_build_and_run(None)
