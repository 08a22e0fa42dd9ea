"""GPU, at BASELINE.json's full size (c4: ogbn-products-shaped, N = 2 449 029, E' = 64 308 169, d = 128,
H = 4, A = 64) where the oracle cannot evaluate the whole graph: size-independent properties of the hot path,
plus an oracle check on sampled rows (a row's output depends only on its own edges, so the oracle on the
sub-graph "all edges of the sampled rows" must reproduce those rows of the full evaluation)."""
import numpy as np
import pytest
import torch

from oracle import funcs
from helpers import assert_close
from test_gpu_partition import make_func

pytestmark = pytest.mark.gpu

D, HEADS, ADIM = 128, 4, 64


@pytest.fixture(scope='module')
def c4():
    import grand_b200
    from grand_b200 import synth, utils
    dev = torch.device('cuda')
    ei, n = synth.make_graph('c4', dev)
    ei, w = utils.get_rw_adj(ei, edge_weight=None, norm_dim=1, fill_value=1.0, num_nodes=n, dtype=torch.float32)
    gen = torch.Generator(device=dev).manual_seed(7)
    x = torch.randn(n, D, device=dev, generator=gen)
    x0 = torch.randn(n, D, device=dev, generator=gen)
    g = torch.randn(n, D, device=dev, generator=gen)
    func = make_func(D, HEADS, ADIM)
    func.edge_index, func.edge_weight, func.x0 = ei, w, x0
    lap = grand_b200.LaplacianODEFunc(D, D, {'hidden_dim': D, 'block': 'constant', 'max_nfe': 10 ** 6,
                                             'no_alpha_sigmoid': False, 'add_source': True}, dev).cuda()
    with torch.no_grad():
        lap.alpha_train.fill_(0.4)
        lap.beta_train.fill_(-0.2)
    lap.edge_index, lap.edge_weight, lap.x0 = ei, w, x0
    return dict(n=n, ei=ei, w=w, x=x, x0=x0, g=g, func=func, lap=lap)


def _params(func):
    L = func.multihead_att_layer
    return [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]


def test_sampled_rows_match_the_oracle(c4):
    """256 random rows (and the 8 longest) of f and of the fused eval_vjp's f against the oracle on their edges"""
    func, ei, n = c4['func'], c4['ei'], c4['n']
    with torch.no_grad():
        f = func(0.0, c4['x'])
    f2, dx, _ = func.eval_vjp(0.0, c4['x'], c4['g'], _params(func))
    assert_close(f2, f, 1e-6, 'f of the adjoint kernel vs the forward kernel')
    deg = torch.bincount(ei[0], minlength=n)
    rows = torch.cat([torch.randperm(n, generator=torch.Generator().manual_seed(3))[:256].cuda(),
                      deg.topk(8).indices])
    mark = torch.zeros(n, dtype=torch.bool, device='cuda')
    mark[rows] = True
    sub = ei[:, mark[ei[0]]]                                   # every edge of the sampled rows
    nodes, inv = torch.unique(torch.cat([rows, sub.reshape(-1)]), return_inverse=True)
    loc_rows, loc_ei = inv[:rows.numel()], inv[rows.numel():].reshape(2, -1)
    L = func.multihead_att_layer
    c = lambda t: t.detach().double().cpu()
    ref = funcs.transformer_f(c(c4['x'][nodes]), loc_ei.cpu(), c(L.Q.weight), c(L.Q.bias), c(L.K.weight),
                              c(L.K.bias), HEADS, c(func.alpha_train), c(func.beta_train), c(c4['x0'][nodes]))
    assert int(deg[rows].max()) > 2048                         # the long-row kernel is among the sampled rows
    assert_close(f[rows], ref[loc_rows.cpu()], 1e-5, 'sampled rows of f')


def test_sampled_columns_of_the_vjp_match_the_oracle(c4):
    """dL/dx_j depends on the rows with an edge into j (their whole softmax rows) and on row j itself: the oracle's
    autograd on the sub-problem "all out-edges of S and of every in-neighbour of S" reproduces dx[S] of the full-size
    fused VJP exactly (as a formula) - 48 random columns plus the column with the most in-edges among 2000 candidates"""
    func, ei, n = c4['func'], c4['ei'], c4['n']
    _, dx, _ = func.eval_vjp(0.0, c4['x'], c4['g'], _params(func))
    cand = torch.randperm(n, generator=torch.Generator().manual_seed(5))[:2000].cuda()
    indeg = torch.bincount(ei[1], minlength=n)
    cols = torch.cat([cand[:48], cand[indeg[cand].argmax()][None]])
    in_s = torch.zeros(n, dtype=torch.bool, device='cuda')
    in_s[cols] = True
    rows = torch.unique(torch.cat([cols, ei[0, in_s[ei[1]]]]))          # S and every row with an edge into S
    in_r = torch.zeros(n, dtype=torch.bool, device='cuda')
    in_r[rows] = True
    sub = ei[:, in_r[ei[0]]]                                            # every out-edge of those rows
    nodes, inv = torch.unique(torch.cat([cols, rows, sub.reshape(-1)]), return_inverse=True)
    loc_cols = inv[:cols.numel()].cpu()
    loc_rows = inv[cols.numel():cols.numel() + rows.numel()].cpu()
    loc_ei = inv[cols.numel() + rows.numel():].reshape(2, -1).cpu()
    assert loc_ei.shape[1] < 3_000_000, loc_ei.shape                    # stays a seconds-sized oracle problem
    L = func.multihead_att_layer
    c = lambda t: t.detach().double().cpu()
    xs = c(c4['x'][nodes]).requires_grad_(True)
    ref = funcs.transformer_f(xs, loc_ei, c(L.Q.weight), c(L.Q.bias), c(L.K.weight), c(L.K.bias), HEADS,
                              c(func.alpha_train), c(func.beta_train), c(c4['x0'][nodes]))
    gsub = torch.zeros_like(ref)
    gsub[loc_rows] = c(c4['g'][rows])                                   # rows outside R do not reach x[S]
    rdx, = torch.autograd.grad(ref, xs, gsub)
    assert_close(dx[cols], rdx[loc_cols], 1e-4, 'sampled columns of dL/dx')
    from helpers import assert_rows_close
    assert_rows_close(dx[cols], rdx[loc_cols], 1e-4, 'sampled columns of dL/dx (per row)')


def test_constant_state_is_a_fixed_point_of_the_attention_diffusion(c4):
    """row softmax is row-stochastic: x_j = v for every node  =>  A(x) x = x  =>  f = beta * x0"""
    func, n = c4['func'], c4['n']
    v = torch.linspace(-1.0, 1.0, D, device='cuda')
    x = v.expand(n, D).contiguous()
    with torch.no_grad():
        f = func(0.0, x)
    err = float((f - func.beta_train.detach() * c4['x0']).abs().max())
    # fp32 sums over up to ~2e4 edges per hub row: |sum_e att_e - 1| ~ 1e-5 at |v| <= 1
    assert err <= 5e-5, f'|f - beta x0| = {err:.3e}'


def test_laplacian_diffusion_conserves_mass(c4):
    """get_rw_adj(norm_dim=1) normalises by the column sums: 1^T A = 1^T  =>  sum_i (f - beta x0)_i = 0"""
    lap, x = c4['lap'], c4['x']
    with torch.no_grad():
        r = (lap(0.0, x) - lap.beta_train.detach() * c4['x0']).double()
    drift, scale = r.sum(0).abs().max(), r.abs().sum(0).max()
    assert float(drift) <= 1e-6 * float(scale), (float(drift), float(scale))


def test_laplacian_is_linear_and_its_vjp_is_its_transpose(c4):
    lap, x, g, x0 = c4['lap'], c4['x'], c4['g'], c4['x0']
    with torch.no_grad():
        src = lap.beta_train.detach() * x0
        f1, f2 = lap(0.0, x) - src, lap(0.0, g) - src
        f12 = lap(0.0, x + 2.0 * g) - src
    assert_close(f12, f1 + 2.0 * f2, 1e-5, 'f(x + 2g) = f(x) + 2 f(g)')
    xg = x.clone().requires_grad_(True)
    f = lap(0.0, xg)
    dx, = torch.autograd.grad(f, xg, g)
    lhs = float((g.double() * f1.double()).sum())              # <g, J x>
    rhs = float((dx.double() * x.double()).sum())              # <J^T g, x>
    assert abs(lhs - rhs) <= 1e-5 * max(abs(lhs), abs(rhs), 1.0) + 1e-3, (lhs, rhs)


def test_vjp_matches_a_directional_difference(c4):
    """<g, f(x + eps v) - f(x - eps v)> / (2 eps) = <vjp_x(g), v>; fp32 forward differences limit this to ~1e-3"""
    func, x, g = c4['func'], c4['x'], c4['g']
    v = torch.randn(x.shape, device='cuda', generator=torch.Generator(device='cuda').manual_seed(11))
    _, dx, _ = func.eval_vjp(0.0, x, g, _params(func))
    eps = 1e-2
    with torch.no_grad():
        fp, fm = func(0.0, x + eps * v), func(0.0, x - eps * v)
    lhs = float((g.double() * (fp.double() - fm.double())).sum()) / (2 * eps)
    rhs = float((dx.double() * v.double()).sum())
    assert abs(lhs - rhs) <= 2e-3 * max(abs(lhs), abs(rhs)), (lhs, rhs)


def test_full_size_results_are_bit_reproducible(c4):
    func, x, g = c4['func'], c4['x'], c4['g']
    a = func.eval_vjp(0.0, x, g, _params(func))
    b = func.eval_vjp(0.0, x, g, _params(func))
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    for p, q in zip(a[2], b[2]):
        assert (p is None and q is None) or torch.equal(p, q)
