"""GPU: the 1-D row-partitioned path (grand_b200/partition.py).  The ranks are emulated one after the other
on one GPU (several processes sharing a GPU must not wait on each other, B200_PROFILING.md): every rank's local
compute runs on the hand-assembled gathered state and the exchange (all-gather / reduce-scatter / all-reduce)
is done with plain tensor sums - the collectives themselves are covered by the gloo test on CPU."""
import numpy as np
import pytest
import torch

from oracle import funcs, graph as og
from helpers import assert_close, dev, rand_graph

pytestmark = pytest.mark.gpu

OPT = {'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 32, 'attention_norm_idx': 0, 'add_source': True,
       'hidden_dim': 32, 'block': 'attention', 'function': 'transformer', 'max_nfe': 10000,
       'no_alpha_sigmoid': False, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
       'square_plus': False, 'mix_features': False, 'multi_modal': False, 'augment': False, 'adjoint': True,
       'tol_scale': 1.0, 'tol_scale_adjoint': 1.0, 'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': 'rk4',
       'step_size': 1.0, 'max_iters': 100, 'adjoint_method': 'rk4', 'adjoint_step_size': 1.0,
       'kinetic_energy': None, 'jacobian_norm2': None, 'total_deriv': None, 'directional_penalty': None}


def make_func(d, heads, adim, seed=3):
    import grand_b200
    opt = dict(OPT, heads=heads, attention_dim=adim, hidden_dim=d)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    g = torch.Generator().manual_seed(seed)
    L = func.multihead_att_layer
    with torch.no_grad():
        for lin in (L.Q, L.K):
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / d ** 0.5).cuda())
            lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=g)).cuda())
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    return func


@pytest.mark.parametrize('world', [1, 2, 3, 8])
@pytest.mark.parametrize('n,d,heads,adim', [(1001, 32, 4, 32), (700, 128, 4, 64), (515, 64, 8, 64)])
def test_nl_ranks_sum_to_single_gpu(world, n, d, heads, adim):
    import grand_b200
    from grand_b200 import partition as P
    ei = rand_graph(n, 9 * n, n, hub=(3, 900))                               # > 256-edge rows and columns
    rng = np.random.default_rng(world)
    x, x0, g = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    func = make_func(d, heads, adim)
    L = func.multihead_att_layer
    func.edge_index, func.x0 = dev(ei), dev(x0)
    params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]
    f_ref, dx_ref, p_ref = func.eval_vjp(0.0, dev(x), dev(g), params)        # single-GPU fused path

    part0 = P.RowPartition(n, world, 0)
    pad = lambda a: torch.cat([dev(a), torch.zeros(part0.n_pad - n, d, device='cuda')])
    x_full, x0_full, g_full = pad(x), pad(x0), pad(g)
    f_parts, dx_share, grads = [], 0, 0
    for r in range(world):
        part = P.RowPartition(n, world, r)
        ei_loc, _ = part.local_edges(dev(ei))
        plan = grand_b200.GraphPlan(ei_loc, part.n_pad)
        sl = slice(r * part.slab, (r + 1) * part.slab)
        x_slab, x0_slab, g_slab = x_full[sl].contiguous(), x0_full[sl].contiguous(), g_full[sl].contiguous()
        out, q = P.nl_local_forward(part, plan, x_full, x_slab, L.Q.weight.detach(), L.Q.bias.detach(),
                                    L.K.weight.detach(), func.alpha_train.detach(), func.beta_train.detach(),
                                    x0_slab, heads, True)
        f2, share, gr = P.nl_local_vjp(part, plan, x_full, q, L.Q.weight.detach(), L.K.weight.detach(),
                                            func.alpha_train.detach(), func.beta_train.detach(), x0_slab, heads, True,
                                            g_slab, True, True)
        assert_close(f2, out, 1e-6, 'f of the adjoint kernel vs the forward kernel')
        assert float(out[part.hi - part.lo:].abs().max() if part.hi - part.lo < part.slab else 0.0) == 0.0
        f_parts.append(out)
        dx_share = dx_share + share                                          # reduce-scatter == sum, then slice
        grads = grads + gr                                                   # all-reduce
    f = torch.cat(f_parts)[:n]
    dx = dx_share[:n]
    assert_close(f, f_ref, 2e-6, 'f')
    assert_close(dx, dx_ref, 1e-5, 'dx')
    a = adim
    dwq, dwk, dbq, dbk, dalpha, dbeta = P.unpack_nl_grads(grads, a, d)
    bias_scale = float(p_ref[1].abs().max())
    for got, ref, name in ((dwq, p_ref[0], 'dWq'), (dbq, p_ref[1], 'dbq'), (dwk, p_ref[2], 'dWk'),
                           (dbk, p_ref[3], 'dbk'), (dalpha, p_ref[4], 'dalpha'), (dbeta, p_ref[5], 'dbeta')):
        assert_close(got, ref, 2e-5, name, floor=bias_scale if name == 'dbk' else 0.0)


@pytest.mark.parametrize('world', [1, 2, 5])
def test_laplacian_ranks_sum_to_single_gpu(world):
    import grand_b200
    from grand_b200 import partition as P
    n, d = 903, 48
    ei, w = og.reset_graph_data(rand_graph(n, 7 * n, 5, self_loops=False, hub=(11, 700)), None, n, 'rw', 1.0)
    rng = np.random.default_rng(1)
    x, x0, g = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    opt = {'hidden_dim': d, 'block': 'constant', 'max_nfe': 100, 'no_alpha_sigmoid': False, 'add_source': True}
    func = grand_b200.LaplacianODEFunc(d, d, opt, torch.device('cuda')).cuda()
    with torch.no_grad():
        func.alpha_train.fill_(0.4)
        func.beta_train.fill_(-0.2)
    func.edge_index, func.edge_weight, func.x0 = dev(ei), dev(w), dev(x0)
    xg = dev(x).requires_grad_(True)
    f_ref = func(0.0, xg)
    dx_ref, da_ref, db_ref = torch.autograd.grad(f_ref, [xg, func.alpha_train, func.beta_train], dev(g))

    part0 = P.RowPartition(n, world, 0)
    pad = lambda a: torch.cat([dev(a), torch.zeros(part0.n_pad - n, d, device='cuda')])
    x_full, x0_full, g_full = pad(x), pad(x0), pad(g)
    f_parts, dx, dparams = [], 0, 0
    for r in range(world):
        part = P.RowPartition(n, world, r)
        ei_loc, w_loc = part.local_edges(dev(ei), dev(w))
        plan = grand_b200.GraphPlan(ei_loc, part.n_pad)
        vals = plan.to_csr_order(w_loc)
        sl = slice(r * part.slab, (r + 1) * part.slab)
        x0_slab, g_slab = x0_full[sl].contiguous(), g_full[sl].contiguous()
        f_parts.append(P.diffusion_local_forward(part, plan, vals, x_full, x0_slab, func.alpha_train.detach(),
                                                 func.beta_train.detach(), True))
        share, dp = P.diffusion_local_vjp(part, plan, vals, x_full, x0_slab, func.alpha_train.detach(), True, g_slab)
        dx, dparams = dx + share, dparams + dp
    assert_close(torch.cat(f_parts)[:n], f_ref, 2e-6, 'f')
    assert_close(dx[:n], dx_ref, 1e-5, 'dx')
    assert_close(dparams[0], da_ref, 2e-5, 'dalpha')
    assert_close(dparams[1], db_ref, 2e-5, 'dbeta')


def test_block_api_world_of_one_matches_unpartitioned():
    """ODEblock.set_partition with a 1-rank partition (LoopbackHalo): same end state and gradients as the
    ordinary block, through odeint_adjoint and through plain autograd"""
    import grand_b200
    n, d = 400, 32
    ei = torch.tensor(og.shaped_graph(n, 2400, 9), device='cuda')
    x = torch.randn(n, d, device='cuda', generator=torch.Generator(device='cuda').manual_seed(0))
    gz = torch.randn(n, d, device='cuda', generator=torch.Generator(device='cuda').manual_seed(1))
    results = []
    for adjoint in (True, False):
        for partitioned in (False, True):
            opt = dict(OPT, adjoint=adjoint, hidden_dim=d)
            torch.manual_seed(0)
            block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, torch.device('cuda'),
                                              t=torch.tensor([0., 2.])).cuda()
            block.train()
            with torch.no_grad():
                for lin in (block.odefunc.multihead_att_layer.Q, block.odefunc.multihead_att_layer.K):
                    lin.weight.copy_(torch.randn(lin.weight.shape, generator=torch.Generator().manual_seed(5)).cuda() / 6)
                block.odefunc.alpha_train.fill_(0.3)
                block.odefunc.beta_train.fill_(0.1)
            if partitioned:
                part = grand_b200.RowPartition(n, 1, 0)
                block.set_partition(part, grand_b200.LoopbackHalo(part))
            data = grand_b200.GraphData()
            data.new_graph(ei, n)
            xg = x.clone().requires_grad_(True)
            block.set_x0(xg)
            z = block(xg, data)
            L = block.odefunc.multihead_att_layer
            params = [xg, L.Q.weight, L.K.weight, block.odefunc.alpha_train, block.odefunc.beta_train]
            results.append([z.detach()] + list(torch.autograd.grad((z * gz).sum(), params)))
    for base, other in ((results[0], results[1]), (results[2], results[3])):
        for a, b, name in zip(base, other, ('z', 'dx', 'dWq', 'dWk', 'dalpha', 'dbeta')):
            assert_close(b, a, 2e-5, name)


def test_an_empty_row_range_owns_no_column():
    """the column pass takes its ownership range from (row_lo, row_hi), and (0, 0) used to read as "own every column"
    (subtract g_j, accumulate dalpha / dbeta for every row) - what a compact-halo rank without rows would pass.
    Column pass of the same prepared row pass with range (0, n) and with (0, 0): the results differ by exactly the
    self term, dx_(0,0) - dx_(0,n) = alpha g."""
    import types
    import grand_b200
    from grand_b200 import partition as P
    from grand_b200.ops import _nl_workspace
    n, d, heads, adim = 640, 32, 4, 32
    func = make_func(d, heads, adim)
    L = func.multihead_att_layer
    ei = dev(rand_graph(n, 6 * n, 5))
    part = P.RowPartition(n, 1, 0)
    assert part.n_pad == n
    plan = grand_b200.GraphPlan(ei, n)
    gen = torch.Generator(device='cuda').manual_seed(4)
    x, g, x0 = (torch.randn(n, d, device='cuda', generator=gen) for _ in range(3))
    _, q = P.nl_local_forward(part, plan, x, x, L.Q.weight.detach(), L.Q.bias.detach(), L.K.weight.detach(),
                              func.alpha_train.detach(), func.beta_train.detach(), x0, heads, True)
    h = P.nl_local_vjp_rows(part, plan, x, q, L.Q.weight.detach(), L.K.weight.detach(), func.alpha_train.detach(),
                            func.beta_train.detach(), x0, heads, True, g, True)
    lib, ws = grand_b200._lib.load(), _nl_workspace(x.device)
    P._nl_adjoint_phase(lib, h, 2, ws)                                       # GB_NL_PHASE_COLS, owner of every column
    dx_all, dparams_all = h.dx.clone(), h.grads[-2:].clone()
    h.part = types.SimpleNamespace(lo=0, hi=0, slab=part.slab, n_pad=n)
    P._nl_adjoint_phase(lib, h, 2, ws)                                       # same edges, owner of none
    alpha = torch.sigmoid(func.alpha_train.detach())
    assert_close(h.dx - dx_all, alpha * g, 1e-6, 'self term')
    assert float((h.grads[-2:] - dparams_all).abs().max()) > 0               # <g, x0> and the -g_j x_j part are the owner's
