"""GPU: the ODE blocks drive the CUDA ODE functions through the integrators (forward and
odeint_adjoint backward); end-of-integration states and gradients vs the fp64 oracle at fixed step
size, rtol 1e-4 (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

from oracle import funcs, graph as og, integrate as oint
from helpers import assert_close, dev

pytestmark = pytest.mark.gpu

OPT = {'hidden_dim': 16, 'block': 'constant', 'function': 'laplacian', 'max_nfe': 10000, 'no_alpha_sigmoid': False,
       'add_source': True, 'augment': False, 'adjoint': False, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0,
       'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': 'rk4', 'step_size': 1.0, 'max_iters': 100,
       'adjoint_method': 'rk4', 'adjoint_step_size': 1.0, 'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 16,
       'attention_norm_idx': 0, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
       'square_plus': False, 'mix_features': False, 'multi_modal': False, 'kinetic_energy': None,
       'jacobian_norm2': None, 'total_deriv': None, 'directional_penalty': None}


def build(opt, n, e, T, seed=0):
    import grand_b200
    block_cls, func_cls = grand_b200.set_block(opt), grand_b200.set_function(opt)
    block = block_cls(func_cls, [], opt, torch.device('cuda'), t=torch.tensor([0., T])).cuda()
    ei = og.shaped_graph(n, e, seed)
    data = grand_b200.GraphData()
    data.new_graph(dev(ei), n)
    return block, data, ei


def rand_qk(layer, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for lin in (layer.Q, layer.K):
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / lin.weight.shape[1] ** 0.5).cuda())
            lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=g)).cuda())


def t64(a):
    return torch.tensor(np.asarray(a.detach().cpu() if torch.is_tensor(a) else a), dtype=torch.float64)


@pytest.mark.parametrize('adjoint', [False, True])
@pytest.mark.parametrize('method', ['rk4', 'euler'])
@pytest.mark.parametrize('n,e', [(400, 3000), (3000, 20000)])    # 3000 x 16 elements: the one-pass stage combination
def test_constant_block_laplacian_rk4(adjoint, method, n, e):      # (integrators._FUSE_MIN) is on the autograd path
    d, T = 16, 3.0
    opt = dict(OPT, adjoint=adjoint, method=method, adjoint_method=method)
    block, data, ei = build(opt, n, e, T)
    block.train()
    func = block.odefunc
    with torch.no_grad():
        func.alpha_train.fill_(0.4)
        func.beta_train.fill_(0.2)
    x = torch.randn(n, d, device='cuda', requires_grad=True)
    gout = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    z = block(x, data)
    steps = {'rk4': 4, 'euler': 1}[method] * 3
    assert z.shape == x.shape and func.nfe == steps
    func.nfe = 0
    gx, ga, gb = torch.autograd.grad(z, (x, func.alpha_train, func.beta_train), gout)
    if adjoint:
        assert func.nfe == steps                                             # one f per adjoint dynamics eval
    # oracle: same graph build, fp64, its own integrator
    ei_n, w = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    xo = t64(x).requires_grad_(True)
    a, b = t64(0.4).requires_grad_(True), t64(0.2).requires_grad_(True)
    x0 = t64(x)
    f = lambda t, y: funcs.laplacian_f(y, torch.tensor(ei_n), t64(w), a, b, x0)
    # adjoint gradients are optimise-then-discretise: compare with the oracle's adjoint, not with backprop
    if adjoint:
        zo = oint.odeint_adjoint(f, xo, [0., T], method, 1.0, adjoint_params=(a, b))[1]
    else:
        zo = oint.odeint(f, xo, [0., T], method, 1.0)[1]
    ro = torch.autograd.grad(zo, (xo, a, b), t64(gout))
    assert_close(z, zo, 1e-4, 'z(T)')
    for g, r, name in zip((gx, ga, gb), ro, ('gx', 'galpha', 'gbeta')):
        assert_close(g, r, 1e-4, name)


@pytest.mark.parametrize('adjoint', [False, True])
@pytest.mark.parametrize('n,e', [(300, 2000), (2000, 14000)])    # 2000 x 16: above integrators._FUSE_MIN
def test_attention_block_transformer_function(adjoint, n, e):
    d, T = 16, 2.0
    opt = dict(OPT, block='attention', function='transformer', adjoint=adjoint)
    block, data, ei = build(opt, n, e, T, seed=2)
    block.train()
    func = block.odefunc
    rand_qk(func.multihead_att_layer, 1)
    rand_qk(block.multihead_att_layer, 2)
    with torch.no_grad():
        func.alpha_train.fill_(-0.3)
        func.beta_train.fill_(0.5)
    x = torch.randn(n, d, device='cuda', requires_grad=True)
    gout = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    z = block(x, data)
    L = func.multihead_att_layer
    params = [x, func.alpha_train, func.beta_train, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias]
    grads = torch.autograd.grad(z, params, gout)
    ei_n, _ = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    leaves = [t64(p).requires_grad_(True) for p in params]
    x0 = t64(x)
    f = lambda t, y: funcs.transformer_f(y, torch.tensor(ei_n), leaves[3], leaves[4], leaves[5], leaves[6], 4,
                                         leaves[1], leaves[2], x0)
    if adjoint:
        zo = oint.odeint_adjoint(f, leaves[0], [0., T], 'rk4', 1.0, adjoint_params=tuple(leaves[1:]))[1]
    else:
        zo = oint.odeint(f, leaves[0], [0., T], 'rk4', 1.0)[1]
    ro = torch.autograd.grad(zo, leaves, t64(gout))
    assert_close(z, zo, 1e-4, 'z(T)')
    bias_scale = max(float(ro[4].abs().max()), float(ro[6].abs().max()))
    for g, r, name in zip(grads, ro, ('gx', 'galpha', 'gbeta', 'gwq', 'gbq', 'gwk', 'gbk')):
        assert_close(g, r, 1e-4, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


def test_attention_block_transformer_function_evaluates_the_block_attention_lazily():
    """block_transformer_attention.py:24-32 computes attention_weights at t = 0 although function 'transformer'
    never reads it; here the same values appear on the first read, and the block forward launches no edge softmax"""
    import grand_b200
    n, e, d, T = 300, 2000, 16, 2.0
    opt = dict(OPT, block='attention', function='transformer', adjoint=False)
    block, data, ei = build(opt, n, e, T, seed=2)
    rand_qk(block.odefunc.multihead_att_layer, 1)
    rand_qk(block.multihead_att_layer, 2)
    x = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    with grand_b200._lib.Profiler() as prof:
        with torch.no_grad():
            block(x, data)
        kernels = prof.results()
        assert 'nl_fwd' in kernels and 'edge_softmax_fwd' not in kernels, sorted(kernels)
        lazy = block.odefunc.attention_weights                       # first read: evaluated now
        assert 'edge_softmax_fwd' in prof.results()
    with torch.no_grad():
        eager = block.get_attention_weights(x)
    assert lazy is block.odefunc.attention_weights                   # evaluated once
    assert torch.equal(lazy, eager) and lazy.shape[-1] == 4
    block.odefunc.attention_weights = None                           # plain assignment still works
    assert block.odefunc.attention_weights is None


@pytest.mark.parametrize('function,blockname,method', [('transformer', 'attention', 'rk4'), ('laplacian', 'constant', 'rk4'),
                                                       ('transformer', 'attention', 'midpoint'), ('laplacian', 'constant', 'euler')])
def test_stage_combinations_fused_into_the_kernel_epilogues_are_bit_identical(function, blockname, method):
    """north_star (a): y + h sum c_j k_j of the next Runge-Kutta stage is written by the kernel that produces f
    (gb_stage epilogue) - no stage_axpy launches in a no-grad solve, and the same bits as the separate pass"""
    import grand_b200
    from grand_b200 import integrators
    n, e, d, T = 3000, 24000, 64, 2.0
    opt = dict(OPT, block=blockname, function=function, adjoint=False, hidden_dim=d, attention_dim=32, method=method)
    block, data, ei = build(opt, n, e, T, seed=4)
    if function == 'transformer':
        rand_qk(block.odefunc.multihead_att_layer, 1)
    with torch.no_grad():
        block.odefunc.alpha_train.fill_(0.4)
        block.odefunc.beta_train.fill_(0.2)
    x = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    out = {}
    for fuse in (True, False):
        integrators._FUSE_STAGES = fuse
        try:
            with grand_b200._lib.Profiler() as prof, torch.no_grad():
                out[fuse] = block(x, data)
                kernels = prof.results()
        finally:
            integrators._FUSE_STAGES = True
        stages = {'rk4': 4, 'midpoint': 2, 'euler': 1}[method]
        assert kernels.get('stage_axpy', {'launches': 0})['launches'] == (0 if fuse else 2 * stages), kernels
    assert torch.equal(out[True], out[False])


def test_attention_block_laplacian_function_trains_block_attention_without_adjoint():
    """GRAND-nl as configured by best_params (block=attention, function=laplacian): attention computed
    once from the block input; without adjoint its Q/K receive gradients."""
    n, e, d, T = 250, 1500, 16, 2.0
    opt = dict(OPT, block='attention', function='laplacian')
    block, data, ei = build(opt, n, e, T, seed=4)
    block.train()
    rand_qk(block.multihead_att_layer, 3)
    func = block.odefunc
    x = torch.randn(n, d, device='cuda', requires_grad=True)
    gout = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    z = block(x, data)
    B = block.multihead_att_layer
    params = [x, B.Q.weight, B.Q.bias, B.K.weight, B.K.bias]
    grads = torch.autograd.grad(z, params, gout)
    ei_n, _ = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    leaves = [t64(p).requires_grad_(True) for p in params]
    att, _ = funcs.attention_layer(leaves[0], torch.tensor(ei_n), leaves[1], leaves[2], leaves[3], leaves[4], 4)
    x0 = t64(x)
    a, b = t64(0.0), t64(0.0)
    f = lambda t, y: funcs.laplacian_f(y, torch.tensor(ei_n), att.mean(1), a, b, x0)
    zo = oint.odeint(f, leaves[0], [0., T], 'rk4', 1.0)[1]
    ro = torch.autograd.grad(zo, leaves, t64(gout))
    assert_close(z, zo, 1e-4, 'z(T)')
    bias_scale = max(float(ro[2].abs().max()), float(ro[4].abs().max()))
    for g, r, name in zip(grads, ro, ('gx', 'gwq', 'gbq', 'gwk', 'gbk')):
        assert_close(g, r, 1e-4, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


def test_dopri5_block_runs_and_matches_oracle_loosely():
    n, e, d, T = 300, 2000, 16, 4.0
    opt = dict(OPT, method='dopri5', tol_scale=100.0)
    block, data, ei = build(opt, n, e, T, seed=5)
    block.eval()
    x = torch.randn(n, d, device='cuda')
    block.set_x0(x)
    with torch.no_grad():
        z = block(x, data)
    ei_n, w = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    x0 = t64(x)
    f = lambda t, y: funcs.laplacian_f(y, torch.tensor(ei_n), t64(w), t64(0.0), t64(0.0), x0)
    zo = oint.odeint(f, x0, [0., T], 'dopri5', 1.0, rtol=1e-9, atol=1e-11)[1]
    assert_close(z, zo, 1e-3, 'z(T) dopri5')
    assert block.odefunc.nfe > 6


def test_regularisers_raise():
    import grand_b200
    with pytest.raises(NotImplementedError):
        grand_b200.base_classes.create_regularization_fns(dict(OPT, kinetic_energy=0.1))
