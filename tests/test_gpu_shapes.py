"""GPU: per-evaluation parity at the C2 (PubMed-shaped, 19 717 nodes, d = 128) and C3 (ogbn-arxiv-shaped, 169 343
nodes, d = 96) sizes of BASELINE.json - whole graph against the fp64 oracle: f and every gradient of GRAND-l
(LaplacianODEFunc) and GRAND-nl (ODEFuncTransformerAtt), with the norm-wise gate of the other tests (1e-5 / 1e-4)
AND the per-row gate of SURVEY.md 8d ("hard part 1"); the three error figures are printed (pytest -s) and written to
gpurun_out/parity_report.json when that directory exists."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import funcs, graph as og
from helpers import assert_close, assert_rows_close, error_report, dev

pytestmark = pytest.mark.gpu
SHAPES = {'c2': (19717, 88648, 128), 'c3': (169343, 1166243, 96)}
HEADS, ADIM = 4, 64
REPORT = {}


def _t64(a):
    return torch.tensor(np.asarray(a.detach().cpu() if torch.is_tensor(a) else a), dtype=torch.float64)


def _note(key, got, ref):
    REPORT[key] = {k: float(f'{v:.3e}') for k, v in error_report(got, ref).items()}
    print(key, REPORT[key])
    out = os.path.join(os.path.dirname(__file__), '..', 'gpurun_out')
    if os.path.isdir(out):
        with open(os.path.join(out, 'parity_report.json'), 'w') as fh:
            json.dump(REPORT, fh, indent=1, sort_keys=True)


@pytest.fixture(scope='module', params=sorted(SHAPES))
def shape(request):
    n, e, d = SHAPES[request.param]
    ei, w = og.reset_graph_data(og.shaped_graph(n, e // 2 * 2, 1234), None, n, 'rw', 1.0)
    rng = np.random.default_rng(11)
    x, x0, g = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    return dict(name=request.param, n=n, d=d, ei=ei, w=np.asarray(w, dtype=np.float32), x=x, x0=x0, g=g)


def test_grand_l_evaluation_and_gradients(shape):
    import grand_b200
    n, d = shape['n'], shape['d']
    opt = {'hidden_dim': d, 'block': 'constant', 'max_nfe': 10 ** 6, 'no_alpha_sigmoid': False, 'add_source': True}
    f = grand_b200.LaplacianODEFunc(d, d, opt, torch.device('cuda')).cuda()
    with torch.no_grad():
        f.alpha_train.fill_(0.3)
        f.beta_train.fill_(0.1)
    f.edge_index, f.edge_weight, f.x0 = dev(shape['ei']), dev(shape['w']), dev(shape['x0'])
    xg = dev(shape['x']).requires_grad_(True)
    out = f(0.0, xg)
    grads = torch.autograd.grad(out, [xg, f.alpha_train, f.beta_train], dev(shape['g']))
    leaves = [_t64(shape['x']).requires_grad_(True), _t64(0.3).requires_grad_(True), _t64(0.1).requires_grad_(True)]
    ref = funcs.laplacian_f(leaves[0], torch.tensor(shape['ei']), _t64(shape['w']), leaves[1], leaves[2], _t64(shape['x0']))
    rg = torch.autograd.grad(ref, leaves, _t64(shape['g']))
    assert_close(out, ref, 1e-5, 'f')
    _note(f"{shape['name']}/GRAND-l/f", out, ref)
    _note(f"{shape['name']}/GRAND-l/dx", grads[0], rg[0])
    assert_rows_close(out, ref, 1e-5, 'f (per row)')
    for got, want, name in zip(grads, rg, ('dx', 'dalpha', 'dbeta')):
        assert_close(got, want, 1e-4, name)
    assert_rows_close(grads[0], rg[0], 1e-4, 'dx (per row)')


@pytest.mark.parametrize('impl', ['3', '2'])
def test_grand_nl_evaluation_and_gradients(shape, impl, monkeypatch):
    """impl 3: k-gather row kernels (default for d > attention_dim), 2: single-gather kernels"""
    import grand_b200
    monkeypatch.setenv('GRANDB200_NL_IMPL', impl)
    n, d = shape['n'], shape['d']
    opt = {'leaky_relu_slope': 0.2, 'heads': HEADS, 'attention_dim': ADIM, 'attention_norm_idx': 0, 'add_source': True,
           'hidden_dim': d, 'block': 'attention', 'function': 'transformer', 'max_nfe': 10 ** 6,
           'no_alpha_sigmoid': False, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
           'square_plus': False, 'mix_features': False, 'multi_modal': False}
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, torch.device('cuda')).cuda()
    L = func.multihead_att_layer
    gen = torch.Generator().manual_seed(3)
    with torch.no_grad():
        for lin in (L.Q, L.K):
            lin.weight.copy_((torch.randn(lin.weight.shape, generator=gen) / d ** 0.5).cuda())
            lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=gen)).cuda())
        func.alpha_train.fill_(0.3)
        func.beta_train.fill_(0.1)
    func.edge_index, func.x0 = dev(shape['ei']), dev(shape['x0'])
    xg = dev(shape['x']).requires_grad_(True)
    params = [xg, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]
    out = func(0.0, xg)
    grads = torch.autograd.grad(out, params, dev(shape['g']))
    leaves = [_t64(p).requires_grad_(True) for p in params]
    ref = funcs.transformer_f(leaves[0], torch.tensor(shape['ei']), leaves[1], leaves[2], leaves[3], leaves[4], HEADS,
                              leaves[5], leaves[6], _t64(shape['x0']))
    rg = torch.autograd.grad(ref, leaves, _t64(shape['g']))
    tag = f"{shape['name']}/GRAND-nl/impl{impl}"
    _note(tag + '/f', out, ref)
    _note(tag + '/dx', grads[0], rg[0])
    assert_close(out, ref, 1e-5, 'f')
    assert_rows_close(out, ref, 2e-5, 'f (per row)')
    bias_scale = max(float(rg[2].abs().max()), float(rg[4].abs().max()))
    for got, want, name in zip(grads, rg, ('dx', 'dwq', 'dbq', 'dwk', 'dbk', 'dalpha', 'dbeta')):
        assert_close(got, want, 1e-4, name, floor=bias_scale if name in ('dbq', 'dbk') else 0.0)
    assert_rows_close(grads[0], rg[0], 1e-4, 'dx (per row)')
    f2, vx, _ = func.eval_vjp(0.0, dev(shape['x']), dev(shape['g']), params[1:])     # the adjoint solver's entry
    assert_close(f2, ref, 1e-5, 'f of eval_vjp')
    assert_close(vx, rg[0], 1e-4, 'dx of eval_vjp')
