"""the CGNN model classes (mirror of src/CGNN.py:68-167) with `torch.spmm` replaced by the fused CSR
SpMM kernel.  The same bodies run on the CPU through the emulated C ABI and on the device (marked gpu); the checker is
the golden produced by the reference's own ODEFunc (tests/golden/ref_cgnn.npz) and the fp64 oracle composition."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import funcs, integrate
from helpers import assert_close

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def _func_against_the_golden(device):
    from grand_b200 import CGNN
    z = np.load(os.path.join(GOLD, 'ref_cgnn.npz'))
    T = lambda a, dt=torch.float32: torch.tensor(a, dtype=dt, device=device)
    n, h = int(z['n']), int(z['hidden'])
    ei = torch.tensor(z['edge_index'], device=device)
    for adj in ((ei, T(z['w']), n), torch.sparse_coo_tensor(ei, T(z['w']), (n, n))):
        func = CGNN.ODEFunc(2 * h, 2 * h, {'alpha': 0.9, 'hidden_dim': h}, adj).to(device)
        assert func.alpha_train.shape == (n,) and func.w.shape == (h, h) and func.d.shape == (h,)
        assert float(func.alpha_train.detach()[0]) == pytest.approx(0.9)
        with torch.no_grad():
            func.alpha_train.copy_(T(z['alpha_train']))
        func.x0 = T(z['x0'])
        x = T(z['x']).requires_grad_(True)
        f = func(0.0, x)
        dx, da = torch.autograd.grad(f, (x, func.alpha_train), T(z['cot']))
        assert func.nfe == 1
        assert_close(f, z['f'], 4e-6, 'f')
        assert_close(dx, z['dx'], 4e-6, 'dx')
        assert_close(da, z['dalpha'], 4e-6, 'dalpha')


def _model_against_the_oracle(device, n=120, e=700, feat=37, hidden=16, classes=5, T_end=1.5):
    from grand_b200 import CGNN
    from helpers import normalized_graph
    ei, w = normalized_graph(n, e, 5)
    ei, w = torch.as_tensor(ei), torch.as_tensor(w)
    opt = dict(num_feature=feat, hidden_dim=hidden, num_class=classes, alpha=0.8, input_dropout=0.0, dropout=0.0)
    torch.manual_seed(3)
    model = CGNN.CGNN(opt, (ei.to(device), w.to(device), n), T_end, device).to(device)
    model.eval()
    x = torch.randn(n, feat)
    y = torch.randint(0, classes, (n,))
    out = model(x.to(device))
    assert out.shape == (n, classes) and model.odeblock.nfe > 0 and model.getNFE() == 0     # the block zeroes the function's count (CGNN.py:110-111)
    loss = torch.nn.functional.cross_entropy(out, y.to(device))
    names = ['m1.weight', 'm1.bias', 'm2.weight', 'm2.bias', 'odeblock.odefunc.alpha_train']
    params = dict(model.named_parameters())
    assert set(params) == set(names) | {'odeblock.odefunc.w', 'odeblock.odefunc.d'}
    grads = torch.autograd.grad(loss, [params[k] for k in names])

    # fp64 oracle composition: encoder, augmentation, dopri5 at the solver defaults, relu, decoder (CGNN.py:146-167)
    P = {k: params[k].detach().double().cpu().requires_grad_(True) for k in names}
    h0 = x.double() @ P['m1.weight'].t() + P['m1.bias']
    h0 = torch.cat([h0, torch.zeros_like(h0)], dim=1)
    x0 = h0.detach()
    f = lambda t, v: funcs.cgnn_f(v, ei.long(), w.double(), P['odeblock.odefunc.alpha_train'], x0)
    zT = integrate.odeint(f, h0, [0.0, T_end], method='dopri5', rtol=1e-7, atol=1e-9)[1]
    logits = torch.relu(zT[:, :hidden]) @ P['m2.weight'].t() + P['m2.bias']
    ref_loss = torch.nn.functional.cross_entropy(logits, y)
    ref_grads = torch.autograd.grad(ref_loss, [P[k] for k in names])
    assert_close(out, logits, 2e-5, 'logits')
    for k, g, r in zip(names, grads, ref_grads):
        assert_close(g, r, 1e-4, k, floor=1e-4)


def test_cgnn_function_matches_the_reference_class(emulated_abi):
    _func_against_the_golden('cpu')
    assert 'gb_spmm_fwd' in emulated_abi.calls and 'gb_spmm_bwd' in emulated_abi.calls


def test_cgnn_model_matches_the_oracle_composition(emulated_abi):
    _model_against_the_oracle(torch.device('cpu'))


def test_get_sym_adj_feeds_the_function(emulated_abi):
    from grand_b200 import CGNN
    from helpers import rand_graph
    n = 40
    ei = torch.as_tensor(rand_graph(n, 200, 2, self_loops=False))
    data = types.SimpleNamespace(edge_index=ei, edge_attr=None, num_nodes=n, x=torch.randn(n, 3))
    edge_index, vals, nn_ = CGNN.get_sym_adj(data, {'self_loop_weight': 1.0}, torch.device('cpu'))
    assert nn_ == n and edge_index.shape[1] == vals.shape[0] >= ei.shape[1] + 1       # self loops added (fill value 1)
    func = CGNN.ODEFunc(8, 8, {'alpha': 0.5, 'hidden_dim': 4}, (edge_index, vals, n))
    func.x0 = torch.zeros(n, 8)
    x = torch.randn(n, 8)
    ref = funcs.cgnn_f(x.double(), edge_index.reshape(2, -1).long(), vals.reshape(-1).double(), func.alpha_train.detach().double(),
                       func.x0.double())
    assert_close(func(0.0, x), ref, 4e-6, 'f')


@pytest.mark.gpu
def test_gpu_cgnn_function_matches_the_reference_class():
    _func_against_the_golden('cuda')


@pytest.mark.gpu
def test_gpu_cgnn_model_matches_the_oracle_composition():
    _model_against_the_oracle(torch.device('cuda'))
    _model_against_the_oracle(torch.device('cuda'), n=2708, e=10556, feat=1433, hidden=16, classes=7, T_end=2.0)
