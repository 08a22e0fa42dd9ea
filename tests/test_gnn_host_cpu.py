"""CPU (emulated C ABI): the model wrappers around the hot path - GNN / GNNEarly (mirrors of the reference's
src/GNN.py, src/GNN_early.py, src/base_classes.py:137-186): encoder -> ODE block -> relu -> decoder against the
oracle, and the early-stopping test-time integrator end to end."""
import types

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import helpers
from helpers import assert_close
from oracle import funcs, graph as og, integrate as oint

T64 = lambda a: torch.tensor(np.asarray(a.detach().cpu() if torch.is_tensor(a) else a), dtype=torch.float64)
D = lambda t: t.to(helpers.device())          # 'cpu' under the emulated ABI, 'cuda' when the opt-in GPU tests call these

OPT = {'hidden_dim': 16, 'block': 'constant', 'function': 'laplacian', 'max_nfe': 10000, 'no_alpha_sigmoid': False,
       'add_source': True, 'augment': False, 'adjoint': False, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0,
       'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': 'rk4', 'step_size': 1.0, 'max_iters': 100,
       'adjoint_method': 'rk4', 'adjoint_step_size': 1.0, 'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 16,
       'attention_norm_idx': 0, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
       'square_plus': False, 'mix_features': False, 'multi_modal': False, 'kinetic_energy': None,
       'jacobian_norm2': None, 'total_deriv': None, 'directional_penalty': None, 'time': 3.0, 'use_mlp': False,
       'use_labels': False, 'fc_out': False, 'batch_norm': False, 'input_dropout': 0.0, 'dropout': 0.0,
       'max_test_steps': 100, 'earlystopxT': 2.0}


@pytest.fixture
def host(emulated_abi, monkeypatch):
    monkeypatch.setattr(helpers, 'DEVICE', 'cpu')
    return emulated_abi


@pytest.mark.parametrize('batched', [False, True])
@pytest.mark.parametrize('adjoint', [False, True])
def test_gnn_logits_and_gradients_match_the_oracle_composition(host, batched, adjoint):
    import grand_b200
    n, e, feats, classes = 120, 800, 20, 4
    torch.manual_seed(0)
    model = grand_b200.GNN(dict(OPT, adjoint=adjoint), classes, feats, helpers.device()).to(helpers.device())
    model.train()
    func = model.odeblock.odefunc
    with torch.no_grad():
        func.alpha_train.fill_(0.4)
        func.beta_train.fill_(0.2)
    ei = og.shaped_graph(n, e, 6)
    x = D(torch.randn(n, feats))
    labels = D(torch.randint(0, classes, (n,)))
    wrap = (lambda t: t[None]) if batched else (lambda t: t)
    logits = model(wrap(x), wrap(D(torch.tensor(ei))))
    assert logits.shape == ((1, n, classes) if batched else (n, classes))
    # getNFE adds odefunc.nfe and reg_odefunc.odefunc.nfe (base_classes.py:174-175): two different function
    # instances, as in the reference (the one inside reg_odefunc is never evaluated without regularisers)
    assert func.nfe == 12 and model.getNFE() == 12
    assert model.odeblock.reg_odefunc.odefunc is not func
    loss = F.cross_entropy(logits.reshape(n, classes), labels)
    params = [model.m1.weight, model.m1.bias, model.m2.weight, model.m2.bias, func.alpha_train, func.beta_train]
    grads = torch.autograd.grad(loss, params)
    # oracle composition in fp64
    ei_n, w = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    leaves = [T64(p).requires_grad_(True) for p in params]
    h0 = T64(x) @ leaves[0].t() + leaves[1]
    x0 = h0.detach()                                       # ODEblock.set_x0: detached clone (base_classes.py:53-55)

    class Ref(torch.nn.Module):
        def forward(self, t, y):
            return funcs.laplacian_f(y, torch.tensor(ei_n), T64(w), leaves[4], leaves[5], x0)

    solve = oint.odeint_adjoint if adjoint else oint.odeint
    kw = dict(adjoint_params=(leaves[4], leaves[5])) if adjoint else {}
    z = solve(Ref(), h0, [0., 3.0], 'rk4', 1.0, **kw)[1]
    ref_logits = F.relu(z) @ leaves[2].t() + leaves[3]
    ref = torch.autograd.grad(F.cross_entropy(ref_logits, labels.cpu()), leaves)
    assert_close(logits.reshape(n, classes), ref_logits, 1e-4, 'logits')
    for g, r, name in zip(grads, ref, ('dm1.w', 'dm1.b', 'dm2.w', 'dm2.b', 'dalpha', 'dbeta')):
        assert_close(g, r, 2e-4, name)
    model.resetNFE()
    assert model.getNFE() == 0


def test_gnn_early_runs_the_early_stop_integrator_at_test_time(host):
    import grand_b200
    n, e, feats, classes = 100, 700, 12, 3
    torch.manual_seed(1)
    model = grand_b200.GNNEarly(dict(OPT, time=2.0), classes, feats, helpers.device()).to(helpers.device())
    ei = D(torch.tensor(og.shaped_graph(n, e, 2)))
    x = D(torch.randn(n, feats))
    labels = D(torch.randint(0, classes, (n,)))
    idx = torch.randperm(n)
    masks = [torch.zeros(n, dtype=torch.bool) for _ in range(3)]
    for m, sl in zip(masks, (idx[:30], idx[30:60], idx[60:])):
        m[sl] = True
    masks = [D(m) for m in masks]
    data = types.SimpleNamespace(y=labels, train_mask=masks[0], val_mask=masks[1], test_mask=masks[2])
    model.set_solver_data(data)
    model.train()
    out_train = model(x, ei)                               # training: the ordinary integrator over [0, T]
    assert model.odeblock.test_integrator.solver is None and model.odeblock.odefunc.nfe == 8
    model.eval()
    model.resetNFE()
    with torch.no_grad():
        out = model(x, ei)                                 # test time: [0, earlystopxT * T] = 4 rk4 steps, decoded after each
    s = model.odeblock.test_integrator.solver
    assert out.shape == out_train.shape == (n, classes) and model.odeblock.odefunc.nfe == 16 and s.n_evaluations == 4
    assert 0 < s.best_time <= 4.0 and 0 <= s.best_test <= 1 and s.best_val > 0
    assert torch.equal(model.odeblock.test_integrator.m2_weight, model.m2.weight.data)


def test_base_gnn_option_branches_build_and_run(host):
    """beltrami (mx | mp), use_mlp, use_labels, fc_out, batch_norm, augment: shapes and hidden_dim bookkeeping"""
    import grand_b200
    n, e, feats, classes = 60, 400, 10, 3
    opt = dict(OPT, beltrami=True, feat_hidden_dim=8, pos_enc_dim=6, pos_enc_hidden_dim=4, use_mlp=True, use_labels=True,
               fc_out=True, batch_norm=True, augment=True, time=1.0)
    model = grand_b200.GNN(opt, classes, feats, helpers.device()).to(helpers.device())
    assert opt['hidden_dim'] == 8 + 4 + classes and model.hidden_dim == 15
    assert model.odeblock.odefunc.in_features == 30                     # augmented: 2 x hidden_dim
    ei = D(torch.tensor(og.shaped_graph(n, e, 3)))
    x = D(torch.cat([torch.randn(n, feats), F.one_hot(torch.randint(0, classes, (n,)), classes).float()], dim=1))
    model.train()
    out = model(x, ei, pos_encoding=D(torch.randn(n, 6)))
    assert out.shape == (n, classes)
    out.sum().backward()
    assert model.mx.weight.grad is not None and model.mp.weight.grad is not None and model.fc.weight.grad is not None


@pytest.mark.parametrize('function,block', [('laplacian', 'constant'), ('transformer', 'attention'), ('laplacian', 'mixed')])
def test_a_few_optimiser_steps_reduce_the_training_loss(host, function, block):
    """the whole training step of run_GNN.py (model.train(); loss = CE(out[train_mask]); backward; Adam step,
    src/graph_datasets/run_GNN.py:84-118) on a planted-partition graph: gradients have the right sign end to end"""
    import grand_b200
    n, classes, feats = 90, 3, 8
    g = torch.Generator().manual_seed(0)
    labels = torch.arange(n) % classes
    same = labels[:, None] == labels[None, :]
    prob = torch.where(same, torch.tensor(0.12), torch.tensor(0.01))
    adj = torch.rand(n, n, generator=g) < prob
    adj = (adj | adj.t()) & ~torch.eye(n, dtype=torch.bool)
    ei = D(adj.nonzero().t().contiguous())
    x = D(torch.randn(n, feats, generator=g) + 0.5 * F.one_hot(labels, feats).float())
    labels = D(labels)
    train_mask = D(torch.rand(n, generator=g) < 0.6)
    torch.manual_seed(0)
    opt = dict(OPT, function=function, block=block, adjoint=(block != 'mixed'), time=2.0, hidden_dim=16)
    model = grand_b200.GNN(opt, classes, feats, helpers.device()).to(helpers.device())
    optim = torch.optim.Adam(model.parameters(), lr=0.02)
    losses = []
    for _ in range(12):
        model.train()
        optim.zero_grad()
        out = model(x, ei)
        loss = F.cross_entropy(out[train_mask], labels[train_mask])
        loss.backward()
        optim.step()
        model.resetNFE()
        losses.append(float(loss.detach()))
    assert all(np.isfinite(losses)) and losses[-1] < 0.8 * losses[0], losses


def test_positional_encoding_files_round_trip(tmp_path):
    """graph_rewiring.apply_beltrami (src/graph_rewiring.py:277-313): same file name and pickle payloads as the
    reference - a plain tensor for GDC, {'data': ...} for the DW* types - plus the optional CSV side file"""
    import os
    import pickle
    import types
    import numpy as np
    from grand_b200 import graph_rewiring as gr
    data = types.SimpleNamespace(num_nodes=7)
    p = torch.randn(7, 5)
    opt = {'dataset': 'Cora', 'pos_enc_type': 'GDC', 'pos_enc_csv': True}
    with pytest.raises(NotImplementedError, match='Cora_GDC.pkl'):
        gr.apply_beltrami(data, opt, str(tmp_path))
    fname = gr.save_pos_encoding(p, opt, str(tmp_path))
    assert fname == str(tmp_path / 'pos_encodings' / 'Cora_GDC.pkl')
    with open(fname, 'rb') as f:
        assert torch.equal(pickle.load(f), p)                                 # what the reference's reader expects
    assert torch.equal(gr.apply_beltrami(data, opt, str(tmp_path)), p)
    table = np.loadtxt(str(tmp_path / 'pos_encodings' / 'Cora_GDC.csv'), delimiter=',')
    assert table.shape == (35, 3) and np.allclose(table[:, 2], p.reshape(-1).numpy(), atol=1e-6)
    opt = {'dataset': 'Photo', 'pos_enc_type': 'DW128'}
    with pytest.raises(FileNotFoundError):
        gr.apply_beltrami(data, opt, str(tmp_path))
    os.makedirs(tmp_path / 'pos_encodings', exist_ok=True)
    with open(tmp_path / 'pos_encodings' / 'Photo_DW128.pkl', 'wb') as f:     # as the reference's DeepWalk files are laid out
        pickle.dump({'data': p.double().numpy()}, f)
    got = gr.apply_beltrami(data, opt, str(tmp_path))
    assert got.dtype == torch.float32 and torch.allclose(got, p)
    with pytest.raises(ValueError, match='7 rows'):
        gr.apply_beltrami(types.SimpleNamespace(num_nodes=8), opt, str(tmp_path))
