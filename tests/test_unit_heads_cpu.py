"""CPU: the algebra behind the training path of attention_type cosine_sim / pearson.

The CUDA VJP kernels implement scaled_dot scores; grand_b200 differentiates the other two types as
scaled_dot of per-(node, head) unit vectors (`unit_heads`, node-level torch ops).  Here that reduction is
checked on the host: the scores are the reference's (oracle `edge_scores`), and the whole evaluation
reproduces the gradients the REFERENCE produced (tests/golden/ref_transformer_{cosine_sim,pearson}_*.npz,
written by oracle/make_golden.py from the unmodified reference modules)."""
import glob
import math
import os

import numpy as np
import pytest
import torch

from oracle import funcs

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def unit_heads():
    from grand_b200.function_transformer_attention import unit_heads as fn
    return fn


@pytest.mark.parametrize('att_type', ['cosine_sim', 'pearson'])
@pytest.mark.parametrize('heads,dk', [(1, 4), (2, 4), (4, 16), (8, 5)])
def test_scores_equal_the_reference_form(att_type, heads, dk):
    g = torch.Generator().manual_seed(heads * 100 + dk)
    n, e, a = 40, 300, heads * dk
    q = torch.randn(n, a, generator=g, dtype=torch.float64)
    k = torch.randn(n, a, generator=g, dtype=torch.float64)
    q[3] = 0.0                                                   # |q| < eps: clamped, score 0
    k[5] *= 1e-7
    ei = torch.randint(0, n, (2, e), generator=g)
    ref = funcs.edge_scores(q, k, ei, heads, att_type)
    u = unit_heads()(torch.cat([q, k], dim=1), heads, centre=att_type == 'pearson')
    got = funcs.edge_scores(u[:, :a].contiguous(), u[:, a:].contiguous(), ei, heads, 'scaled_dot')
    torch.testing.assert_close(got, ref, rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize('att_type', ['cosine_sim', 'pearson'])
def test_gradients_through_unit_heads_equal_autograd_of_the_reference_form(att_type):
    g = torch.Generator().manual_seed(3)
    n, e, heads, dk = 30, 200, 2, 8
    a = heads * dk
    ei = torch.randint(0, n, (2, e), generator=g)
    w = torch.randn(e, heads, generator=g, dtype=torch.float64)
    q0 = torch.randn(n, a, generator=g, dtype=torch.float64)
    k0 = torch.randn(n, a, generator=g, dtype=torch.float64)

    def grads(via_unit):
        q, k = q0.clone().requires_grad_(True), k0.clone().requires_grad_(True)
        if via_unit:
            u = unit_heads()(torch.cat([q, k], dim=1), heads, centre=att_type == 'pearson')
            s = funcs.edge_scores(u[:, :a], u[:, a:], ei, heads, 'scaled_dot')
        else:
            s = funcs.edge_scores(q, k, ei, heads, att_type)
        return torch.autograd.grad((s * w).sum(), (q, k))

    for got, ref in zip(grads(True), grads(False)):
        torch.testing.assert_close(got, ref, rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_transformer_cosine_sim_*.npz')) +
                                        glob.glob(os.path.join(GOLD, 'ref_transformer_pearson_*.npz'))))
def test_composite_evaluation_reproduces_the_reference_gradients(path):
    """projection -> unit_heads -> scaled_dot scores -> segment softmax -> head mean -> SpMM -> epilogue, i.e. the
    exact composition ODEFuncTransformerAtt runs on the GPU when a gradient is required, in fp64 on the host
    with the oracle standing in for the CUDA kernels; targets: f and all gradients of the reference itself"""
    z = np.load(path)
    T = lambda k: torch.tensor(z[k], dtype=torch.float64)
    att_type, heads, norm_idx = str(z['attention_type']), int(z['heads']), int(z['attention_norm_idx'])
    ei = torch.tensor(z['edge_index'])
    leaves = {k: T(k).requires_grad_(True) for k in ('x', 'wq', 'bq', 'wk', 'bk')}
    alpha = torch.tensor(float(z['alpha_train']), dtype=torch.float64, requires_grad=True)
    beta = torch.tensor(float(z['beta_train']), dtype=torch.float64, requires_grad=True)
    x = leaves['x']
    a = leaves['wq'].shape[0]
    qk = torch.cat([x @ leaves['wq'].t() + leaves['bq'], x @ leaves['wk'].t() + leaves['bk']], dim=1)
    u = unit_heads()(qk, heads, centre=att_type == 'pearson')
    prods = funcs.edge_scores(u[:, :a], u[:, a:], ei, heads, 'scaled_dot')
    att = funcs.segment_softmax(prods, ei[norm_idx], x.shape[0])
    f = funcs.laplacian_f(x, ei, att.mean(dim=1), alpha, beta, T('x0'))
    grads = torch.autograd.grad(f, [x, alpha, beta, leaves['wq'], leaves['bq'], leaves['wk'], leaves['bk']], T('gout'))
    pairs = [(prods, 'prods'), (att, 'attention'), (f, 'f')] + list(
        zip(grads, ('gx', 'galpha', 'gbeta', 'gwq', 'gbq', 'gwk', 'gbk')))
    for got, key in pairs:
        ref = T(key)                                             # the reference computed these in fp32
        assert torch.allclose(got, ref, rtol=5e-6, atol=5e-6 * float(ref.abs().max())), (key, att_type)


def test_unit_heads_handles_the_sqrt_dk_factor_and_layout():
    heads, dk, n = 2, 9, 5
    a = heads * dk
    qk = torch.randn(n, 2 * a, dtype=torch.float64)
    u = unit_heads()(qk, heads, centre=False).view(n, 2 * heads, dk)
    norms = u.norm(dim=2)
    torch.testing.assert_close(norms[:, :heads], torch.full((n, heads), math.sqrt(dk), dtype=torch.float64))
    torch.testing.assert_close(norms[:, heads:], torch.ones(n, heads, dtype=torch.float64))
