"""CPU: the algebra of grand_b200.ops.exp_kernel_score_vjp (the backward of attention_type exp_kernel) with
an index_add standing in for the multi-head SpMM kernels, against autograd of the oracle's score and against
the gradients the REFERENCE produced (tests/golden/ref_transformer_exp_kernel_*.npz)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import funcs, graph as og

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def csr_layout(ei, n):
    """oracle plan of a [2,E] graph: everything exp_kernel_score_vjp reads from a GraphPlan"""
    rowptr, colidx, eperm = og.coo_to_csr(ei, n)
    colptr, rowidx, tperm = og.csr_to_csc(rowptr, colidx, n)
    rows_csr = np.repeat(np.arange(n), np.diff(rowptr))
    T = lambda a: torch.tensor(np.asarray(a).astype(np.int64))
    return T(rowptr), T(colptr), T(tperm), T(rows_csr), T(colidx), T(eperm)


def vjp_on_host(ds_csr, prods_csr, q, k, kparams, lay, heads):
    from grand_b200.ops import exp_kernel_score_vjp
    rowptr, colptr, tperm, rows, cols, _ = lay
    n, a = q.shape
    dk = a // heads

    def mh(c):                                                   # what gb_edge_score_bwd computes (without 1/sqrt(dk))
        cw = c[:, :, None]
        ck = torch.zeros(n, heads, dk, dtype=q.dtype).index_add_(0, rows, cw * k.view(n, heads, dk)[cols])
        cq = torch.zeros(n, heads, dk, dtype=q.dtype).index_add_(0, cols, cw * q.view(n, heads, dk)[rows])
        return ck.view(n, a), cq.view(n, a)

    return exp_kernel_score_vjp(ds_csr, prods_csr, q, k, kparams, rowptr.int(), colptr.int(), tperm.int(), heads, mh)


@pytest.mark.parametrize('heads,dk,n,e', [(1, 4, 20, 90), (2, 8, 50, 400), (4, 16, 64, 700)])
def test_vjp_equals_autograd_of_the_score(heads, dk, n, e):
    g = torch.Generator().manual_seed(heads + dk)
    a = heads * dk
    ei = torch.randint(0, n, (2, e), generator=g).numpy()
    lay = csr_layout(ei, n)
    rows, cols = lay[3], lay[4]
    q = (0.5 * torch.randn(n, a, generator=g, dtype=torch.float64)).requires_grad_(True)
    k = (0.5 * torch.randn(n, a, generator=g, dtype=torch.float64)).requires_grad_(True)
    kp = torch.tensor([1.3, 0.8], dtype=torch.float64, requires_grad=True)
    ds = torch.randn(e, heads, generator=g, dtype=torch.float64)
    s = funcs.edge_scores(q, k, torch.stack([rows, cols]), heads, 'exp_kernel', output_var=kp[0], lengthscale=kp[1])
    rq, rk, rkp = torch.autograd.grad((s * ds).sum(), (q, k, kp))
    dq, dk_, dkp = vjp_on_host(ds, s.detach(), q.detach(), k.detach(), kp.detach(), lay, heads)
    torch.testing.assert_close(dq, rq, rtol=1e-10, atol=1e-12)
    torch.testing.assert_close(dk_, rk, rtol=1e-10, atol=1e-12)
    torch.testing.assert_close(dkp, rkp, rtol=1e-10, atol=1e-12)


def test_underflowed_scores_give_zero_not_nan():
    n, heads, dk = 4, 1, 2
    ei = np.array([[0, 1, 2, 3], [1, 2, 3, 0]])
    lay = csr_layout(ei, n)
    q = torch.zeros(n, 2, dtype=torch.float32)
    k = torch.full((n, 2), 100.0)
    prods = torch.zeros(4, 1)                                     # exp(-1e4 / ..) underflows to exactly 0 in fp32
    dq, dk_, dkp = vjp_on_host(torch.ones(4, 1), prods, q, k, torch.tensor([1.0, 1.0]), lay, heads)
    assert torch.isfinite(dq).all() and torch.isfinite(dk_).all() and torch.isfinite(dkp).all()
    assert float(dq.abs().max()) == 0.0 and float(dkp.abs().max()) == 0.0


@pytest.mark.parametrize('path', sorted(glob.glob(os.path.join(GOLD, 'ref_transformer_exp_kernel_*.npz'))))
def test_composite_backward_reproduces_the_reference_gradients(path):
    """forward by the oracle, backward by hand along the product's route (SpMM VJP -> SDDMM -> softmax backward ->
    exp_kernel_score_vjp -> projection transposes), compared with the reference's own gradients"""
    z = np.load(path)
    T = lambda k: torch.tensor(z[k], dtype=torch.float64)
    heads, norm_idx = int(z['heads']), int(z['attention_norm_idx'])
    ei = z['edge_index']
    n = z['x'].shape[0]
    lay = csr_layout(ei, n)
    rowptr, colptr, tperm, rows, cols, eperm = lay
    ei_csr = torch.stack([rows, cols])
    x, wq, bq, wk, bk, x0, gout = (T(k) for k in ('x', 'wq', 'bq', 'wk', 'bk', 'x0', 'gout'))
    kp = torch.tensor([float(z['output_var']), float(z['lengthscale'])], dtype=torch.float64)
    alpha = torch.sigmoid(torch.tensor(float(z['alpha_train']), dtype=torch.float64))
    q, k = x @ wq.t() + bq, x @ wk.t() + bk
    prods = funcs.edge_scores(q, k, ei_csr, heads, 'exp_kernel', output_var=kp[0], lengthscale=kp[1])
    grp = ei_csr[norm_idx]
    att = funcs.segment_softmax(prods, grp, n)
    # hand-written backward, CSR-slot order throughout
    dabar = alpha * (gout[rows] * x[cols]).sum(dim=1)                                  # SDDMM
    datt = (dabar / heads)[:, None].expand(-1, heads)
    inner = torch.zeros(n, heads, dtype=torch.float64).index_add_(0, grp, att * datt)
    ds = att * (datt - inner[grp])                                                     # softmax backward
    dq, dk_, dkp = vjp_on_host(ds, prods, q, k, kp, lay, heads)
    gx = alpha * (torch.zeros_like(x).index_add_(0, cols, att.mean(dim=1)[:, None] * gout[rows]) - gout)
    gx = gx + dq @ wq + dk_ @ wk
    checks = [(gx, 'gx'), (dq.t() @ x, 'gwq'), (dq.sum(0), 'gbq'), (dk_.t() @ x, 'gwk'), (dk_.sum(0), 'gbk'),
              (dkp[0], 'goutput_var'), (dkp[1], 'glengthscale')]
    for got, key in checks:
        ref = T(key).reshape(got.shape)
        assert torch.allclose(got, ref, rtol=2e-5, atol=2e-5 * max(float(ref.abs().max()), 1e-3)), \
            (key, float((got - ref).abs().max()), float(ref.abs().max()))
