"""Generate tests/golden/*.npz by RUNNING THE REFERENCE'S OWN MODULES (TEST INFRASTRUCTURE ONLY).

Run in the build container, where /root/reference exists:

    python oracle/make_golden.py            # rewrites tests/golden/ref_*.npz

The reference's hot-path files (`src/utils.py`, `src/base_classes.py`,
`src/function_laplacian_diffusion.py`, `src/function_transformer_attention.py`) are imported
UNMODIFIED from /root/reference/src; the only additions are the import-only stubs in
`oracle/ref_stubs/` for third-party modules that cannot be installed here (torch_scatter,
torch_geometric.nn.conv.MessagePassing, data_multi).  Inputs are seeded; tensors use the fork's
batched layout with B=1 ([1,N,d], [1,2,E], [1,E], [1,E,H]) and are stored un-batched.

The fixtures pin `oracle/funcs.py` (and through it the CUDA path) to the reference:
  ref_softmax.npz            utils.softmax                               (utils.py:116-127)
  ref_squareplus.npz         utils.squareplus                            (utils.py:129-140)
  ref_laplacian_*.npz        LaplacianODEFunc.forward + autograd grads   (function_laplacian_diffusion.py:39-77)
  ref_transformer_*.npz      ODEFuncTransformerAtt.forward + grads for cosine_sim / pearson /
                             exp_kernel (per-edge in the fork) and scaled_dot (fork-literal E x E sum)
  ref_gat_n*.npz             SpGraphAttentionLayer.forward + ODEFuncAtt.forward (function_GAT_attention.py); FORWARD
                             values only: the reference multiplies with `sparse_att.permute(0, 3, 1, 2).to_dense()`
                             (:46-47) and torch 2.11's backward of a permuted sparse tensor does not un-permute the
                             indices (dense-then-permute gives the finite-difference value, sparse-permute does not),
                             so the reference's own gradients are wrong under this torch; the oracle's gradients are
                             pinned by torch.autograd.gradcheck instead (tests/test_oracle_golden.py)
  ref_cgnn.npz               CGNN.ODEFunc.forward (torch.spmm on coo2tensor's sparse tensor, per-node alpha) + autograd
                             grads w.r.t. x and alpha_train           (CGNN.py:68-91, :269-277); needs the import-only
                             stubs ray / torchdiffeq / data (the fork ships no src/data.py) / torch_geometric.utils.convert
  ref_k3.npz                 the K3 / x=ones known answer: attention == 0.5 exactly
                             (test/test_transformer_attention.py:98-106)
The GPU box has no /root/reference: tests only read the .npz files.
"""
import os
import sys
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = '/root/reference/src'
GOLD = os.path.join(HERE, '..', 'tests', 'golden')


def load_reference():
    """Import the reference modules (unmodified) with the stubs on sys.path."""
    if not os.path.isdir(REF_SRC):
        raise RuntimeError('/root/reference is not present (only the build container has it)')
    for p in (REF_SRC, os.path.join(HERE, 'ref_stubs')):    # stubs end up first on sys.path
        if p not in sys.path:
            sys.path.insert(0, p)
    import utils as ref_utils                                    # noqa: E402
    import function_laplacian_diffusion as ref_lap               # noqa: E402
    import function_transformer_attention as ref_att             # noqa: E402
    return ref_utils, ref_lap, ref_att


BASE_OPT = {'self_loop_weight': 1, 'leaky_relu_slope': 0.2, 'heads': 2, 'attention_norm_idx': 0,
            'add_source': True, 'hidden_dim': 8, 'attention_dim': 8, 'block': 'constant',
            'function': 'laplacian', 'max_nfe': 1000, 'no_alpha_sigmoid': False,
            'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
            'square_plus': False, 'mix_features': False, 'multi_modal': False,
            'feat_hidden_dim': 6, 'pos_enc_hidden_dim': 2}


def toy_graph(n, e, seed, self_loops=True):
    g = torch.Generator().manual_seed(seed)
    row = torch.randint(0, n, (e,), generator=g)
    col = torch.randint(0, n, (e,), generator=g)
    if self_loops:
        row = torch.cat([row, torch.arange(n)])
        col = torch.cat([col, torch.arange(n)])
    return torch.stack([row, col]).long()


def np_(t):
    return t.detach().cpu().numpy()


def gen_softmax(ref_utils):
    g = torch.Generator().manual_seed(11)
    n, e, h = 9, 40, 3
    src = torch.randn(1, e, h, generator=g) * 3
    index = torch.randint(0, n, (1, e), generator=g)
    out = ref_utils.softmax(src, index, n)
    np.savez(os.path.join(GOLD, 'ref_softmax.npz'), src=np_(src[0]), index=np_(index[0]), n=n, out=np_(out[0]))


def gen_squareplus(ref_utils):
    """utils.squareplus itself (utils.py:129-140) on the fork's batched layout; its only call site
    (function_transformer_attention.py:264) passes two arguments and cannot run, the function can"""
    g = torch.Generator().manual_seed(12)
    n, e, h = 9, 40, 3
    src = torch.randn(1, e, h, generator=g) * 3
    index = torch.randint(0, n, (1, e), generator=g)
    out = ref_utils.squareplus(src, index, n)
    np.savez(os.path.join(GOLD, 'ref_squareplus.npz'), src=np_(src[0]), index=np_(index[0]), n=n, out=np_(out[0]))


def gen_laplacian(ref_lap):
    dev = torch.device('cpu')
    for name, block in (('constant', 'constant'), ('attention', 'attention')):
        for no_sig in (False, True):
            opt = dict(BASE_OPT, block=block, no_alpha_sigmoid=no_sig)
            n, d, h = 12, opt['hidden_dim'], 3
            ei = toy_graph(n, 30, 5)
            e = ei.shape[1]
            g = torch.Generator().manual_seed(7)
            func = ref_lap.LaplacianODEFunc(d, d, opt, dev)
            x = torch.randn(1, n, d, generator=g).requires_grad_(True)
            x0 = torch.randn(1, n, d, generator=g)
            w = torch.rand(1, e, generator=g)
            att = torch.rand(1, e, h, generator=g)
            gout = torch.randn(1, n, d, generator=g)
            with torch.no_grad():
                func.alpha_train.fill_(0.3)
                func.beta_train.fill_(-0.2)
            func.edge_index, func.edge_weight, func.attention_weights, func.x0 = ei[None], w, att, x0
            f = func(torch.tensor(0.5), x)
            gx, ga, gb = torch.autograd.grad(f, (x, func.alpha_train, func.beta_train), gout)
            np.savez(os.path.join(GOLD, f'ref_laplacian_{name}_{"raw" if no_sig else "sig"}.npz'),
                     edge_index=np_(ei), x=np_(x[0]), x0=np_(x0[0]), edge_weight=np_(w[0]), attention=np_(att[0]),
                     alpha_train=0.3, beta_train=-0.2, block=block, no_alpha_sigmoid=no_sig, gout=np_(gout[0]),
                     f=np_(f[0]), gx=np_(gx[0]), galpha=np_(ga), gbeta=np_(gb), nfe=func.nfe)


def gen_transformer(ref_att):
    dev = torch.device('cpu')
    for att_type in ('cosine_sim', 'pearson', 'exp_kernel', 'scaled_dot'):
        for norm_idx in (0, 1):
            opt = dict(BASE_OPT, block='attention', function='transformer', attention_type=att_type,
                       attention_norm_idx=norm_idx, heads=2, attention_dim=8)
            n, d = 10, opt['hidden_dim']
            ei = toy_graph(n, 24, 9)
            g = torch.Generator().manual_seed(13)
            func = ref_att.ODEFuncTransformerAtt(d, d, opt, dev)
            layer = func.multihead_att_layer
            with torch.no_grad():
                for lin in (layer.Q, layer.K):
                    lin.weight.copy_(torch.randn(lin.weight.shape, generator=g) / d ** 0.5)
                    lin.bias.copy_(torch.randn(lin.bias.shape, generator=g) * 0.1)
                func.alpha_train.fill_(-0.4)
                func.beta_train.fill_(0.7)
                if att_type == 'exp_kernel':
                    layer.output_var.fill_(1.3)
                    layer.lengthscale.fill_(0.8)
            x = torch.randn(1, n, d, generator=g).requires_grad_(True)
            x0 = torch.randn(1, n, d, generator=g)
            gout = torch.randn(1, n, d, generator=g)
            func.edge_index, func.x0 = ei[None], x0
            func.y = None                         # what ODEblock.reset_graph_data does (base_classes.py:90)
            attention, (_, prods) = layer(x, ei[None])
            f = func(torch.tensor(0.0), x)
            params = [func.alpha_train, func.beta_train, layer.Q.weight, layer.Q.bias, layer.K.weight, layer.K.bias]
            if att_type == 'exp_kernel':
                params += [layer.output_var, layer.lengthscale]
            grads = torch.autograd.grad(f, [x] + params, gout)
            extra = {}
            if att_type == 'exp_kernel':
                extra = dict(output_var=1.3, lengthscale=0.8, goutput_var=np_(grads[7]), glengthscale=np_(grads[8]))
            np.savez(os.path.join(GOLD, f'ref_transformer_{att_type}_n{norm_idx}.npz'),
                     edge_index=np_(ei), x=np_(x[0]), x0=np_(x0[0]), gout=np_(gout[0]), heads=2,
                     wq=np_(layer.Q.weight), bq=np_(layer.Q.bias), wk=np_(layer.K.weight), bk=np_(layer.K.bias),
                     alpha_train=-0.4, beta_train=0.7, attention_type=att_type, attention_norm_idx=norm_idx,
                     attention=np_(attention[0]), prods=np_(prods[0]), f=np_(f[0]),
                     gx=np_(grads[0][0]), galpha=np_(grads[1]), gbeta=np_(grads[2]), gwq=np_(grads[3]),
                     gbq=np_(grads[4]), gwk=np_(grads[5]), gbk=np_(grads[6]),
                     fork_literal=(att_type == 'scaled_dot'), **extra)


def gen_gat():
    """function_GAT_attention.py, imported unmodified like the other modules (load_reference() put the paths in place)"""
    import function_GAT_attention as ref_gat                     # noqa: E402
    dev = torch.device('cpu')
    for norm_idx in (0, 1):
        opt = dict(BASE_OPT, block='constant', function='GAT', attention_norm_idx=norm_idx, heads=2, attention_dim=8)
        n, d = 11, opt['hidden_dim']
        ei = toy_graph(n, 28, 15)
        torch.manual_seed(21 + norm_idx)                          # xavier init of W, Wout, a
        func = ref_gat.ODEFuncAtt(d, d, opt, dev)
        layer = func.multihead_att_layer
        g = torch.Generator().manual_seed(17)
        with torch.no_grad():
            func.alpha_train.fill_(0.25)
            func.beta_train.fill_(-0.3)
        x = torch.randn(1, n, d, generator=g).requires_grad_(True)
        x0 = torch.randn(1, n, d, generator=g)
        gout = torch.randn(1, n, d, generator=g)
        func.edge_index, func.x0 = ei[None], x0
        func.y = None
        attention, wx = layer(x, ei[None])
        f = func(torch.tensor(0.0), x)
        np.savez(os.path.join(GOLD, f'ref_gat_n{norm_idx}.npz'),
                 edge_index=np_(ei), x=np_(x[0]), x0=np_(x0[0]), gout=np_(gout[0]), heads=2, slope=opt['leaky_relu_slope'],
                 w=np_(layer.W), a=np_(layer.a), alpha_train=0.25, beta_train=-0.3, attention_norm_idx=norm_idx,
                 attention=np_(attention[0]), wx=np_(wx[0]), f=np_(f[0]), nfe=func.nfe)


def gen_k3(ref_att):
    """test/test_transformer_attention.py:98-106 with the reference's default init."""
    opt = dict(BASE_OPT, heads=2, attention_dim=32, hidden_dim=6)
    edge1 = torch.tensor([[[0, 0, 1, 1, 2, 2], [1, 2, 0, 2, 0, 1]]])
    x1 = torch.ones((1, 3, 2))
    torch.manual_seed(3)
    layer = ref_att.SpGraphTransAttentionLayer(2, 2, opt, torch.device('cpu'), concat=True)
    attention, _ = layer(x1, edge1)
    assert torch.all(torch.eq(attention, 0.5 * torch.ones((1, 6, 2))))
    np.savez(os.path.join(GOLD, 'ref_k3.npz'), edge_index=np_(edge1[0]), x=np_(x1[0]),
             wq=np_(layer.Q.weight), bq=np_(layer.Q.bias), wk=np_(layer.K.weight), bk=np_(layer.K.bias),
             attention=np_(attention[0]), heads=2)


def gen_cgnn():
    import scipy.sparse
    import CGNN as ref_cgnn                                      # noqa: E402  (the reference's src/CGNN.py, unmodified)
    n, e, h = 23, 90, 4
    ei = toy_graph(n, e, 31)                                     # duplicates + self loops
    g = torch.Generator().manual_seed(32)
    w = torch.rand(ei.shape[1], generator=g) * 0.3
    coo = scipy.sparse.coo_matrix((np_(w), (np_(ei[0]), np_(ei[1]))), shape=(n, n))
    adj = ref_cgnn.coo2tensor(coo, torch.device('cpu'))
    opt = {'alpha': 0.9, 'hidden_dim': h}
    func = ref_cgnn.ODEFunc(2 * h, 2 * h, opt, adj)
    with torch.no_grad():
        func.alpha_train.copy_(torch.randn(n, generator=g))
    x = torch.randn(n, 2 * h, generator=g, requires_grad=True)
    func.x0 = torch.randn(n, 2 * h, generator=g)
    f = func(0.0, x)
    cot = torch.randn(n, 2 * h, generator=g)
    dx, dalpha = torch.autograd.grad(f, [x, func.alpha_train], cot)
    np.savez(os.path.join(GOLD, 'ref_cgnn.npz'), edge_index=np_(ei), w=np_(w), n=n, hidden=h, alpha_train=np_(func.alpha_train),
             x=np_(x), x0=np_(func.x0), f=np_(f), cot=np_(cot), dx=np_(dx), dalpha=np_(dalpha))


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref_utils, ref_lap, ref_att = load_reference()
    if sys.argv[1:] == ['cgnn']:                                 # only this fixture
        gen_cgnn()
        return
    if sys.argv[1:] == ['gat']:                                  # only the GAT fixtures (leaves the others untouched)
        gen_gat()
        return
    gen_gat()
    gen_softmax(ref_utils)
    gen_squareplus(ref_utils)
    gen_laplacian(ref_lap)
    gen_transformer(ref_att)
    gen_k3(ref_att)
    gen_cgnn()
    print('wrote', sorted(os.listdir(GOLD)))


if __name__ == '__main__':
    main()
