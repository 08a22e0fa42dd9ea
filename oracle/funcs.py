"""Oracle: the per-evaluation math of the two ODE functions (TEST INFRASTRUCTURE ONLY).

Plain torch on the CPU, differentiable (autograd supplies the gradient oracle), dtype chosen by the
caller: float64 is the truth the tolerances are judged against, float32 reproduces the reference's
own rounding.  Un-batched tensors: x [N,d], edge_index [2,E] (row = edge_index[0] = the node the
message is summed INTO, col = edge_index[1] = the node whose features are gathered; this is the
orientation of `sparse_coo_tensor(indices=[row, col]) @ x` in
`src/function_laplacian_diffusion.py:41-57` and of upstream's `torch_sparse.spmm`).

Follows:
  segment_softmax        src/utils.py:116-127
  attention_layer        src/function_transformer_attention.py:159-267 (line 249 replaced by the
                         per-edge dot product unless fork_literal=True)
  beltrami_attention_layer  src/function_transformer_attention.py:164-216 (beltrami + exp_kernel: product of a
                         feature kernel and a positional kernel).  PARITY UNPINNED for this branch: the fork
                         slices `x[:, f:label]`, which on its batched [B,N,d] layout cuts the NODE axis; restated
                         with the upstream intent (the feature axis), no reference golden exists.
  gat_attention_layer    src/function_GAT_attention.py:117-141 (literal: gather both endpoints' [H,dk] projections,
                         a * edge_h summed over 2 dk, leaky relu, softmax);  gat_f  :53-69 (mix_features=False)
  laplacian_f            src/function_laplacian_diffusion.py:39-77
  transformer_f          src/function_transformer_attention.py:20-59 (mix_features=False branch)
  cgnn_f                 src/CGNN.py:85-91 (per-node alpha, fixed adjacency, source x0 with unit weight)
  decode_eval            src/early_stop_solver.py:108-128 (evaluate) + :88-98 (test): relu, m2, masked arg-max counts
"""
import math
import torch


def segment_softmax(src, index, num_nodes):
    """src [E,H], index int64 [E] -> [E,H]; out = exp(s - max_g) / (sum_g exp(s - max_g) + 1e-16)."""
    e, h = src.shape
    idx = index[:, None].expand(e, h)
    mx = torch.zeros(num_nodes, h, dtype=src.dtype).scatter_reduce(0, idx, src, reduce='amax', include_self=False)
    out = (src - mx[index]).exp()
    den = torch.zeros(num_nodes, h, dtype=src.dtype).index_add_(0, index, out)
    return out / (den[index] + 1e-16)


def squareplus(src, index, num_nodes):
    """Intended behaviour of `src/utils.py:129-140` (the fork's call site passes 2 args to a 3-arg
    function and its amax is over the whole tensor; this keeps the fork's global-max form)."""
    out = src - src.amax()
    out = (out + torch.sqrt(out ** 2 + 4)) / 2
    den = torch.zeros(num_nodes, src.shape[1], dtype=src.dtype).index_add_(0, index, out)
    return out / (den[index] + 1e-16)


def edge_scores(q, k, edge_index, heads, attention_type='scaled_dot', fork_literal=False,
                output_var=None, lengthscale=None):
    """q,k [N,A] -> prods [E,H].  Head h owns the contiguous slice [h*dk,(h+1)*dk) of the A axis
    (`.view(N, h, d_k)` at function_transformer_attention.py:230-232)."""
    n, a = q.shape
    dk = a // heads
    row, col = edge_index[0], edge_index[1]
    src = q.view(n, heads, dk)[row]        # [E,H,dk]
    dst = k.view(n, heads, dk)[col]
    if attention_type == 'scaled_dot':
        if fork_literal:
            # line 249: matmul([H,E,dk],[H,dk,E]/sqrt(dk)).sum(dim=3): every edge sees the SUM of all dst rows
            return torch.matmul(src.permute(1, 0, 2), dst.permute(1, 2, 0) / math.sqrt(dk)).sum(dim=2).permute(1, 0)
        return (src * dst).sum(dim=2) / math.sqrt(dk)
    if attention_type == 'cosine_sim':
        return torch.nn.functional.cosine_similarity(src, dst, dim=2, eps=1e-5)
    if attention_type == 'pearson':
        src = src - src.mean(dim=2, keepdim=True)
        dst = dst - dst.mean(dim=2, keepdim=True)
        return torch.nn.functional.cosine_similarity(src, dst, dim=2, eps=1e-5)
    if attention_type == 'exp_kernel':
        return output_var ** 2 * torch.exp(-((src - dst) ** 2).sum(dim=2) / (2 * lengthscale ** 2))
    raise NotImplementedError(attention_type)


def attention_layer(x, edge_index, wq, bq, wk, bk, heads, attention_type='scaled_dot',
                    attention_norm_idx=0, edge_weight=None, reweight_attention=False,
                    square_plus=False, fork_literal=False, output_var=None, lengthscale=None):
    """SpGraphTransAttentionLayer.forward (non-beltrami-exp_kernel branch).  Returns
    (attention [E,H], prods [E,H]).  V is computed by the reference and never used when
    mix_features=False, so it is omitted (its gradient is None in the reference too)."""
    n = x.shape[0]
    q = x @ wq.t() + bq
    k = x @ wk.t() + bk
    prods = edge_scores(q, k, edge_index, heads, attention_type, fork_literal, output_var, lengthscale)
    if reweight_attention and edge_weight is not None:
        prods = prods * edge_weight[:, None]
    grp = edge_index[attention_norm_idx]
    if square_plus:
        return squareplus(prods, grp, n), prods
    return segment_softmax(prods, grp, n), prods


def beltrami_attention_layer(x, edge_index, w, heads, feat_dim, pos_dim, attention_norm_idx=0, edge_weight=None,
                             reweight_attention=False):
    """SpGraphTransAttentionLayer.forward, beltrami + exp_kernel branch (:164-216).  x = [features (feat_dim) |
    positional encoding (pos_dim) | labels ...]; `w` maps 'Qx.weight', 'Qx.bias', 'Kx.*', 'Qp.*', 'Kp.*',
    'output_var_x', 'lengthscale_x', 'output_var_p', 'lengthscale_p' to tensors.  Returns (attention, prods) [E,H]."""
    n = x.shape[0]
    lab = feat_dim + pos_dim
    p = x[:, feat_dim:lab]
    xf = torch.cat((x[:, :feat_dim], x[:, lab:]), dim=1)
    row, col = edge_index[0], edge_index[1]

    def heads_of(t):
        return t.view(n, heads, t.shape[1] // heads)

    src_x = heads_of(xf @ w['Qx.weight'].t() + w['Qx.bias'])[row]
    dst_x = heads_of(xf @ w['Kx.weight'].t() + w['Kx.bias'])[col]
    src_p = heads_of(p @ w['Qp.weight'].t() + w['Qp.bias'])[row]
    dst_p = heads_of(p @ w['Kp.weight'].t() + w['Kp.bias'])[col]
    prods = w['output_var_x'] ** 2 * torch.exp(-((src_x - dst_x) ** 2).sum(dim=2) / (2 * w['lengthscale_x'] ** 2)) \
        * w['output_var_p'] ** 2 * torch.exp(-((src_p - dst_p) ** 2).sum(dim=2) / (2 * w['lengthscale_p'] ** 2))
    if reweight_attention and edge_weight is not None:
        prods = prods * edge_weight[:, None]
    return segment_softmax(prods, edge_index[attention_norm_idx], n), prods


def gat_attention_layer(x, edge_index, w, a, heads, slope, attention_norm_idx=0):
    """SpGraphAttentionLayer.forward (function_GAT_attention.py:117-141): w [d, A], a [2 dk] (the reference keeps it
    as [1, 2 dk, 1, 1]).  Returns (attention [E,H], raw scores after the leaky relu [E,H])."""
    n = x.shape[0]
    wx = x @ w
    dk = wx.shape[1] // heads
    h = wx.view(n, heads, dk)
    edge_h = torch.cat((h[edge_index[0]], h[edge_index[1]]), dim=2)           # [E, H, 2 dk]
    edge_e = torch.nn.functional.leaky_relu((a.reshape(1, 1, 2 * dk) * edge_h).sum(dim=2), slope)
    return segment_softmax(edge_e, edge_index[attention_norm_idx], n), edge_e


def gat_f(x, edge_index, w, a, heads, slope, alpha_train, beta_train, x0, add_source=True, no_alpha_sigmoid=False,
          attention_norm_idx=0):
    """ODEFuncAtt.forward (function_GAT_attention.py:53-69, mix_features=False): mean over heads of the attention
    applied to x, same alpha / beta epilogue as the other functions."""
    att, _ = gat_attention_layer(x, edge_index, w, a, heads, slope, attention_norm_idx)
    return laplacian_f(x, edge_index, att.mean(dim=1), alpha_train, beta_train, x0, add_source, no_alpha_sigmoid)


def spmm(edge_index, values, x):
    """ax[i] = sum_{e: row_e = i} values_e * x[col_e]; duplicates add."""
    row, col = edge_index[0], edge_index[1]
    return torch.zeros_like(x).index_add_(0, row, values[:, None] * x[col])


def _alpha(alpha_train, no_alpha_sigmoid):
    return alpha_train if no_alpha_sigmoid else torch.sigmoid(alpha_train)


def laplacian_f(x, edge_index, values, alpha_train, beta_train, x0, add_source=True,
                no_alpha_sigmoid=False):
    """LaplacianODEFunc.forward. `values` [E] is edge_weight (constant block) or the head-mean of
    attention_weights (attention block, function_laplacian_diffusion.py:45-49)."""
    ax = spmm(edge_index, values, x)
    f = _alpha(alpha_train, no_alpha_sigmoid) * (ax - x)
    if add_source:
        f = f + beta_train * x0
    return f


def transformer_f(x, edge_index, wq, bq, wk, bk, heads, alpha_train, beta_train, x0,
                  add_source=True, no_alpha_sigmoid=False, **att_kwargs):
    """ODEFuncTransformerAtt.forward: attention recomputed from the current x, head-mean, A @ x
    (applied to x, not v), same alpha/beta epilogue."""
    att, _ = attention_layer(x, edge_index, wq, bq, wk, bk, heads, **att_kwargs)
    ax = spmm(edge_index, att.mean(dim=1), x)
    f = _alpha(alpha_train, no_alpha_sigmoid) * (ax - x)
    if add_source:
        f = f + beta_train * x0
    return f


def decode_eval(z, w2, b2, y, masks):
    """EarlyStopDopri5.evaluate + test (src/early_stop_solver.py:108-128, :88-98; same in EarlyStopRK4 :210-226):
    z [N, >= d] state after a solver step (an augmented state keeps only the first d = w2.shape[1] columns, :113-114),
    logits = relu(z) w2^T + b2, and for each mask (train, val, test): pred = logits[mask].max(-1)[1],
    correct = pred.eq(y[mask]).sum(), total = mask.sum().  Returns (logits, pred [N], [(correct, total)] per mask);
    the reference's accuracy is correct / total.  Un-batched restatement of the fork's [1, N, d] layout."""
    d = w2.shape[1]
    zz = torch.relu(z[:, :d])
    logits = zz @ w2.t()
    if b2 is not None:
        logits = logits + b2
    pred = logits.max(dim=-1)[1]
    counts = []
    for m in masks:
        if m is None:
            counts.append((0, 0))
            continue
        m = m.bool()
        counts.append((int(pred[m].eq(y[m]).sum().item()), int(m.sum().item())))
    return logits, pred, counts


def cgnn_f(x, edge_index, values, alpha_train, x0):
    """CGNN's ODEFunc.forward (src/CGNN.py:85-91): sigmoid(alpha_train)[:, None] * 0.5 * (A x - x) + x0 with
    alpha_train [N] (one per node) and A = the fixed normalised adjacency (torch.spmm(adj, x), :89)"""
    alph = torch.sigmoid(alpha_train).unsqueeze(dim=1)
    return alph * 0.5 * (spmm(edge_index, values, x) - x) + x0
