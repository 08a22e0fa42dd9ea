"""autograd.Functions over the C ABI: the three composable pieces of one ODE-function evaluation.

  QKProjFn        q|k = x [Wq;Wk]^T + [bq;bk]                (function_transformer_attention.py:224-225)
  EdgeSoftmaxFn   att[E,H], abar[E] from q,k over the plan    (…:240-266 + utils.py:116-127)
  DiffusionFn     f = alpha (A x - x) + beta x0                (function_laplacian_diffusion.py:60-77)

Each `backward` is a hand-written VJP (the reference has none: it relies on torch autograd through
its dense restatement, and `odeint_adjoint` calls autograd.grad on the same forward).  Double
backward (`create_graph=True`, used only by regularized_ODE_function.py:39-75) is not supported and
raises.  Every tensor must live on a CUDA device: there is no CPU path.
"""
import os

import torch
from torch.autograd.function import once_differentiable

from . import _lib


def _f32c(t):
    if t.dtype != torch.float32:
        raise TypeError(f'grand_b200 computes in float32 (reference dtype); got {t.dtype}')
    return t if t.is_contiguous() else t.contiguous()


class QKProjFn(torch.autograd.Function):
    """x [N,d]; wq,wk [A,d]; bq,bk [A]  ->  qk [N,2A] (q = qk[:, :A], k = qk[:, A:])"""

    @staticmethod
    def forward(ctx, x, wq, bq, wk, bk):
        lib = _lib.load()
        _lib.require_cuda(x, wq, bq, wk, bk)
        x = _f32c(x)
        n, d = x.shape
        a = wq.shape[0]
        w = torch.cat([wq, wk], dim=0).contiguous()
        b = torch.cat([bq, bk], dim=0).contiguous()
        out = torch.empty(n, 2 * a, dtype=torch.float32, device=x.device)
        _lib.check(lib.gb_linear_fwd(x.data_ptr(), d, n, d, w.data_ptr(), b.data_ptr(), 2 * a, out.data_ptr(),
                                     2 * a, _lib.stream()), 'gb_linear_fwd')
        ctx.save_for_backward(x, w)
        ctx.a = a
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dqk):
        lib = _lib.load()
        x, w = ctx.saved_tensors
        n, d = x.shape
        a = ctx.a
        dqk = _f32c(dqk)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _lib.check(lib.gb_linear_bwd_x(dqk.data_ptr(), 2 * a, n, 2 * a, w.data_ptr(), d, dx.data_ptr(), d, 0,
                                           _lib.stream()), 'gb_linear_bwd_x')
        if any(ctx.needs_input_grad[1:]):
            dw = torch.empty(2 * a, d, dtype=torch.float32, device=x.device)
            db = torch.empty(2 * a, dtype=torch.float32, device=x.device)
            ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n, d, 2 * a)), dtype=torch.uint8, device=x.device)
            _lib.check(lib.gb_linear_bwd_w(dqk.data_ptr(), 2 * a, x.data_ptr(), d, n, d, 2 * a, dw.data_ptr(),
                                           db.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
            return dx, dw[:a], db[:a], dw[a:], db[a:]
        return dx, None, None, None, None


class LinearFn(torch.autograd.Function):
    """y = x w^T + b for x [..., d], w [nout, d] (torch's Linear.weight layout), b [nout] or None: the encoder /
    decoder layers of GNN.forward (src/GNN.py:30-38, :66-74; src/base_classes.py:149-160) on the library's own GEMMs -
    tcgen05 3xTF32 where the shape qualifies (K slices for the [N, 1433] x [1433, d] encoder), exact-fp32 CUDA-core
    tiles otherwise (the [d] -> [classes] decoder)."""

    @staticmethod
    def forward(ctx, x, w, b):
        lib = _lib.load()
        _lib.require_cuda(x, w, b)
        shape = x.shape
        x2 = _f32c(x.reshape(-1, shape[-1]))
        w = _f32c(w)
        n, d = x2.shape
        nout = w.shape[0]
        out = torch.empty(n, nout, dtype=torch.float32, device=x.device)
        if n > 0:
            _lib.check(lib.gb_linear_fwd(x2.data_ptr(), d, n, d, w.data_ptr(), _lib.ptr(None if b is None else _f32c(b)),
                                         nout, out.data_ptr(), nout, _lib.stream()), 'gb_linear_fwd')
        ctx.save_for_backward(x2, w)
        ctx.shape, ctx.has_bias = shape, b is not None
        return out.reshape(*shape[:-1], nout)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        lib = _lib.load()
        x2, w = ctx.saved_tensors
        n, d = x2.shape
        nout = w.shape[0]
        dy = _f32c(dy.reshape(-1, nout))
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x2)
            if n > 0:
                _lib.check(lib.gb_linear_bwd_x(dy.data_ptr(), nout, n, nout, w.data_ptr(), d, dx.data_ptr(), d, 0,
                                               _lib.stream()), 'gb_linear_bwd_x')
            dx = dx.reshape(ctx.shape)
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dw = torch.zeros(nout, d, dtype=torch.float32, device=x2.device)
            db = torch.zeros(nout, dtype=torch.float32, device=x2.device)
            if n > 0:
                ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n, d, nout)), dtype=torch.uint8, device=x2.device)
                _lib.check(lib.gb_linear_bwd_w(dy.data_ptr(), nout, x2.data_ptr(), d, n, d, nout, dw.data_ptr(),
                                               db.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
            if not ctx.has_bias:
                db = None
        return dx, dw, db


class Linear(torch.nn.Linear):
    """torch.nn.Linear with the same parameters (`weight`, `bias`: reference checkpoints load unchanged) whose forward
    and backward run on the library's GEMM kernels instead of cuBLAS"""

    def forward(self, x):
        return LinearFn.apply(x, self.weight, self.bias)


def decode_eval(z, w2, b2, y, masks, want_logits=False, want_pred=False):
    """EarlyStopInt's per-step evaluation (src/early_stop_solver.py:108-128 + :88-98) in one kernel: z [N, >= d] (or
    the fork's [1, N, .]) state, w2 [C, d], b2 [C]; y [N] int64; masks = (train, val, test) bool [N] (entries may be
    None).  Returns (counts int32[6] ON THE DEVICE: correct x3, total x3, logits or None, pred or None)."""
    lib = _lib.load()
    _lib.require_cuda(z, w2, b2, y, *masks)
    z2 = _f32c(z.reshape(-1, z.shape[-1]))
    w2 = _f32c(w2.reshape(-1, w2.shape[-1]))
    n, ldz = z2.shape
    classes, d = w2.shape
    if d > ldz:
        raise ValueError(f'decoder expects {d} features, the state has {ldz}')
    y = y.reshape(-1)
    if y.dtype != torch.int64 or y.numel() != n:
        raise TypeError(f'labels must be int64 [N={n}]; got {y.dtype} {tuple(y.shape)}')
    y = y.contiguous()
    ms = []
    for m in masks:
        if m is not None:
            m = m.reshape(-1)
            if m.dtype != torch.bool or m.numel() != n:
                raise TypeError(f'masks must be bool [N={n}]; got {m.dtype} {tuple(m.shape)}')
            m = m.contiguous().view(torch.uint8)
        ms.append(m)
    counts = torch.empty(6, dtype=torch.int32, device=z.device)
    logits = torch.empty(n, classes, dtype=torch.float32, device=z.device) if want_logits else None
    pred = torch.empty(n, dtype=torch.int64, device=z.device) if want_pred else None
    _lib.check(lib.gb_decode_eval(z2.data_ptr(), ldz, n, d, w2.data_ptr(), d, _lib.ptr(None if b2 is None else _f32c(b2)),
                                  classes, y.data_ptr(), _lib.ptr(ms[0]), _lib.ptr(ms[1]), _lib.ptr(ms[2]),
                                  counts.data_ptr(), _lib.ptr(logits), classes, _lib.ptr(pred), _lib.stream()),
               'gb_decode_eval')
    return counts, logits, pred


def _segment_sums(vals, ptr):
    """[E,H] values laid out segment after segment, ptr [n+1] -> [n,H] sums.  Prefix sums in fp64 (deterministic on
    every backend; a difference of two prefixes loses ~1e-16 * E relative to the running total)"""
    if vals.requires_grad:                       # autograd path (opt['square_plus']): no out= argument
        cs = torch.cat([torch.zeros(1, vals.shape[1], dtype=torch.float64, device=vals.device),
                        torch.cumsum(vals.double(), dim=0)], dim=0)
    else:
        cs = torch.zeros(vals.shape[0] + 1, vals.shape[1], dtype=torch.float64, device=vals.device)
        torch.cumsum(vals.double(), dim=0, out=cs[1:])
    ptr = ptr.long()
    return (cs[ptr[1:]] - cs[ptr[:-1]]).to(vals.dtype)


def exp_kernel_score_vjp(ds, prods, q, k, kparams, rowptr, colptr, tperm, heads, mh):
    """VJP of the exp_kernel score  s[e,h] = ov^2 exp(-|q[row_e,h] - k[col_e,h]|^2 / (2 l^2))
    (function_transformer_attention.py:247-248) given ds = dL/ds [E,H] (CSR-slot order) and the scores themselves:

        c_eh  = ds_eh s_eh / l^2
        dq[i,h] = sum_{e in row i} c_eh k[col_e,h] - q[i,h] sum_{e in row i} c_eh
        dk[j,h] = sum_{e in col j} c_eh q[row_e,h] - k[j,h] sum_{e in col j} c_eh
        d ov = 2/ov sum ds s            d l = sum ds s |q-k|^2 / l^3,  |q-k|^2 = -2 l^2 log(s / ov^2)

    `mh(c)` returns the two gathered sums ([N,A] each) - on the GPU the multi-head SpMM kernels behind
    gb_edge_score_bwd (scaled_dot form), in the CPU test of this algebra an index_add.  Everything else is
    elementwise / prefix-sum torch work on [E,H] and [N,A] tensors."""
    n, a = q.shape
    dk = a // heads
    ov, ls = kparams[0], kparams[1]
    w = ds * prods
    c = w / (ls * ls)
    ck, cq = mh(c)
    rsum = _segment_sums(c, rowptr)
    csum = _segment_sums(c[tperm.long()], colptr)
    dq = ck.view(n, heads, dk) - q.view(n, heads, dk) * rsum[:, :, None]
    dkk = cq.view(n, heads, dk) - k.view(n, heads, dk) * csum[:, :, None]
    logt = torch.where(prods > 0, torch.log(prods / (ov * ov)), torch.zeros_like(prods))
    d_ov = w.double().sum() * 2.0 / ov.double()
    d_ls = (w.double() * (-2.0 * logt.double())).sum() / ls.double()
    return dq.reshape(n, a), dkk.reshape(n, a), torch.stack([d_ov, d_ls]).to(ds.dtype)


def _score_vjp(plan, qk, heads, dk, att_type, ds, prods, kparams, edge_w, want_kparams):
    """VJP of the per-edge scores w.r.t. [q | k] (and the exp_kernel parameters) given ds = dL/d(raw score) [E,H] in
    CSR-slot order.  `prods` are the scores as the forward kernel stored them, i.e. AFTER the optional
    reweight_attention factor `edge_w`; the exp_kernel VJP needs the raw ones (prods / edge_w)."""
    lib = _lib.load()
    n, a2 = qk.shape
    a = a2 // 2
    if att_type not in (_lib.GB_ATT['scaled_dot'], _lib.GB_ATT['exp_kernel']):
        raise NotImplementedError('cosine_sim / pearson are differentiated as scaled_dot of unit vectors '
                                  '(function_transformer_attention.unit_heads); call through the layer')
    dqk = torch.empty_like(qk)

    def mh(c, scale=1.0):
        """sum_{e in row i} c_eh k[col_e,h] -> dqk[:, :A], sum_{e in col j} c_eh q[row_e,h] -> dqk[:, A:]; the
        kernels multiply by 1/sqrt(dk) (scaled_dot VJP), `scale` undoes it where the caller needs plain sums"""
        c = _f32c(c * scale if scale != 1.0 else c)
        _lib.check(lib.gb_edge_score_bwd(plan.c, c.data_ptr(), qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2, heads,
                                         dk, _lib.GB_ATT['scaled_dot'], None, None, dqk.data_ptr(), a2,
                                         dqk.data_ptr() + 4 * a, a2, None, None, _lib.stream()),
                   'gb_edge_score_bwd')
        return dqk[:, :a], dqk[:, a:]

    if att_type == _lib.GB_ATT['scaled_dot']:
        mh(ds)
        return dqk, None
    raw = prods
    if edge_w is not None:                       # s_raw = prods / w; w == 0 -> ds == 0 there, any finite value does
        w = edge_w.reshape(-1, 1)
        raw = torch.where(w != 0, prods / torch.where(w != 0, w, torch.ones_like(w)), torch.zeros_like(prods))
    dq, dkk, dkparams = exp_kernel_score_vjp(ds, raw, qk[:, :a], qk[:, a:], kparams, plan.rowptr, plan.colptr,
                                             plan.tperm, heads, lambda c: mh(c, float(dk) ** 0.5))
    return torch.cat([dq, dkk], dim=1), (dkparams if want_kparams else None)


class EdgeScoreFn(torch.autograd.Function):
    """qk [N,2A] -> prods [E,H] (CSR-slot order): the per-edge scores alone, differentiable - for normalisations
    other than the fused segment softmax (opt['square_plus'], utils.squareplus)."""

    @staticmethod
    def forward(ctx, qk, plan, heads, att_type, edge_w, kparams):
        lib = _lib.load()
        _lib.require_cuda(qk)
        qk = _f32c(qk)
        n, a2 = qk.shape
        a = a2 // 2
        dk = a // heads
        f32 = dict(dtype=torch.float32, device=qk.device)
        att, abar, prods = torch.empty(plan.e, heads, **f32), torch.empty(plan.e, **f32), torch.empty(plan.e, heads, **f32)
        _lib.check(lib.gb_edge_softmax_fwd(plan.c, 0, qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2, heads, dk, att_type,
                                           _lib.ptr(edge_w), _lib.ptr(kparams), att.data_ptr(), abar.data_ptr(),
                                           prods.data_ptr(), _lib.stream()), 'gb_edge_softmax_fwd')
        exp_kernel = att_type == _lib.GB_ATT['exp_kernel']
        ctx.save_for_backward(qk, prods if exp_kernel else None, kparams if exp_kernel else None)
        ctx.plan, ctx.heads, ctx.dk, ctx.att_type, ctx.edge_w = plan, heads, dk, att_type, edge_w
        return prods

    @staticmethod
    @once_differentiable
    def backward(ctx, dprods):
        qk, prods, kparams = ctx.saved_tensors
        ds = _f32c(dprods)
        if ctx.edge_w is not None:
            ds = ds * ctx.edge_w.reshape(-1, 1)
        dqk, dkparams = _score_vjp(ctx.plan, qk, ctx.heads, ctx.dk, ctx.att_type, ds, prods, kparams, ctx.edge_w,
                                   ctx.needs_input_grad[5])
        return dqk, None, None, None, None, dkparams


def squareplus_csr(prods, plan, norm_idx):
    """utils.squareplus (utils.py:129-140; upstream semantics - the fork's call site at
    function_transformer_attention.py:264 passes two arguments to this three-argument function):
    out = s - max(s over ALL edges and heads); sp = (out + sqrt(out^2 + 4)) / 2; attention = sp / (group sum + 1e-16),
    groups = rows (norm_idx 0) or columns (1) per head.  prods in CSR-slot order -> (att [E,H], abar [E]).
    [E,H]-sized torch glue (autograd-friendly; group sums are fp64 prefix sums: deterministic)."""
    out = prods - prods.amax()                   # differentiated through, as the src.amax of utils.py:132 is
    sp = (out + torch.sqrt(out * out + 4.0)) * 0.5
    counts = (plan.rowptr[1:] - plan.rowptr[:-1]).long()
    if norm_idx == 0:
        den = torch.repeat_interleave(_segment_sums(sp, plan.rowptr), counts, dim=0)
    else:
        den = _segment_sums(sp[plan.tperm.long()], plan.colptr)[plan.colidx.long()]
    att = sp / (den + 1e-16)
    return att, att.mean(dim=1)


class EdgeSoftmaxFn(torch.autograd.Function):
    """qk [N,2A] -> (att [E,H], abar [E], prods [E,H]) in CSR-slot order."""

    @staticmethod
    def forward(ctx, qk, plan, heads, att_type, norm_idx, edge_w, kparams):
        lib = _lib.load()
        _lib.require_cuda(qk)
        qk = _f32c(qk)
        n, a2 = qk.shape
        a = a2 // 2
        dk = a // heads
        att = torch.empty(plan.e, heads, dtype=torch.float32, device=qk.device)
        abar = torch.empty(plan.e, dtype=torch.float32, device=qk.device)
        prods = torch.empty(plan.e, heads, dtype=torch.float32, device=qk.device)
        if plan.long_threshold == 256 and lib.gb_att_fused_supported(qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2,
                                                                     heads, dk, None, 0, 0):
            _lib.check(lib.gb_att_fwd_fused(plan.c, norm_idx, qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2, heads, dk,
                                            att_type, _lib.ptr(edge_w), _lib.ptr(kparams), att.data_ptr(),
                                            abar.data_ptr(), prods.data_ptr(), 1, None, 0, None, 0, None, 0, None, 0,
                                            None, 0, None, _lib.stream()), 'gb_att_fwd_fused')
        else:
            _lib.check(lib.gb_edge_softmax_fwd(plan.c, norm_idx, qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2, heads,
                                               dk, att_type, _lib.ptr(edge_w), _lib.ptr(kparams), att.data_ptr(),
                                               abar.data_ptr(), prods.data_ptr(), _lib.stream()),
                       'gb_edge_softmax_fwd')
        exp_kernel = att_type == _lib.GB_ATT['exp_kernel']
        ctx.save_for_backward(qk, att, prods if exp_kernel else None, kparams if exp_kernel else None)
        ctx.plan, ctx.heads, ctx.dk, ctx.att_type, ctx.norm_idx, ctx.edge_w = plan, heads, dk, att_type, norm_idx, edge_w
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(prods)
        return att, abar, prods

    @staticmethod
    @once_differentiable
    def backward(ctx, datt, dabar, _dprods):
        lib = _lib.load()
        qk, att, prods, kparams = ctx.saved_tensors
        plan, heads, dk = ctx.plan, ctx.heads, ctx.dk
        n, a2 = qk.shape
        a = a2 // 2
        if datt is None and dabar is None:
            return (None,) * 7
        datt = None if datt is None else _f32c(datt)
        dabar = None if dabar is None else _f32c(dabar)
        ds = torch.empty(plan.e, heads, dtype=torch.float32, device=qk.device)
        _lib.check(lib.gb_edge_softmax_bwd(plan.c, ctx.norm_idx, att.data_ptr(), _lib.ptr(dabar), _lib.ptr(datt), heads,
                                           _lib.ptr(ctx.edge_w), ds.data_ptr(), _lib.stream()), 'gb_edge_softmax_bwd')
        dqk, dkparams = _score_vjp(plan, qk, heads, dk, ctx.att_type, ds, prods, kparams, ctx.edge_w,
                                   ctx.needs_input_grad[6])
        return dqk, None, None, None, None, None, dkparams


class GatSoftmaxFn(torch.autograd.Function):
    """s [N,2H] = [s_src | s_dst] per-node score halves -> (att [E,H], abar [E], prods [E,H]) in CSR-slot order:
    prods = leaky_relu(s_src[row] + s_dst[col]), att = segment softmax (function_GAT_attention.py:136-141).
    Backward: softmax VJP kernel (gb_edge_softmax_bwd), leaky_relu' elementwise, then the two segment sums of the
    [E,H] score gradients (prefix sums, deterministic) - per-node SCALARS, so there is no row gather to do."""

    @staticmethod
    def forward(ctx, s, plan, heads, slope, norm_idx):
        lib = _lib.load()
        _lib.require_cuda(s)
        s = _f32c(s)
        if s.shape[1] != 2 * heads:
            raise ValueError(f'GatSoftmaxFn: s must be [N, {2 * heads}], got {tuple(s.shape)}')
        att = torch.empty(plan.e, heads, dtype=torch.float32, device=s.device)
        abar = torch.empty(plan.e, dtype=torch.float32, device=s.device)
        prods = torch.empty(plan.e, heads, dtype=torch.float32, device=s.device)
        _lib.check(lib.gb_gat_softmax_fwd(plan.c, norm_idx, s.data_ptr(), s.data_ptr() + 4 * heads, 2 * heads, heads,
                                          float(slope), att.data_ptr(), abar.data_ptr(), prods.data_ptr(),
                                          _lib.stream()), 'gb_gat_softmax_fwd')
        ctx.save_for_backward(att, prods)
        ctx.plan, ctx.heads, ctx.slope, ctx.norm_idx = plan, heads, float(slope), norm_idx
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(prods)
        return att, abar, prods

    @staticmethod
    @once_differentiable
    def backward(ctx, datt, dabar, _dprods):
        lib = _lib.load()
        att, prods = ctx.saved_tensors
        plan, heads = ctx.plan, ctx.heads
        if datt is None and dabar is None:
            return (None,) * 5
        datt = None if datt is None else _f32c(datt)
        dabar = None if dabar is None else _f32c(dabar)
        ds = torch.empty(plan.e, heads, dtype=torch.float32, device=att.device)
        _lib.check(lib.gb_edge_softmax_bwd(plan.c, ctx.norm_idx, att.data_ptr(), _lib.ptr(dabar), _lib.ptr(datt), heads,
                                           None, ds.data_ptr(), _lib.stream()), 'gb_edge_softmax_bwd')
        dz = torch.where(prods > 0, ds, ds * ctx.slope)
        d_src = _segment_sums(dz, plan.rowptr)
        d_dst = _segment_sums(dz[plan.tperm.long()], plan.colptr)
        return torch.cat([d_src, d_dst], dim=1), None, None, None, None


class SegmentSoftmaxFn(torch.autograd.Function):
    """utils.softmax on ready-made scores: src [E,H], ptr [n+1] / perm [E] from the sorted group ids (utils.softmax
    builds them) -> out [E,H].  Backward: ds = out * (dout - sum_g out * dout), the group sums as prefix sums in
    sorted order (deterministic)."""

    @staticmethod
    def forward(ctx, src, ptr, perm, index, long_groups):
        lib = _lib.load()
        _lib.require_cuda(src)
        src = _f32c(src)
        e, h = src.shape
        out = torch.empty_like(src)
        if e:
            _lib.check(lib.gb_segment_softmax_fwd(ptr.data_ptr(), perm.data_ptr(), ptr.numel() - 1, h, src.data_ptr(),
                                                  out.data_ptr(), None, _lib.ptr(long_groups) if long_groups.numel() else None,
                                                  long_groups.numel(), SegmentSoftmaxFn.LONG, _lib.stream()),
                       'gb_segment_softmax_fwd')
        ctx.save_for_backward(out, ptr, perm, index)
        return out

    LONG = 256          # groups with more members get one CTA each

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        out, ptr, perm, index = ctx.saved_tensors
        prod = out * dout
        inner = _segment_sums(prod[perm.long()], ptr)
        return out * (dout - inner[index]), None, None, None, None


class DiffusionFn(torch.autograd.Function):
    """f = alpha (A x - x) + beta x0 with A given by `vals` [E] in CSR-slot order.

    alpha = sigmoid(alpha_train) unless no_alpha_sigmoid; beta_train/x0 None -> no source term.
    Gradients: x, vals (SDDMM, only when requested), alpha_train, beta_train.  x0 is detached by the
    caller (base_classes.py:53-55) and gets none."""

    @staticmethod
    def forward(ctx, x, vals, alpha_train, beta_train, x0, plan, alpha_sigmoid, stage):
        lib = _lib.load()
        _lib.require_cuda(x, vals, alpha_train, beta_train, x0)
        x = _f32c(x)
        vals = _f32c(vals)
        n, d = x.shape
        if n != plan.n:
            raise ValueError(f'x has {n} rows but the graph has {plan.n} nodes')
        source = beta_train is not None and x0 is not None
        if source:
            x0 = _f32c(x0)
        out = torch.empty_like(x)
        _lib.check(lib.gb_spmm_fwd(plan.c, vals.data_ptr(), x.data_ptr(), d, _lib.ptr(x0) if source else None, d,
                                   alpha_train.data_ptr(), int(alpha_sigmoid),
                                   beta_train.data_ptr() if source else None, d, out.data_ptr(), d, stage,
                                   _lib.stream()), 'gb_spmm_fwd')
        ctx.save_for_backward(x, vals, alpha_train, x0 if source else None)
        ctx.plan, ctx.alpha_sigmoid, ctx.source = plan, int(alpha_sigmoid), source
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        lib = _lib.load()
        x, vals, alpha_train, x0 = ctx.saved_tensors
        plan = ctx.plan
        n, d = x.shape
        gout = _f32c(gout)
        dx = torch.empty_like(x)
        want_params = ctx.needs_input_grad[2] or ctx.needs_input_grad[3]
        partials = torch.empty(int(lib.gb_spmm_bwd_partials(plan.c, d)), dtype=torch.float32, device=x.device)
        dparams = torch.empty(2, dtype=torch.float32, device=x.device) if want_params else None
        _lib.check(lib.gb_spmm_bwd(plan.c, vals.data_ptr(), gout.data_ptr(), d, x.data_ptr(), d, _lib.ptr(x0), d,
                                   alpha_train.data_ptr(), ctx.alpha_sigmoid, d, dx.data_ptr(), d, 0,
                                   partials.data_ptr(), _lib.ptr(dparams), _lib.stream()), 'gb_spmm_bwd')
        dvals = None
        if ctx.needs_input_grad[1]:
            dvals = torch.empty_like(vals)
            _lib.check(lib.gb_sddmm(plan.c, gout.data_ptr(), d, x.data_ptr(), d, d, alpha_train.data_ptr(),
                                    ctx.alpha_sigmoid, dvals.data_ptr(), _lib.stream()), 'gb_sddmm')
        dalpha = dparams[0].reshape(alpha_train.shape) if ctx.needs_input_grad[2] else None
        dbeta = None
        if ctx.needs_input_grad[3] and ctx.source:
            dbeta = dparams[1].reshape(())
        return (dx if ctx.needs_input_grad[0] else None), dvals, dalpha, dbeta, None, None, None, None


def fused_transformer_supported(plan, x, heads, adim, att_type, norm_idx, edge_w):
    """can FusedTransformerFn take this evaluation? (else the composite QKProj/EdgeSoftmax/Diffusion path runs)"""
    if norm_idx != 0 or edge_w is not None or plan.long_threshold != 256 or adim % heads != 0:
        return False
    if x.dtype != torch.float32 or not x.is_cuda:
        return False
    n, d = x.shape
    # q|k live in one [N, 2A] buffer: row stride 2A floats, k offset A floats
    fake = 256                                       # any 16-byte aligned address: only alignment is inspected
    return bool(_lib.load().gb_att_fused_supported(fake, 2 * adim, fake + 4 * adim, 2 * adim, heads, adim // heads,
                                                   x.data_ptr() if x.is_contiguous() else fake, d, d))


class FusedTransformerFn(torch.autograd.Function):
    """One GRAND-nl evaluation (ODEFuncTransformerAtt.forward, function_transformer_attention.py:44-59):
    projection GEMM + ONE fused graph kernel (score, segment softmax, head mean, SpMM, epilogue).
    The VJP is projection-GEMM transposes + two fused graph kernels (rows: SDDMM, softmax backward,
    dq; columns: transpose SpMM, dk, dalpha/dbeta partial sums)."""

    @staticmethod
    def forward(ctx, x, wq, bq, wk, bk, alpha_train, beta_train, x0, plan, heads, att_type, alpha_sigmoid, kparams,
                wb_cache, stage):
        lib = _lib.load()
        _lib.require_cuda(x, wq, bq, wk, bk, alpha_train, beta_train, x0)
        x = _f32c(x)
        n, d = x.shape
        a = wq.shape[0]
        dk = a // heads
        if wb_cache is not None:
            w, b = wb_cache
        else:
            w = torch.cat([wq, wk], dim=0).contiguous()
            b = torch.cat([bq, bk], dim=0).contiguous()
        source = beta_train is not None and x0 is not None
        if source:
            x0 = _f32c(x0)
        qk = torch.empty(n, 2 * a, dtype=torch.float32, device=x.device)
        _lib.check(lib.gb_linear_fwd(x.data_ptr(), d, n, d, w.data_ptr(), b.data_ptr(), 2 * a, qk.data_ptr(),
                                     2 * a, _lib.stream()), 'gb_linear_fwd')
        store = any(ctx.needs_input_grad[:7])
        need_buf = store or plan.long_rows.numel() > 0
        att = torch.empty(plan.e, heads, dtype=torch.float32, device=x.device) if need_buf else None
        abar = torch.empty(plan.e, dtype=torch.float32, device=x.device) if need_buf else None
        out = torch.empty_like(x)
        _lib.check(lib.gb_att_fwd_fused(plan.c, 0, qk.data_ptr(), 2 * a, qk.data_ptr() + 4 * a, 2 * a, heads, dk,
                                        att_type, None, _lib.ptr(kparams), _lib.ptr(att), _lib.ptr(abar), None,
                                        int(store), x.data_ptr(), d, _lib.ptr(x0) if source else None, d,
                                        alpha_train.data_ptr(), int(alpha_sigmoid),
                                        beta_train.data_ptr() if source else None, d, out.data_ptr(), d, stage,
                                        _lib.stream()), 'gb_att_fwd_fused')
        if store:
            ctx.save_for_backward(x, w, qk, att, abar, alpha_train, x0 if source else None)
        ctx.plan, ctx.heads, ctx.dk, ctx.att_type, ctx.alpha_sigmoid, ctx.source = \
            plan, heads, dk, att_type, int(alpha_sigmoid), source
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        lib = _lib.load()
        if ctx.att_type != _lib.GB_ATT['scaled_dot']:
            raise NotImplementedError('backward of attention types other than scaled_dot is not implemented')
        x, w, qk, att, abar, alpha_train, x0 = ctx.saved_tensors
        plan, heads, dk = ctx.plan, ctx.heads, ctx.dk
        n, d = x.shape
        a2 = qk.shape[1]
        a = a2 // 2
        gout = _f32c(gout)
        dev = x.device
        ds = torch.empty(plan.e, heads, dtype=torch.float32, device=dev)
        dabar = torch.empty(plan.e, dtype=torch.float32, device=dev) if plan.long_rows.numel() > 0 else None
        dqk = torch.empty_like(qk)
        dx = torch.empty_like(x)
        want_params = ctx.needs_input_grad[5] or ctx.needs_input_grad[6]
        partials = torch.empty(int(lib.gb_att_bwd_fused_partials(plan.c)), dtype=torch.float32, device=dev)
        dparams = torch.empty(2, dtype=torch.float32, device=dev) if want_params else None
        _lib.check(lib.gb_att_bwd_fused(plan.c, qk.data_ptr(), a2, qk.data_ptr() + 4 * a, a2, heads, dk,
                                        att.data_ptr(), abar.data_ptr(), gout.data_ptr(), d, x.data_ptr(), d,
                                        _lib.ptr(x0), d, alpha_train.data_ptr(), ctx.alpha_sigmoid, d, ds.data_ptr(),
                                        _lib.ptr(dabar), dqk.data_ptr(), a2, dqk.data_ptr() + 4 * a, a2,
                                        dx.data_ptr(), d, partials.data_ptr(), _lib.ptr(dparams), _lib.stream()),
                   'gb_att_bwd_fused')
        # dense tail: dx += [dq|dk] [Wq;Wk] ; d[Wq;Wk] = [dq|dk]^T x ; d[bq;bk] = column sums
        _lib.check(lib.gb_linear_bwd_x(dqk.data_ptr(), a2, n, a2, w.data_ptr(), d, dx.data_ptr(), d, 1,
                                       _lib.stream()), 'gb_linear_bwd_x')
        dwq = dbq = dwk = dbk = None
        if any(ctx.needs_input_grad[1:5]):
            dw = torch.empty(a2, d, dtype=torch.float32, device=dev)
            db = torch.empty(a2, dtype=torch.float32, device=dev)
            ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n, d, a2)), dtype=torch.uint8, device=dev)
            _lib.check(lib.gb_linear_bwd_w(dqk.data_ptr(), a2, x.data_ptr(), d, n, d, a2, dw.data_ptr(), db.data_ptr(),
                                           ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
            dwq, dbq, dwk, dbk = dw[:a], db[:a], dw[a:], db[a:]
        dalpha = dparams[0].reshape(alpha_train.shape) if ctx.needs_input_grad[5] else None
        dbeta = dparams[1].reshape(()) if (ctx.needs_input_grad[6] and ctx.source) else None
        return (dx if ctx.needs_input_grad[0] else None, dwq, dbq, dwk, dbk, dalpha, dbeta) + (None,) * 8


class HeadMeanToCsrFn(torch.autograd.Function):
    """attention_weights [E,H] in the caller's COO order -> mean over heads in CSR-slot order"""

    @staticmethod
    def forward(ctx, att_coo, plan):
        ctx.plan, ctx.heads = plan, att_coo.reshape(plan.e, -1).shape[1]
        return plan.head_mean_to_csr(att_coo)

    @staticmethod
    @once_differentiable
    def backward(ctx, dabar):
        plan = ctx.plan
        g = (dabar / ctx.heads)[:, None].expand(plan.e, ctx.heads)
        out = torch.empty(plan.e, ctx.heads, dtype=dabar.dtype, device=dabar.device)
        out[plan.eperm.long()] = g
        return out, None


class ToCooOrderFn(torch.autograd.Function):
    """[E,W] CSR-slot order -> caller's COO order (what SpGraphTransAttentionLayer returns)"""

    @staticmethod
    def forward(ctx, rows_csr, plan):
        ctx.plan = plan
        return plan.to_coo_order(rows_csr)

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        plan = ctx.plan
        if plan.identity_order:
            return g, None
        return g[plan.eperm.long()], None


# ---- single-gather GRAND-nl path (csrc/nl_kernels.cuh) --------------------------------------------
_NL_WS = {}


def _nl_workspace(device):
    ws = _NL_WS.get(device)
    if ws is None:
        ws = torch.zeros(int(_lib.load().gb_nl_workspace_bytes()), dtype=torch.uint8, device=device)
        _NL_WS[device] = ws
    return ws


def nl_supported(plan, x, heads, adim, att_type, norm_idx, edge_w):
    """can the single-gather kernels take this evaluation?  (scaled_dot, attention_norm_idx 0, no edge
    reweighting, d <= 128; otherwise the k-gather kernels of fused_att.cu / attention.cu run)"""
    if norm_idx != 0 or edge_w is not None or att_type != _lib.GB_ATT['scaled_dot'] or adim % heads != 0:
        return False
    if plan.long_threshold != 256 or x.dtype != torch.float32 or not x.is_cuda or x.dim() != 2:
        return False
    return bool(_lib.load().gb_nl_supported(heads, adim // heads, x.shape[1]))


def nlk_eligible(heads, adim, d):
    """k-gather row kernels (csrc/nlk_kernels.cuh) instead of the single-gather ones: they fetch k[col] next to x[col]
    (4A more bytes per edge) and drop the per-row Wk work; measured faster whenever d > attention_dim
    (profiles/r02*).  GRANDB200_NL_IMPL=1|2 keeps the single-gather generations for A/B timing."""
    if os.environ.get('GRANDB200_NL_IMPL', '3') != '3' or d < adim:
        return False
    return bool(_lib.load().gb_nlk_supported(heads, adim // heads, d))


def _nl_qk(x, wq, bq, wk, bk):
    """[q | k] = x [Wq; Wk]^T + [bq; bk] in one projection pass ([N, 2A])"""
    lib = _lib.load()
    n, d = x.shape
    a2 = 2 * wq.shape[0]
    w_cat, b_cat = torch.cat([wq, wk], dim=0), torch.cat([bq, bk], dim=0)
    qk = torch.empty(n, a2, dtype=torch.float32, device=x.device)
    _lib.check(lib.gb_linear_fwd(x.data_ptr(), d, n, d, w_cat.data_ptr(), b_cat.data_ptr(), a2, qk.data_ptr(), a2,
                                 _lib.stream()), 'gb_linear_fwd')
    return qk


def _nl_vjp(plan, rows, x, q, wq, wk, x0, alpha_train, beta_train, alpha_sigmoid, heads, gout, want_f, want_params,
            x_row_offset=0, kgather=False):
    """shared by NlTransformerFn.backward and nl_eval_vjp.  Returns (f|None, dx_direct [Ncols,d], dqk [Nrows?,2A],
    dparams|None).  All row-indexed tensors are full-height here (one GPU)."""
    lib = _lib.load()
    n, d = x.shape
    a = wq.shape[0]
    dk = a // heads
    dev = x.device
    ds = torch.empty(plan.e, heads, dtype=torch.float32, device=dev)
    dabar = torch.empty(plan.e, dtype=torch.float32, device=dev)
    abar = torch.empty(plan.e, dtype=torch.float32, device=dev)
    dqk = torch.empty(n, 2 * a, dtype=torch.float32, device=dev)
    dx = torch.empty(n, d, dtype=torch.float32, device=dev)
    out = torch.empty(n, d, dtype=torch.float32, device=dev) if want_f else None
    partials = torch.empty(int(lib.gb_att_bwd_fused_partials(plan.c)), dtype=torch.float32, device=dev)
    dparams = torch.empty(2, dtype=torch.float32, device=dev) if want_params else None
    source = beta_train is not None and x0 is not None
    if kgather:                                              # q is [q | k] ([N, 2A]) here
        _lib.check(lib.gb_nlk_adjoint_phase(
            plan.c, rows[0], rows[1], x.data_ptr(), d, q.data_ptr(), 2 * a, q.data_ptr() + 4 * a, 2 * a, heads, dk, d,
            _lib.ptr(x0) if source else None, d, alpha_train.data_ptr(), int(alpha_sigmoid),
            beta_train.data_ptr() if source else None, gout.data_ptr(), d, _lib.ptr(out), d, ds.data_ptr(),
            dabar.data_ptr(), abar.data_ptr(), dqk.data_ptr(), 2 * a, dqk.data_ptr() + 4 * a, 2 * a, dx.data_ptr(), d,
            partials.data_ptr(), _lib.ptr(dparams), _nl_workspace(dev).data_ptr(), 0, _lib.stream()),
            'gb_nlk_adjoint_phase')
        return out, dx, dqk, dparams
    _lib.check(lib.gb_nl_adjoint(plan.c, rows[0], rows[1], x.data_ptr(), d, q.data_ptr(), a, wk.data_ptr(), d, heads, dk,
                                 d, _lib.ptr(x0) if source else None, d, alpha_train.data_ptr(), int(alpha_sigmoid),
                                 beta_train.data_ptr() if source else None, gout.data_ptr(), d, _lib.ptr(out), d,
                                 ds.data_ptr(), dabar.data_ptr(), abar.data_ptr(), dqk.data_ptr(), 2 * a,
                                 dqk.data_ptr() + 4 * a, 2 * a, dx.data_ptr(), d, partials.data_ptr(),
                                 _lib.ptr(dparams), _nl_workspace(dev).data_ptr(), _lib.stream()), 'gb_nl_adjoint')
    return out, dx, dqk, dparams


def _nl_dense_tail(x, dqk, w_cat, dx, want_w):
    """dx += [dq|dk] [Wq;Wk];  d[Wq;Wk] = [dq|dk]^T x, d[bq;bk] = column sums"""
    lib = _lib.load()
    n, d = x.shape
    a2 = dqk.shape[1]
    _lib.check(lib.gb_linear_bwd_x(dqk.data_ptr(), a2, n, a2, w_cat.data_ptr(), d, dx.data_ptr(), d, 1,
                                   _lib.stream()), 'gb_linear_bwd_x')
    if not want_w:
        return None, None
    dw = torch.empty(a2, d, dtype=torch.float32, device=x.device)
    db = torch.empty(a2, dtype=torch.float32, device=x.device)
    ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n, d, a2)), dtype=torch.uint8, device=x.device)
    _lib.check(lib.gb_linear_bwd_w(dqk.data_ptr(), a2, x.data_ptr(), d, n, d, a2, dw.data_ptr(), db.data_ptr(),
                                   ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
    return dw, db


def _nl_q(x, wq, bq):
    lib = _lib.load()
    n, d = x.shape
    a = wq.shape[0]
    q = torch.empty(n, a, dtype=torch.float32, device=x.device)
    _lib.check(lib.gb_linear_fwd(x.data_ptr(), d, n, d, wq.data_ptr(), bq.data_ptr(), a, q.data_ptr(), a,
                                 _lib.stream()), 'gb_linear_fwd')
    return q


class NlTransformerFn(torch.autograd.Function):
    """One GRAND-nl evaluation (ODEFuncTransformerAtt.forward, function_transformer_attention.py:44-59) as the
    Q projection GEMM + ONE single-gather graph kernel; nothing per-edge is stored for the backward pass,
    which re-runs the row pass fused with the VJP (the x[col] gather is needed by the SDDMM anyway)."""

    @staticmethod
    def forward(ctx, x, wq, bq, wk, bk, alpha_train, beta_train, x0, plan, heads, alpha_sigmoid, stage):
        lib = _lib.load()
        _lib.require_cuda(x, wq, bq, wk, bk, alpha_train, beta_train, x0)
        x = _f32c(x)
        wq, wk, bq = _f32c(wq), _f32c(wk), _f32c(bq)
        n, d = x.shape
        a = wq.shape[0]
        source = beta_train is not None and x0 is not None
        if source:
            x0 = _f32c(x0)
        out = torch.empty_like(x)
        ctx.kgather = nlk_eligible(heads, a, d)
        if ctx.kgather:
            q = _nl_qk(x, wq, bq, wk, _f32c(bk))           # [q | k]
            _lib.check(lib.gb_nlk_fwd(plan.c, 0, n, x.data_ptr(), d, q.data_ptr(), 2 * a, q.data_ptr() + 4 * a, 2 * a,
                                      heads, a // heads, d, _lib.ptr(x0) if source else None, d,
                                      alpha_train.data_ptr(), int(alpha_sigmoid),
                                      beta_train.data_ptr() if source else None, out.data_ptr(), d, stage,
                                      _nl_workspace(x.device).data_ptr(), _lib.stream()), 'gb_nlk_fwd')
        else:
            q = _nl_q(x, wq, bq)
            _lib.check(lib.gb_nl_fwd(plan.c, 0, n, x.data_ptr(), d, q.data_ptr(), a, wk.data_ptr(), d, heads, a // heads,
                                     d, _lib.ptr(x0) if source else None, d, alpha_train.data_ptr(), int(alpha_sigmoid),
                                     beta_train.data_ptr() if source else None, out.data_ptr(), d, stage,
                                     _nl_workspace(x.device).data_ptr(), _lib.stream()), 'gb_nl_fwd')
        ctx.save_for_backward(x, wq, wk, q, alpha_train, beta_train if source else None, x0 if source else None)
        ctx.plan, ctx.heads, ctx.alpha_sigmoid = plan, heads, int(alpha_sigmoid)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, gout):
        x, wq, wk, q, alpha_train, beta_train, x0 = ctx.saved_tensors
        n = x.shape[0]
        a = wq.shape[0]
        gout = _f32c(gout)
        want_params = ctx.needs_input_grad[5] or ctx.needs_input_grad[6]
        _, dx, dqk, dparams = _nl_vjp(ctx.plan, (0, n), x, q, wq, wk, x0, alpha_train, beta_train, ctx.alpha_sigmoid,
                                      ctx.heads, gout, False, want_params, kgather=ctx.kgather)
        w_cat = torch.cat([wq, wk], dim=0)
        dw, db = _nl_dense_tail(x, dqk, w_cat, dx, any(ctx.needs_input_grad[1:5]))
        dwq = dbq = dwk = dbk = None
        if dw is not None:
            dwq, dbq, dwk, dbk = dw[:a], db[:a], dw[a:], db[a:]
        dalpha = dparams[0].reshape(alpha_train.shape) if ctx.needs_input_grad[5] else None
        dbeta = dparams[1].reshape(()) if (ctx.needs_input_grad[6] and beta_train is not None) else None
        return (dx if ctx.needs_input_grad[0] else None, dwq, dbq, dwk, dbk, dalpha, dbeta) + (None,) * 5


def nl_eval_vjp(x, gout, wq, bq, wk, bk, alpha_train, beta_train, x0, plan, heads, alpha_sigmoid, want_w=True):
    """f(x) and the VJP of f at x with cotangent gout in ONE fused pass (what an adjoint solver needs at every
    stage: torchdiffeq's augmented dynamics evaluate f and then call autograd.grad on it).  Returns
    (f, dx, {name: grad}) with the parameter gradients keyed like the autograd inputs."""
    x, gout = _f32c(x.detach()), _f32c(gout.detach())
    wq_, wk_, bq_ = _f32c(wq.detach()), _f32c(wk.detach()), _f32c(bq.detach())
    n = x.shape[0]
    a = wq_.shape[0]
    source = beta_train is not None and x0 is not None
    kgather = nlk_eligible(heads, a, x.shape[1])
    q = _nl_qk(x, wq_, bq_, wk_, _f32c(bk.detach())) if kgather else _nl_q(x, wq_, bq_)
    f, dx, dqk, dparams = _nl_vjp(plan, (0, n), x, q, wq_, wk_, _f32c(x0) if source else None, alpha_train.detach(),
                                  beta_train.detach() if source else None, int(alpha_sigmoid), heads, gout, True, True,
                                  kgather=kgather)
    dw, db = _nl_dense_tail(x, dqk, torch.cat([wq_, wk_], dim=0), dx, want_w)
    grads = {'alpha_train': dparams[0].reshape(alpha_train.shape), 'beta_train': dparams[1].reshape(()) if source else None}
    if dw is not None:
        grads.update({'Q.weight': dw[:a], 'Q.bias': db[:a], 'K.weight': dw[a:], 'K.bias': db[a:]})
    return f, dx, grads
