"""TEST INFRASTRUCTURE: a host-memory emulation of the C ABI of libgrandb200.so (include/grandb200.h), so that the
Python host side of the product (autograd Functions, plan / normalisation plumbing, ODE functions, blocks,
solvers) can be exercised end to end on a machine without a GPU.

It is NOT a CPU fallback of the product: nothing under graph-neural-pde_b200/ imports it, it only exists inside
the `emulated_abi` pytest fixture (tests/conftest.py), which swaps `_lib.load()` for an `Emulator` and lifts the
"CUDA tensors only" guard for the duration of one test.  Every entry point restates the contract written in the
header (numpy / torch fp64 arithmetic on the caller's buffers, rounded to fp32 on store); entry points whose
kernels are selected by `*_supported()` probes answer 0, so evaluations take the general composite route
(projection -> edge softmax -> SpMM), the one that all attention types, both softmax groupings and the
block-level attention share.  Fidelity of the emulation itself is pinned by running the flows that ARE verified
on hardware (scaled_dot forward + every gradient, reference goldens, bit-exact graph normalisation) through it
and comparing with the oracle, exactly like tests/test_gpu_*.py do on the B200.
"""
import ctypes as C
import math

import numpy as np
import torch

from oracle import graph as og

GB_ERR_UNSUPPORTED = 3


def _arr(ptr, ctype, count):
    if not ptr or count == 0:
        return np.zeros(0, dtype=np.dtype(ctype))
    return np.ctypeslib.as_array((ctype * int(count)).from_address(int(ptr)))


def f32(ptr, rows, cols, ld=None):
    """[rows, cols] float32 view with leading dimension ld (None: dense) of caller memory"""
    ld = cols if ld is None else ld
    if rows == 0 or cols == 0:
        return np.zeros((rows, cols), dtype=np.float32)
    flat = _arr(ptr, C.c_float, (rows - 1) * ld + cols)
    return np.lib.stride_tricks.as_strided(flat, shape=(rows, cols), strides=(4 * ld, 4))


def vec(ptr, n):
    return _arr(ptr, C.c_float, n)


def i32(ptr, n):
    return _arr(ptr, C.c_int32, n)


def i64(ptr, n):
    return _arr(ptr, C.c_int64, n)


def scalar(ptr):
    return float(_arr(ptr, C.c_float, 1)[0])


def _p(x):
    """pointer argument as an int (ints, None, ctypes c_void_p)"""
    if x is None:
        return 0
    if isinstance(x, int):
        return x
    return int(getattr(x, 'value', 0) or 0)


class Graph:
    """numpy view of a gb_graph passed as ctypes.byref(GbGraph)"""

    def __init__(self, ref):
        g = ref._obj
        self.n, self.e = int(g.n), int(g.e)
        self.rowptr = i32(g.rowptr, self.n + 1).astype(np.int64)
        self.colidx = i32(g.colidx, self.e).astype(np.int64)
        self.eperm = i32(g.eperm, self.e).astype(np.int64)
        self.colptr = i32(g.colptr, self.n + 1).astype(np.int64)
        self.rowidx = i32(g.rowidx, self.e).astype(np.int64)
        self.tperm = i32(g.tperm, self.e).astype(np.int64)
        self.rows = np.repeat(np.arange(self.n, dtype=np.int64), np.diff(self.rowptr))   # row of CSR slot s


def _alpha(alpha_train, alpha_sigmoid):
    a = scalar(alpha_train)
    return 1.0 / (1.0 + math.exp(-a)) if alpha_sigmoid else a


class Emulator:
    """one method per C entry point the Python host calls; arguments arrive exactly as the host passes them to
    ctypes (ints for pointers, None for NULL, ctypes.byref objects for structs / host out-parameters)"""

    def __init__(self):
        self.launches = 0
        self.error = b''
        self.calls = []                       # names of the entry points called, in order (tests assert on routes)

    _quiet = False                 # set while an entry point is emulated through another one (one call = one note)

    def _note(self, name):
        if self._quiet:
            return
        self.calls.append(name)
        self.launches += 1

    # ---- library --------------------------------------------------------------------------------------------
    def gb_abi_version(self):
        return 2

    def gb_last_error(self):
        return self.error

    def gb_launch_count(self):
        return self.launches

    def _fail(self, msg):
        self.error = msg.encode()
        return GB_ERR_UNSUPPORTED

    # ---- plan / normalisation -------------------------------------------------------------------------------
    def gb_plan_workspace_bytes(self, n, e):
        return 256

    def gb_plan_build(self, row, col, e, n, rowptr, colidx, eperm, colptr, rowidx, tperm, ws, ws_bytes, stream):
        self._note('gb_plan_build')
        ei = np.stack([i64(row, e), i64(col, e)]) if e else np.zeros((2, 0), dtype=np.int64)
        rp, ci, ep = og.coo_to_csr(ei, n)
        cp, ri, tp = og.csr_to_csc(rp, ci, n)
        i32(rowptr, n + 1)[:] = rp
        i32(colptr, n + 1)[:] = cp
        if e:
            i32(colidx, e)[:], i32(eperm, e)[:], i32(rowidx, e)[:], i32(tperm, e)[:] = ci, ep, ri, tp
        return 0

    def gb_plan_long_segments(self, ptr, n, threshold, out, cap, count_ref, ws, ws_bytes, stream):
        p = i32(ptr, n + 1).astype(np.int64)
        ids = np.nonzero(np.diff(p) > threshold)[0]
        i32(out, cap)[:ids.size] = ids
        count_ref._obj.value = int(ids.size)
        return 0

    def gb_degree_workspace_bytes(self, n, e):
        return 16

    def gb_degree(self, idx, w, e, n, deg, ws, ws_bytes, stream):
        self._note('gb_degree')
        wv = vec(w, e) if _p(w) else np.ones(e, dtype=np.float32)
        vec(deg, n)[:] = og._seq_scatter_add_f32(np.asarray(wv, dtype=np.float32), i64(idx, e), n)
        return 0

    def gb_rw_normalize(self, idx, w, deg, e, w_out, stream):
        self._note('gb_rw_normalize')
        n = int(i64(idx, e).max()) + 1 if e else 0
        inv = np.float32(1.0) / vec(deg, n)                       # reciprocal, THEN multiply (utils.py:231-232)
        vec(w_out, e)[:] = vec(w, e) * inv[i64(idx, e)]
        return 0

    def gb_gcn_normalize(self, row, col, w, deg, e, w_out, stream):
        self._note('gb_gcn_normalize')
        r, c = i64(row, e), i64(col, e)
        n = int(max(r.max(), c.max())) + 1 if e else 0
        with np.errstate(divide='ignore'):
            dis = (np.float32(1.0) / np.sqrt(vec(deg, n), dtype=np.float32)).astype(np.float32)   # deg.pow_(-0.5)
        dis[np.isinf(dis)] = 0
        vec(w_out, e)[:] = (dis[r] * vec(w, e)) * dis[c]
        return 0

    def gb_gather_f32(self, src, perm, e, dst, stream):
        self._note('gb_gather_f32')
        vec(dst, e)[:] = vec(src, e)[i32(perm, e)]
        return 0

    def gb_head_mean_gather(self, att_coo, heads, eperm, e, abar_csr, stream):
        self._note('gb_head_mean_gather')
        vec(abar_csr, e)[:] = f32(att_coo, e, heads)[i32(eperm, e)].astype(np.float64).mean(axis=1)
        return 0

    def gb_scatter_rows_f32(self, src_csr, width, eperm, e, dst_coo, stream):
        self._note('gb_scatter_rows_f32')
        f32(dst_coo, e, width)[i32(eperm, e)] = f32(src_csr, e, width)
        return 0

    # ---- K1 / K5a: SpMM, transpose SpMM, SDDMM ----------------------------------------------------------------
    def gb_spmm_fwd(self, g, vals, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, beta_train, d, out, ldo, stage, stream):
        self._note('gb_spmm_fwd')
        if stage is not None:
            return self._fail('emulator: Runge-Kutta stage fusion is not emulated')
        G = Graph(g)
        X = f32(x, G.n, d, ldx).astype(np.float64)
        ax = np.zeros((G.n, d))
        np.add.at(ax, G.rows, vec(vals, G.e).astype(np.float64)[:, None] * X[G.colidx])
        f = _alpha(alpha_train, alpha_sigmoid) * (ax - X)
        if _p(x0) and _p(beta_train):
            f = f + scalar(beta_train) * f32(x0, G.n, d, ldx0)
        f32(out, G.n, d, ldo)[:] = f
        return 0

    def gb_spmm_bwd_partials(self, g, d):
        return 4

    def gb_spmm_bwd(self, g, vals, gout, ldg, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, d, dx, lddx, accumulate,
                    partials, dparams, stream):
        self._note('gb_spmm_bwd')
        G = Graph(g)
        go = f32(gout, G.n, d, ldg).astype(np.float64)
        t = -go.copy()
        np.add.at(t, G.colidx, vec(vals, G.e).astype(np.float64)[:, None] * go[G.rows])
        alpha = _alpha(alpha_train, alpha_sigmoid)
        dxv = f32(dx, G.n, d, lddx)
        dxv[:] = (dxv if accumulate else 0) + alpha * t
        if _p(dparams):
            X = f32(x, G.n, d, ldx).astype(np.float64)
            dp = vec(dparams, 2)
            dp[0] = (alpha * (1 - alpha) if alpha_sigmoid else 1.0) * float((X * t).sum())
            dp[1] = float((go * f32(x0, G.n, d, ldx0)).sum()) if _p(x0) else 0.0
        return 0

    def gb_sddmm(self, g, gout, ldg, x, ldx, d, alpha_train, alpha_sigmoid, dabar, stream):
        self._note('gb_sddmm')
        G = Graph(g)
        go, X = f32(gout, G.n, d, ldg).astype(np.float64), f32(x, G.n, d, ldx).astype(np.float64)
        vec(dabar, G.e)[:] = _alpha(alpha_train, alpha_sigmoid) * (go[G.rows] * X[G.colidx]).sum(axis=1)
        return 0

    # ---- K3: projections --------------------------------------------------------------------------------------
    def gb_linear_fwd(self, x, ldx, n, d, w, b, nout, out, ldo, stream):
        self._note('gb_linear_fwd')
        y = f32(x, n, d, ldx).astype(np.float64) @ f32(w, nout, d).astype(np.float64).T
        f32(out, n, nout, ldo)[:] = y + (vec(b, nout) if _p(b) else 0)
        return 0

    def gb_linear_bwd_x(self, dy, lddy, n, nout, w, d, dx, lddx, accumulate, stream):
        self._note('gb_linear_bwd_x')
        v = f32(dy, n, nout, lddy).astype(np.float64) @ f32(w, nout, d).astype(np.float64)
        dxv = f32(dx, n, d, lddx)
        dxv[:] = (dxv if accumulate else 0) + v
        return 0

    def gb_linear_bwd_w_workspace_bytes(self, n, d, nout):
        return 16

    def gb_linear_bwd_w(self, dy, lddy, x, ldx, n, d, nout, dw, db, ws, ws_bytes, stream):
        self._note('gb_linear_bwd_w')
        DY = f32(dy, n, nout, lddy).astype(np.float64)
        f32(dw, nout, d)[:] = DY.T @ f32(x, n, d, ldx).astype(np.float64)
        vec(db, nout)[:] = DY.sum(axis=0)
        return 0

    # ---- K2 + K4 / K5b: edge scores, segment softmax and their VJP ----------------------------------------------
    def gb_att_fused_supported(self, q, ldq, k, ldk, heads, dk, x, ldx, d):
        return 0                                 # the warp-per-row fused kernels are not emulated: composite route

    def gb_nl_workspace_bytes(self):
        return 16

    def gb_edge_softmax_fwd(self, g, norm_idx, q, ldq, k, ldk, heads, dk, att_type, edge_w, kparams, att, abar, prods,
                            stream):
        self._note('gb_edge_softmax_fwd')
        from oracle import funcs
        G = Graph(g)
        a = heads * dk
        Q = torch.tensor(f32(q, G.n, a, ldq).astype(np.float64))
        K = torch.tensor(f32(k, G.n, a, ldk).astype(np.float64))
        ei = torch.tensor(np.stack([G.rows, G.colidx]))
        name = ['scaled_dot', 'cosine_sim', 'pearson', 'exp_kernel'][att_type]
        kw = {}
        if name == 'exp_kernel':
            kp = vec(kparams, 2)
            kw = dict(output_var=torch.tensor(float(kp[0]), dtype=torch.float64),
                      lengthscale=torch.tensor(float(kp[1]), dtype=torch.float64))
        s = funcs.edge_scores(Q, K, ei, heads, name, **kw)
        if _p(edge_w):
            s = s * torch.tensor(vec(edge_w, G.e).astype(np.float64))[:, None]
        p = funcs.segment_softmax(s, ei[norm_idx], G.n)
        f32(att, G.e, heads)[:] = p.numpy()
        vec(abar, G.e)[:] = p.mean(dim=1).numpy()
        if _p(prods):
            f32(prods, G.e, heads)[:] = s.numpy()
        return 0

    def gb_decode_eval_max_classes(self):
        return 64

    def gb_decode_eval(self, z, ldz, n, d, w2, ldw, b2, classes, y, train_mask, val_mask, test_mask, counts, logits,
                       ldl, pred, stream):
        self._note('gb_decode_eval')
        if classes > 64:
            return 3
        from oracle import funcs
        masks = [torch.tensor(_arr(_p(m), C.c_uint8, n).copy()) if _p(m) else None for m in (train_mask, val_mask, test_mask)]
        yv = torch.tensor(i64(y, n).copy()) if _p(y) else torch.zeros(n, dtype=torch.long)
        lg, pr, cnt = funcs.decode_eval(torch.tensor(f32(z, n, d, ldz).astype(np.float64)),
                                        torch.tensor(f32(w2, classes, d, ldw).astype(np.float64)),
                                        torch.tensor(vec(b2, classes).astype(np.float64)) if _p(b2) else None, yv, masks)
        if _p(counts):
            i32(counts, 6)[:] = [c for c, _ in cnt] + [t for _, t in cnt]
        if _p(logits):
            f32(logits, n, classes, ldl)[:] = lg.numpy()
        if _p(pred):
            i64(pred, n)[:] = pr.numpy()
        return 0

    def gb_gat_softmax_fwd(self, g, norm_idx, s_src, s_dst, lds, heads, slope, att, abar, prods, stream):
        self._note('gb_gat_softmax_fwd')
        from oracle import funcs
        G = Graph(g)
        src = f32(s_src, G.n, heads, lds).astype(np.float64)
        dst = f32(s_dst, G.n, heads, lds).astype(np.float64)
        z = src[G.rows] + dst[G.colidx]
        s = torch.tensor(np.where(z > 0, z, float(slope) * z))
        grp = torch.tensor(G.rows if norm_idx == 0 else G.colidx)
        p = funcs.segment_softmax(s, grp.long(), G.n)
        f32(att, G.e, heads)[:] = p.numpy()
        vec(abar, G.e)[:] = p.mean(dim=1).numpy()
        if _p(prods):
            f32(prods, G.e, heads)[:] = s.numpy()
        return 0

    def gb_segment_softmax_fwd(self, ptr, perm, n, heads, scores, att, abar, long_list, n_long, long_threshold, stream):
        self._note('gb_segment_softmax_fwd')
        if n_long:                                   # the listed groups are exactly the ones above the threshold
            sizes = np.diff(i32(ptr, n + 1))
            assert sorted(i32(long_list, n_long).tolist()) == np.nonzero(sizes > long_threshold)[0].tolist()
        from oracle import funcs
        bounds = i32(ptr, n + 1)
        e = int(bounds[-1])
        order = i32(perm, e).astype(np.int64) if _p(perm) else np.arange(e)
        grp = np.empty(e, dtype=np.int64)
        grp[order] = np.repeat(np.arange(n), np.diff(bounds))
        p = funcs.segment_softmax(torch.tensor(f32(scores, e, heads).astype(np.float64)), torch.tensor(grp), n)
        f32(att, e, heads)[:] = p.numpy()
        if _p(abar):
            vec(abar, e)[:] = p.mean(dim=1).numpy()
        return 0

    def gb_stage_combine(self, y, ks, cs, nk, count, out, stream):
        self._note('gb_stage_combine')
        acc = vec(y, count).astype(np.float64)
        for j in range(nk):
            acc = acc + float(cs[j]) * vec(ks[j], count)
        vec(out, count)[:] = acc
        return 0

    def gb_stage_combine_multi(self, y, ks, cs, nk, count, outs, n_out, stream):
        self._note('gb_stage_combine_multi')
        acc = vec(y, count).astype(np.float64)
        for j in range(nk):
            acc = acc + float(cs[j]) * vec(ks[j], count)
        for o in range(n_out):
            vec(outs[o], count)[:] = acc
        return 0

    def gb_edge_softmax_bwd(self, g, norm_idx, att, dabar, datt, heads, edge_w, ds, stream):
        self._note('gb_edge_softmax_bwd')
        G = Graph(g)
        A = f32(att, G.e, heads).astype(np.float64)
        da = np.zeros((G.e, heads))
        if _p(dabar):
            da += vec(dabar, G.e).astype(np.float64)[:, None] / heads
        if _p(datt):
            da += f32(datt, G.e, heads)
        grp = G.rows if norm_idx == 0 else G.colidx
        inner = np.zeros((G.n, heads))
        np.add.at(inner, grp, A * da)
        out = A * (da - inner[grp])
        if _p(edge_w):                            # scores were prods * edge_w: chain to the raw score
            out = out * vec(edge_w, G.e).astype(np.float64)[:, None]
        f32(ds, G.e, heads)[:] = out
        return 0

    def gb_edge_score_bwd(self, g, ds, q, ldq, k, ldk, heads, dk, att_type, kparams, prods, dq, lddq, dk_out, lddk,
                          dkp_partials, dkp, stream):
        self._note('gb_edge_score_bwd')
        if att_type != 0:
            return self._fail(f'gb_edge_score_bwd: backward of attention_type {att_type} is not implemented (scaled_dot only)')
        G = Graph(g)
        a = heads * dk
        DS = f32(ds, G.e, heads).astype(np.float64)[:, :, None] / math.sqrt(dk)
        Q = f32(q, G.n, a, ldq).astype(np.float64).reshape(G.n, heads, dk)
        K = f32(k, G.n, a, ldk).astype(np.float64).reshape(G.n, heads, dk)
        gq, gk = np.zeros((G.n, heads, dk)), np.zeros((G.n, heads, dk))
        np.add.at(gq, G.rows, DS * K[G.colidx])
        np.add.at(gk, G.colidx, DS * Q[G.rows])
        f32(dq, G.n, a, lddq)[:] = gq.reshape(G.n, a)
        f32(dk_out, G.n, a, lddk)[:] = gk.reshape(G.n, a)
        return 0

    # ---- row-range forms (1-D row partition) and the single-gather GRAND-nl entry points ------------------------
    # Matrices indexed by the row (q, x0, out, gout, dq) arrive as "virtual" base pointers (slab pointer minus
    # lo * ld floats, partition._virt): only rows [lo, hi) may be touched.  x is indexed by global column id.
    @staticmethod
    def _rows(ptr, lo, hi, cols, ld):
        return f32(_p(ptr) + 4 * lo * ld, hi - lo, cols, ld)

    nl_enabled = False                            # tests flip this to route evaluations to gb_nl_fwd / gb_nl_adjoint

    def gb_nl_supported(self, heads, dk, d):
        return int(self.nl_enabled and heads in (1, 2, 4, 8) and dk % 4 == 0 and d % 4 == 0 and d <= 128 and heads * dk <= 128)

    def gb_spmm_fwd_part(self, g, lo, hi, vals, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, beta_train, d, out, ldo,
                         stage, stream):
        self._note('gb_spmm_fwd_part')
        G = Graph(g)
        X = f32(x, G.n, d, ldx).astype(np.float64)
        ax = np.zeros((G.n, d))
        np.add.at(ax, G.rows, vec(vals, G.e).astype(np.float64)[:, None] * X[G.colidx])
        f = _alpha(alpha_train, alpha_sigmoid) * (ax - X)[lo:hi]
        if _p(x0) and _p(beta_train):
            f = f + scalar(beta_train) * self._rows(x0, lo, hi, d, ldx0)
        self._rows(out, lo, hi, d, ldo)[:] = f
        return 0

    def gb_spmm_bwd_part(self, g, own_lo, own_hi, vals, gout, ldg, x, ldx, x0, ldx0, alpha_train, alpha_sigmoid, d, dx,
                         lddx, accumulate, partials, dparams, stream):
        self._note('gb_spmm_bwd_part')
        G = Graph(g)
        if own_lo == 0 and own_hi == 0:
            own_lo, own_hi = 0, G.n
        own_hi = max(own_hi, own_lo)              # (lo, -1): a rank that owns no rows
        go = np.zeros((G.n, d))
        go[own_lo:own_hi] = self._rows(gout, own_lo, own_hi, d, ldg)      # a rank's edges only touch its rows of g
        t = np.zeros((G.n, d))
        np.add.at(t, G.colidx, vec(vals, G.e).astype(np.float64)[:, None] * go[G.rows])
        t[own_lo:own_hi] -= go[own_lo:own_hi]
        alpha = _alpha(alpha_train, alpha_sigmoid)
        dxv = f32(dx, G.n, d, lddx)
        dxv[:] = (dxv if accumulate else 0) + alpha * t
        if _p(dparams):
            dp = vec(dparams, 2)
            dp[0] = (alpha * (1 - alpha) if alpha_sigmoid else 1.0) * float((f32(x, G.n, d, ldx) * t).sum())
            dp[1] = float((go[own_lo:own_hi] * self._rows(x0, own_lo, own_hi, d, ldx0)).sum()) if _p(x0) else 0.0
        return 0

    def gb_att_bwd_fused_partials(self, g):
        return 4

    def _nl_attention(self, G, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d):
        """row-softmax scaled_dot attention of the plan's edges with k = Wk x (the bias cancels in the row softmax)"""
        a = heads * dk
        X = f32(x, G.n, d, ldx).astype(np.float64)
        Q = np.zeros((G.n, a))
        Q[lo:hi] = self._rows(q, lo, hi, a, ldq)
        if isinstance(wk, tuple):                      # ('k', pointer): the k-gather entry points pass k itself
            K = f32(wk[1], G.n, a, ldwk).astype(np.float64)
        else:
            K = X @ f32(wk, a, d, ldwk).astype(np.float64).T
        s = (Q.reshape(G.n, heads, dk)[G.rows] * K.reshape(G.n, heads, dk)[G.colidx]).sum(axis=2) / math.sqrt(dk)
        m = np.full((G.n, heads), -np.inf)
        np.maximum.at(m, G.rows, s)
        p = np.exp(s - m[G.rows])
        den = np.zeros((G.n, heads))
        np.add.at(den, G.rows, p)
        return X, Q, K, p / den[G.rows]

    def gb_nl_fwd(self, g, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train, alpha_sigmoid,
                  beta_train, out, ldo, stage, workspace, stream):
        self._note('gb_nl_fwd')
        if stage is not None:
            return self._fail('emulator: Runge-Kutta stage fusion is not emulated')
        G = Graph(g)
        X, _, _, att = self._nl_attention(G, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d)
        ax = np.zeros((G.n, d))
        np.add.at(ax, G.rows, att.mean(axis=1)[:, None] * X[G.colidx])
        f = _alpha(alpha_train, alpha_sigmoid) * (ax - X)[lo:hi]
        if _p(x0) and _p(beta_train):
            f = f + scalar(beta_train) * self._rows(x0, lo, hi, d, ldx0)
        self._rows(out, lo, hi, d, ldo)[:] = f
        return 0

    # k-gather entry points (csrc/nlk_kernels.cuh): same functions with k = x Wk^T + bk given instead of Wk
    def gb_nlk_supported(self, heads, dk, d):
        return int(self.gb_nl_supported(heads, dk, d) and heads * dk <= 128)

    def gb_nlk_fwd(self, g, lo, hi, x, ldx, q, ldq, k, ldk, heads, dk, d, x0, ldx0, alpha_train, alpha_sigmoid,
                   beta_train, out, ldo, stage, workspace, stream):
        self._note('gb_nlk_fwd')
        self._quiet = True
        try:
            return self.gb_nl_fwd(g, lo, hi, x, ldx, q, ldq, ('k', k), ldk, heads, dk, d, x0, ldx0, alpha_train,
                                  alpha_sigmoid, beta_train, out, ldo, stage, workspace, stream)
        finally:
            self._quiet = False

    def gb_nlk_adjoint_phase(self, g, lo, hi, x, ldx, q, ldq, k, ldk, heads, dk, d, x0, ldx0, alpha_train,
                             alpha_sigmoid, beta_train, gout, ldg, out, ldo, ds, dabar, abar, dq, lddq, dk_out, lddk,
                             dx, lddx, partials, dparams, workspace, phase, stream):
        self._note({0: 'gb_nlk_adjoint', 1: 'gb_nlk_adjoint_rows', 2: 'gb_nlk_adjoint_cols'}[phase])
        self._quiet = True
        try:
            return self.gb_nl_adjoint_phase(g, lo, hi, x, ldx, q, ldq, ('k', k), ldk, heads, dk, d, x0, ldx0,
                                            alpha_train, alpha_sigmoid, beta_train, gout, ldg, out, ldo, ds, dabar,
                                            abar, dq, lddq, dk_out, lddk, dx, lddx, partials, dparams, workspace,
                                            phase, stream)
        finally:
            self._quiet = False

    def gb_nl_adjoint(self, g, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train, alpha_sigmoid,
                      beta_train, gout, ldg, out, ldo, ds, dabar, abar, dq, lddq, dk_out, lddk, dx, lddx, partials,
                      dparams, workspace, stream):
        return self.gb_nl_adjoint_phase(g, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train,
                                        alpha_sigmoid, beta_train, gout, ldg, out, ldo, ds, dabar, abar, dq, lddq, dk_out,
                                        lddk, dx, lddx, partials, dparams, workspace, 0, stream)

    def gb_nl_adjoint_phase(self, g, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d, x0, ldx0, alpha_train,
                            alpha_sigmoid, beta_train, gout, ldg, out, ldo, ds, dabar, abar, dq, lddq, dk_out, lddk, dx,
                            lddx, partials, dparams, workspace, phase, stream):
        """phase 0: both halves; 1: row pass (f, dq, ds, abar); 2: column pass, which READS abar / ds back from the
        caller's buffers (so a host that hands the second half different buffers is caught)"""
        self._note({0: 'gb_nl_adjoint', 1: 'gb_nl_adjoint_rows', 2: 'gb_nl_adjoint_cols'}[phase])
        G = Graph(g)
        a = heads * dk
        alpha = _alpha(alpha_train, alpha_sigmoid)
        X = f32(x, G.n, d, ldx).astype(np.float64)
        go = np.zeros((G.n, d))
        go[lo:hi] = self._rows(gout, lo, hi, d, ldg)
        if phase != 2:
            _, Q, K, att = self._nl_attention(G, lo, hi, x, ldx, q, ldq, wk, ldwk, heads, dk, d)
            ab = att.mean(axis=1)
            if _p(out):
                ax = np.zeros((G.n, d))
                np.add.at(ax, G.rows, ab[:, None] * X[G.colidx])
                f = alpha * (ax - X)[lo:hi]
                if _p(x0) and _p(beta_train):
                    f = f + scalar(beta_train) * self._rows(x0, lo, hi, d, ldx0)
                self._rows(out, lo, hi, d, ldo)[:] = f
            dab = alpha * (go[G.rows] * X[G.colidx]).sum(axis=1)
            da = dab[:, None] / heads
            inner = np.zeros((G.n, heads))
            np.add.at(inner, G.rows, att * da)
            f32(ds, G.e, heads)[:] = att * (da - inner[G.rows])
            vec(abar, G.e)[:] = ab
            if _p(dabar):
                vec(dabar, G.e)[:] = dab
            gq = np.zeros((G.n, heads, dk))
            np.add.at(gq, G.rows, f32(ds, G.e, heads).astype(np.float64)[:, :, None] / math.sqrt(dk)
                      * K.reshape(G.n, heads, dk)[G.colidx])
            self._rows(dq, lo, hi, a, lddq)[:] = gq.reshape(G.n, a)[lo:hi]
        if phase == 1:
            return 0
        ab = vec(abar, G.e).astype(np.float64)
        w = f32(ds, G.e, heads).astype(np.float64)[:, :, None] / math.sqrt(dk)
        Q = np.zeros((G.n, a))
        Q[lo:hi] = self._rows(q, lo, hi, a, ldq)
        gk = np.zeros((G.n, heads, dk))
        np.add.at(gk, G.colidx, w * Q.reshape(G.n, heads, dk)[G.rows])
        f32(dk_out, G.n, a, lddk)[:] = gk.reshape(G.n, a)
        t = np.zeros((G.n, d))
        np.add.at(t, G.colidx, ab[:, None] * go[G.rows])
        t[lo:hi] -= go[lo:hi]
        f32(dx, G.n, d, lddx)[:] = alpha * t
        if _p(dparams):
            dp = vec(dparams, 2)
            dp[0] = (alpha * (1 - alpha) if alpha_sigmoid else 1.0) * float((X * t).sum())
            dp[1] = float((go[lo:hi] * self._rows(x0, lo, hi, d, ldx0)).sum()) if _p(x0) else 0.0
        return 0
