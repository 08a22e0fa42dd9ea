"""`ODEFuncTransformerAtt` and `SpGraphTransAttentionLayer` - drop-ins for the reference's
src/function_transformer_attention.py:9-270.

    attention, (v, prods) = multihead_att_layer(x, edge_index)       # recomputed every evaluation
    f(t, x) = alpha * (mean_h(attention) applied as A to x  -  x) + beta * x0

One evaluation = three kernels: Q|K projection GEMM, fused per-edge score + segment softmax + head
mean (csrc/attention.cu), fused SpMM + epilogue (csrc/spmm.cu).  `scaled_dot` scores are per-edge
q[row].k[col]/sqrt(dk) (upstream / BASELINE north_star semantics; the fork's line 249 sums k over
ALL edges, SURVEY.md F4 - reproduced only by the oracle's fork_literal switch).

Parameter names/shapes match the reference (`Q`, `K`, `V`, `Wout` Linear layers initialised to the
constant 1e-5, default biases; `output_var`, `lengthscale` for exp_kernel).  `V` and `Wout` are
only read by the `mix_features` branch, which is broken in the fork (tuple passed where a tensor is
expected) and raises NotImplementedError here; their gradients are None, as in the reference.
"""
import torch
from torch import nn

from .base_classes import ODEFunc
from . import _lib
from . import plan as _plan
from .ops import (QKProjFn, EdgeSoftmaxFn, EdgeScoreFn, squareplus_csr, DiffusionFn, ToCooOrderFn, FusedTransformerFn, fused_transformer_supported,
                  NlTransformerFn, nl_supported, nl_eval_vjp)


def unit_heads(qk, heads, centre):
    """[N, 2A] projections q|k -> per-(node, head) unit vectors (mean-centred first when `centre`), the q half
    multiplied by sqrt(dk).  The scaled_dot kernels applied to the result give the reference's cosine_sim /
    pearson scores: torch.nn.CosineSimilarity(eps=1e-5) is  <q / max(|q|, eps), k / max(|k|, eps)>  per head
    (function_transformer_attention.py:251-259), and the kernels' 1/sqrt(dk) cancels the factor put on q.
    Node-level torch ops ([N, 2A] elementwise), differentiated by autograd: this is how the backward pass of
    these two attention types is formed (the per-edge work stays in the hand-written scaled_dot VJP kernels)."""
    n, a2 = qk.shape
    dk = a2 // (2 * heads)
    v = qk.reshape(n, 2 * heads, dk)
    if centre:
        v = v - v.mean(dim=2, keepdim=True)
    v = v / v.norm(dim=2, keepdim=True).clamp_min(1e-5)
    scale = torch.ones(2 * heads, dtype=qk.dtype, device=qk.device)
    scale[:heads] = float(dk) ** 0.5
    return (v * scale[None, :, None]).reshape(n, a2)


class SpGraphTransAttentionLayer(nn.Module):
    """function_transformer_attention.py:65-270"""

    def __init__(self, in_features, out_features, opt, device, concat=True, edge_weights=None):
        super(SpGraphTransAttentionLayer, self).__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.alpha = opt['leaky_relu_slope']
        self.concat = concat
        self.device = device
        self.opt = opt
        self.h = int(opt['heads'])
        self.edge_weights = edge_weights
        try:
            self.attention_dim = opt['attention_dim']
        except KeyError:
            self.attention_dim = out_features
        assert self.attention_dim % self.h == 0, "Number of heads ({}) must be a factor of the dimension size ({})".format(
            self.h, self.attention_dim)
        self.d_k = self.attention_dim // self.h
        self.beltrami_kernel = bool(opt.get('beltrami', False)) and opt.get('attention_type') == 'exp_kernel'
        # one warp holds a q / k row as 32 lanes x 4 floats in every score kernel (the beltrami kernel product works on
        # per-head vectors of twice the width); the reference's tuned configurations use 8 .. 128 (best_params.py)
        width = self.attention_dim * (2 if self.beltrami_kernel else 1)
        if width > 128:
            raise NotImplementedError(f"opt['attention_dim'] = {self.attention_dim}" +
                                      (' with beltrami + exp_kernel' if self.beltrami_kernel else '') +
                                      ': the score kernels hold at most 128 floats of q / k per node')
        if self.beltrami_kernel:
            fd = opt['hidden_dim'] - opt['pos_enc_hidden_dim']
            for name in ('output_var_x', 'lengthscale_x', 'output_var_p', 'lengthscale_p'):
                setattr(self, name, nn.Parameter(torch.ones(1)))
            for name, width in (('Qx', fd), ('Vx', fd), ('Kx', fd), ('Qp', opt['pos_enc_hidden_dim']),
                                ('Vp', opt['pos_enc_hidden_dim']), ('Kp', opt['pos_enc_hidden_dim'])):
                lin = nn.Linear(width, self.attention_dim)
                self.init_weights(lin)
                setattr(self, name, lin)
        else:
            if opt.get('attention_type') == 'exp_kernel':
                self.output_var = nn.Parameter(torch.ones(1))
                self.lengthscale = nn.Parameter(torch.ones(1))
            self.Q = nn.Linear(in_features, self.attention_dim)
            self.init_weights(self.Q)
            self.V = nn.Linear(in_features, self.attention_dim)
            self.init_weights(self.V)
            self.K = nn.Linear(in_features, self.attention_dim)
            self.init_weights(self.K)
        if opt.get('multi_modal', False):
            raise NotImplementedError("opt['multi_modal'] is out of scope (SURVEY.md 2.1)")
        self.activation = nn.Sigmoid()
        self.Wout = nn.Linear(self.d_k, in_features)
        self.init_weights(self.Wout)

    def init_weights(self, m):
        if type(m) == nn.Linear:
            nn.init.constant_(m.weight, 1e-5)      # function_transformer_attention.py:153-157

    def _check_supported(self):
        att_type = self.opt.get('attention_type', 'scaled_dot')
        if att_type not in _lib.GB_ATT:
            raise NotImplementedError(f"attention_type '{att_type}'")
        return _lib.GB_ATT[att_type]

    @property
    def square_plus(self):
        """opt['square_plus']: utils.squareplus instead of the segment softmax (utils.py:129-140).  The fork's call
        site (function_transformer_attention.py:264) passes two arguments to that three-argument function and
        cannot run; the function itself is well defined (pinned by tests/golden/ref_squareplus.npz) and is applied
        with the group ids the softmax would use."""
        return bool(self.opt.get('square_plus', False))

    @property
    def _UNIT_TYPES(self):
        return (_lib.GB_ATT['cosine_sim'], _lib.GB_ATT['pearson'])

    def needs_composite_backward(self, att_type, x):
        """evaluations of the other attention types that will be differentiated run as projection -> edge softmax
        -> SpMM (the fused forward kernels have no VJP for them): cosine_sim / pearson as scaled_dot of unit_heads,
        exp_kernel through ops.exp_kernel_score_vjp"""
        if self.beltrami_kernel or self.square_plus:
            return True                                    # always the composite route (two projections, one kernel)
        if att_type == _lib.GB_ATT['scaled_dot'] or not torch.is_grad_enabled():
            return False
        params = [self.Q.weight, self.Q.bias, self.K.weight, self.K.bias]
        if att_type == _lib.GB_ATT['exp_kernel']:
            params += [self.output_var, self.lengthscale]
        return x.requires_grad or any(p.requires_grad for p in params)

    def qk_weights(self):
        """[Wq;Wk], [bq;bk] concatenated once per parameter version (parameters change once per
        optimiser step, evaluations happen NFE times per step)"""
        key = (self.Q.weight._version, self.Q.bias._version, self.K.weight._version, self.K.bias._version,
               self.Q.weight.data_ptr(), self.K.weight.data_ptr())
        hit = getattr(self, '_wb_cache', None)
        if hit is None or hit[0] != key:
            with torch.no_grad():
                w = torch.cat([self.Q.weight, self.K.weight], dim=0).contiguous()
                b = torch.cat([self.Q.bias, self.K.bias], dim=0).contiguous()
            hit = (key, (w, b))
            self._wb_cache = hit
        return hit[1]

    def _beltrami_qk(self, x2):
        """beltrami + exp_kernel (function_transformer_attention.py:164-216): the score is the PRODUCT of an
        exponential kernel on the feature columns and one on the positional columns,
            sx^2 exp(-|qx - kx|^2 / 2 lx^2) * sp^2 exp(-|qp - kp|^2 / 2 lp^2)
              = (sx sp)^2 exp(-|[qx/lx, qp/lp] - [kx/lx, kp/lp]|^2 / 2),
        i.e. ONE exponential kernel with output_var sx*sp and lengthscale 1 on per-head vectors of width 2 d_k.
        Returns ([q' | k'] [N, 4A], kparams [2]).  The fork slices `x[:, f:label]` on its batched layout (the node
        axis); the feature axis is meant."""
        f = int(self.opt['feat_hidden_dim'])
        lab = f + int(self.opt['pos_enc_hidden_dim'])
        n = x2.shape[0]
        p = x2[:, f:lab].contiguous()
        xf = torch.cat((x2[:, :f], x2[:, lab:]), dim=1)
        qk_x = QKProjFn.apply(xf, self.Qx.weight, self.Qx.bias, self.Kx.weight, self.Kx.bias)
        qk_p = QKProjFn.apply(p, self.Qp.weight, self.Qp.bias, self.Kp.weight, self.Kp.bias)
        a, h, dk = self.attention_dim, self.h, self.d_k
        halves = []
        for lo in (0, a):                                  # q' then k'
            hx = qk_x[:, lo:lo + a].reshape(n, h, dk) / self.lengthscale_x
            hp = qk_p[:, lo:lo + a].reshape(n, h, dk) / self.lengthscale_p
            halves.append(torch.cat((hx, hp), dim=2).reshape(n, 2 * a))
        kparams = torch.cat([self.output_var_x * self.output_var_p, torch.ones_like(self.lengthscale_x)])
        return torch.cat(halves, dim=1).contiguous(), kparams.to(torch.float32).contiguous()

    def attention_csr(self, x2, plan):
        """(att [E,H], abar [E], prods [E,H]) in CSR-slot order for an un-batched state x2 [N,d]"""
        att_type = self._check_supported()
        edge_w = None
        if self.opt.get('reweight_attention', False) and self.edge_weights is not None:
            edge_w = _plan.cached_per_edge(plan, self.edge_weights, 'rw', lambda t: plan.to_csr_order(t.detach()))
        if self.beltrami_kernel:
            qk, kparams = self._beltrami_qk(x2)
            if self.square_plus:
                return self._squareplus(qk, plan, att_type, edge_w, kparams)
            return EdgeSoftmaxFn.apply(qk, plan, self.h, att_type, int(self.opt['attention_norm_idx']), edge_w, kparams)
        qk = QKProjFn.apply(x2, self.Q.weight, self.Q.bias, self.K.weight, self.K.bias)
        kparams = None
        if att_type == _lib.GB_ATT['exp_kernel']:            # differentiable: EdgeSoftmaxFn returns d(output_var, lengthscale)
            kparams = torch.cat([self.output_var, self.lengthscale]).to(torch.float32).contiguous()
        if att_type in self._UNIT_TYPES and torch.is_grad_enabled() and qk.requires_grad:
            # training: the VJP kernels implement scaled_dot; cosine_sim / pearson are scaled_dot of unit vectors
            qk = unit_heads(qk, self.h, centre=att_type == _lib.GB_ATT['pearson'])
            att_type = _lib.GB_ATT['scaled_dot']
        if self.square_plus:
            return self._squareplus(qk, plan, att_type, edge_w, kparams)
        return EdgeSoftmaxFn.apply(qk, plan, self.h, att_type, int(self.opt['attention_norm_idx']), edge_w, kparams)

    def _squareplus(self, qk, plan, att_type, edge_w, kparams):
        """scores from the edge kernel (EdgeScoreFn, differentiable), squareplus normalisation as [E,H] torch glue"""
        prods = EdgeScoreFn.apply(qk, plan, self.h, att_type, edge_w, kparams)
        att, abar = squareplus_csr(prods, plan, int(self.opt['attention_norm_idx']))
        return att, abar, prods.detach()

    def forward(self, x, edge, y=None):
        """returns (attention [B,E,H] in the caller's edge order, (v, prods)); v is None because the
        value projection is dead code unless mix_features (see module docstring)."""
        batched = x.dim() == 3
        x2 = ODEFunc._as_2d(x)
        plan = _plan.get_plan(edge, x2.shape[0])
        att, _abar, prods = self.attention_csr(x2, plan)
        att = ToCooOrderFn.apply(att, plan)
        prods = plan.to_coo_order(prods)
        if batched:
            att, prods = att[None], prods[None]
        return att, (None, prods)

    def __repr__(self):
        return self.__class__.__name__ + ' (' + str(self.in_features) + ' -> ' + str(self.out_features) + ')'


class ODEFuncTransformerAtt(ODEFunc):
    """function_transformer_attention.py:9-62"""

    def __init__(self, in_features, out_features, opt, device):
        super(ODEFuncTransformerAtt, self).__init__(opt, device)
        self.device = device
        self.in_features = in_features
        self.out_features = out_features
        self.multihead_att_layer = SpGraphTransAttentionLayer(in_features, out_features, opt, device,
                                                              edge_weights=self.edge_weight).to(device)
        self.y = None

    def forward(self, t, x):  # t is needed when called by the integrator
        self._check_nfe()
        if self.opt.get('mix_features', False):
            raise NotImplementedError("opt['mix_features'] is broken in the reference "
                                      '(function_transformer_attention.py:23-32,49-50) and not implemented')
        x2 = self._as_2d(x)
        if self.partition is not None:
            return self._partitioned(x, x2, None)
        plan = self._plan(x2.shape[0])
        source = bool(self.opt['add_source'])
        x0 = self._as_2d(self.x0) if (source and self.x0 is not None) else None
        L = self.multihead_att_layer
        att_type = L._check_supported()
        edge_w = L.edge_weights if self.opt.get('reweight_attention', False) else None
        if (not L.square_plus and nl_supported(plan, x2, L.h, L.attention_dim, att_type, int(self.opt['attention_norm_idx']), edge_w)):
            f = NlTransformerFn.apply(x2, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, self.alpha_train,
                                      self.beta_train if x0 is not None else None, x0, plan, L.h,
                                      not self.opt['no_alpha_sigmoid'], self._take_stage())
        elif not L.needs_composite_backward(att_type, x2) and fused_transformer_supported(
                plan, x2, L.h, L.attention_dim, att_type, int(self.opt['attention_norm_idx']), edge_w):
            kparams = None
            if att_type == _lib.GB_ATT['exp_kernel']:
                kparams = torch.cat([L.output_var.detach(), L.lengthscale.detach()]).to(torch.float32).contiguous()
            f = FusedTransformerFn.apply(x2, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, self.alpha_train,
                                         self.beta_train if x0 is not None else None, x0, plan, L.h, att_type,
                                         not self.opt['no_alpha_sigmoid'], kparams, L.qk_weights(), self._take_stage())
        else:
            _att, abar, _prods = L.attention_csr(x2, plan)
            f = DiffusionFn.apply(x2, abar, self.alpha_train, self.beta_train if x0 is not None else None, x0, plan,
                                  not self.opt['no_alpha_sigmoid'], self._take_stage())
        return f.view_as(x)

    def _partitioned(self, x, x2, vjp):
        """row-partitioned evaluation (partition.py): x2 is this rank's [S, d] slab.  vjp = None: autograd
        forward; vjp = (cot, params): fused f + VJP for the adjoint solver."""
        from . import partition as P
        part, halo = self.partition
        L = self.multihead_att_layer
        att_type = L._check_supported()
        edge_w = L.edge_weights if self.opt.get('reweight_attention', False) else None
        if x2.shape[0] != part.slab:
            raise ValueError(f'partitioned state must be the [{part.slab}, d] slab of this rank, got {tuple(x2.shape)}')
        plan = self._plan(part.n_pad)
        if not (not L.square_plus and nl_supported(plan, x2, L.h, L.attention_dim, att_type, int(self.opt['attention_norm_idx']), edge_w)):
            raise NotImplementedError('the row-partitioned path supports scaled_dot attention with attention_norm_idx 0, '
                                      'heads in {1,2,4,8}, d <= 128 (single-gather kernels)')
        source = bool(self.opt['add_source'])
        x0 = self._as_2d(self.x0).contiguous() if (source and self.x0 is not None) else None
        sig = not self.opt['no_alpha_sigmoid']
        if vjp is None:
            f = P.PartNlFn.apply(x2, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, self.alpha_train,
                                 self.beta_train if x0 is not None else None, x0, part, halo, plan, L.h, sig)
            return f.view_as(x)
        cot, params = vjp
        self._check_nfe()
        a, d = L.Q.weight.shape
        f, dx, grads = P.nl_eval_vjp_part(part, halo, plan, x2.detach(), self._as_2d(cot).detach(),
                                          L.Q.weight.detach().contiguous(), L.Q.bias.detach().contiguous(),
                                          L.K.weight.detach().contiguous(), self.alpha_train.detach(),
                                          self.beta_train.detach() if x0 is not None else None, x0, L.h, sig)
        halo.all_reduce_(grads)
        dwq, dwk, dbq, dbk, dalpha, dbeta = P.unpack_nl_grads(grads, a, d)
        named = ((self.alpha_train, dalpha.reshape(self.alpha_train.shape)), (self.beta_train, dbeta.reshape(())),
                 (L.Q.weight, dwq), (L.Q.bias, dbq), (L.K.weight, dwk), (L.K.bias, dbk))
        out = []
        for p in params:
            g = None
            for q, gq in named:
                if p is q:
                    g = gq
            out.append(g)
        return f.view_as(x), dx.view_as(x), out

    # ---- two-phase evaluation + VJP for the row-partitioned adjoint solve ----------------------------------------
    # The next Runge-Kutta stage's state depends only on f, which the ROW pass produces; its all-gather can therefore
    # run while the column pass, the dense tail and the reduce-scatter of this evaluation are still at work
    # (DESIGN.md "Multi-GPU").  integrators._adjoint_fixed_phased drives this protocol when phased_ok() says so.
    def phased_ok(self, x):
        """two-phase protocol available?  Row-partitioned, single-gather kernels, and not switched off: the overlap is
        verified over NCCL (profiles/r02_partition_nccl_n2.log) and on by default (GRANDB200_OVERLAP=0 turns it off)"""
        import os
        if self.partition is None or os.environ.get('GRANDB200_OVERLAP', '1') != '1' or self.opt.get('mix_features', False):
            return False
        part, _halo = self.partition
        L = self.multihead_att_layer
        x2 = self._as_2d(x)
        edge_w = L.edge_weights if self.opt.get('reweight_attention', False) else None
        plan = self._plan(part.n_pad)
        return x2.shape[0] == part.slab and not L.square_plus and nl_supported(plan, x2, L.h, L.attention_dim, L._check_supported(),
                                                         int(self.opt['attention_norm_idx']), edge_w)

    def prefetch_state(self, x):
        """start gathering the slabs of a state that a later eval_vjp_phased(…, prefetched=) call will evaluate"""
        part, halo = self.partition
        halo.prepare(self._plan(part.n_pad))
        pending = halo.all_gather_rows_async(self._as_2d(x).detach())
        pending.src = x                                            # eval_vjp_phased checks identity against its `x`
        return pending

    def eval_vjp_phased(self, t, x, cot, params, prefetched=None):
        """first half of eval_vjp on a row-partitioned graph: gather (or take the prefetched gather of) the state,
        project q, run the row pass.  Returns (f, finish): f is this rank's slab of f(t, x); finish() runs the column
        pass, the dense tail and the exchanges and returns (vjp_x, [vjp_p or None ...]) like eval_vjp."""
        from . import partition as P
        from .ops import _nl_q
        part, halo = self.partition
        L = self.multihead_att_layer
        x2 = self._as_2d(x).detach().contiguous()
        self._check_nfe()
        halo.prepare(self._plan(part.n_pad))
        if prefetched is not None and prefetched.src is x:
            x_full = prefetched.wait()
        else:
            if prefetched is not None:
                prefetched.wait()                                  # a gather of some other state: complete it, drop it
            x_full = halo.all_gather_rows(x2)
        plan = self._plan(part.n_pad)
        source = bool(self.opt['add_source'])
        x0 = self._as_2d(self.x0).contiguous() if (source and self.x0 is not None) else None
        wq, bq, wk = (L.Q.weight.detach().contiguous(), L.Q.bias.detach().contiguous(), L.K.weight.detach().contiguous())
        q = _nl_q(x2, wq, bq)
        half = P.nl_local_vjp_rows(part, plan, x_full, q, wq, wk, self.alpha_train.detach(),
                                   self.beta_train.detach() if x0 is not None else None, x0, L.h,
                                   not self.opt['no_alpha_sigmoid'], self._as_2d(cot).detach().contiguous(), True,
                                   dx_out=P.share_buffer(halo, x_full))
        a, d = wq.shape

        def finish():
            dx_share = P.nl_local_vjp_cols(half)
            pending = halo.reduce_scatter_rows_async(dx_share)      # runs under the weight-gradient GEMM
            grads = P.nl_local_vjp_weights(half)
            dx = pending.wait()
            halo.all_reduce_(grads)
            dwq, dwk, dbq, dbk, dalpha, dbeta = P.unpack_nl_grads(grads, a, d)
            named = ((self.alpha_train, dalpha.reshape(self.alpha_train.shape)), (self.beta_train, dbeta.reshape(())),
                     (L.Q.weight, dwq), (L.Q.bias, dbq), (L.K.weight, dwk), (L.K.bias, dbk))
            out = []
            for p in params:
                g = None
                for q_, gq in named:
                    if p is q_:
                        g = gq
                out.append(g)
            return dx.view_as(x), out

        return half.f.view_as(x), finish

    def eval_vjp(self, t, x, cot, params):
        """f(t, x) and the vector-Jacobian products cot^T df/d(x, params) in one fused pass - what the adjoint
        solver's augmented dynamics need (the reference gets them from `func(t, y)` followed by
        `autograd.grad`, src/base_classes.py:45-48 via torchdiffeq).  `params`: the tensors the solver wants
        gradients for; returns (f, vjp_x, [vjp_p or None ...]) or None when the fused path does not apply
        (the solver then falls back to autograd on forward())."""
        if self.opt.get('mix_features', False):
            return None
        x2 = self._as_2d(x)
        if self.partition is not None:
            return self._partitioned(x, x2, (cot, params))
        plan = self._plan(x2.shape[0])
        L = self.multihead_att_layer
        att_type = L._check_supported()
        edge_w = L.edge_weights if self.opt.get('reweight_attention', False) else None
        if not (not L.square_plus and nl_supported(plan, x2, L.h, L.attention_dim, att_type, int(self.opt['attention_norm_idx']), edge_w)):
            return None
        self._check_nfe()
        source = bool(self.opt['add_source'])
        x0 = self._as_2d(self.x0) if (source and self.x0 is not None) else None
        named = {'alpha_train': self.alpha_train, 'beta_train': self.beta_train, 'Q.weight': L.Q.weight,
                 'Q.bias': L.Q.bias, 'K.weight': L.K.weight, 'K.bias': L.K.bias}
        want_w = any(p is q for p in params for q in (L.Q.weight, L.Q.bias, L.K.weight, L.K.bias))
        f, dx, grads = nl_eval_vjp(x2, self._as_2d(cot), L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, self.alpha_train,
                                   self.beta_train if x0 is not None else None, x0, plan, L.h,
                                   not self.opt['no_alpha_sigmoid'], want_w=want_w)
        out = []
        for p in params:
            g = None
            for name, q in named.items():
                if p is q:
                    g = grads.get(name)
            out.append(g)
        return f.view_as(x), dx.view_as(x), out

    def __repr__(self):
        return self.__class__.__name__ + ' (' + str(self.in_features) + ' -> ' + str(self.out_features) + ')'
