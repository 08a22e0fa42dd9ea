"""Mirror of the reference's `src/base_classes.py` for the hot path: `ODEFunc` (plugin state
container, base_classes.py:116-134), `ODEblock` (base_classes.py:33-98) and `GraphData`
(base_classes.py:100-113).  Same constructor signatures, attribute names and parameter names
(`alpha_train`, `beta_train`, `alpha_sc`, `beta_sc`) so reference checkpoints load with
strict=True and the reference's blocks / solvers can drive these classes unchanged.

Differences, all deliberate:
  * `ODEFunc` derives from nn.Module (the reference derives from PyG MessagePassing only for
    nn.Module-ness; no propagate() call exists).
  * graph normalisation goes through the un-broken helpers in utils.py and the result is cached
    while `data.edge_index` is the same tensor (the reference redoes it every block forward).
  * integration uses torchdiffeq when it is importable, else the restated solvers in
    integrators.py (same call signature).
"""
import torch
from torch import nn

from . import ops, plan as _plan
from .utils import get_rw_adj, gcn_norm_fill_val, add_remaining_self_loops, Meter


def _integrators(adjoint):
    try:
        if adjoint:
            from torchdiffeq import odeint_adjoint as odeint
        else:
            from torchdiffeq import odeint
        return odeint
    except ImportError:
        from . import integrators
        return integrators.odeint_adjoint if adjoint else integrators.odeint


REG_KEYS = ('kinetic_energy', 'jacobian_norm2', 'total_deriv', 'directional_penalty')


def create_regularization_fns(args):
    """base_classes.py:19-30.  The RNODE regularisers need double backward through the ODE
    function (regularized_ODE_function.py:39-75), which the CUDA operators do not provide."""
    active = [k for k in REG_KEYS if args.get(k) is not None]
    if active:
        raise NotImplementedError(f'regularisation terms {active} need create_graph=True through the ODE '
                                  'function; the B200 operators implement first-order VJPs only')
    return [], []


class ODEFunc(nn.Module):
    """base_classes.py:116-134"""

    def __init__(self, opt, device):
        super(ODEFunc, self).__init__()
        self.opt = opt
        self.device = device
        self.edge_index = None
        self.edge_weight = None
        self.attention_weights = None
        self.alpha_train = nn.Parameter(torch.tensor(0.0))
        self.beta_train = nn.Parameter(torch.tensor(0.0))
        self.x0 = None
        self.nfe = 0
        self.alpha_sc = nn.Parameter(torch.ones(1))
        self.beta_sc = nn.Parameter(torch.ones(1))
        # (RowPartition, Halo) when the graph is split by rows over several GPUs (partition.py); the
        # state is then this rank's [S, d] slab and edge_index holds the edges whose row it owns
        self.partition = None

    # `attention_weights` is a plain attribute for every caller (the blocks assign it, LaplacianODEFunc and users read
    # it); a block may also park a zero-argument callable with `defer_attention_weights`, evaluated on the first read.
    # AttODEblock does that when opt['function'] == 'transformer': the reference computes the block-level attention
    # there too (block_transformer_attention.py:24-32) but nothing on the path reads it.
    @property
    def attention_weights(self):
        thunk = self.__dict__.get('_attention_thunk')
        if thunk is not None:
            self.__dict__['_attention_thunk'] = None
            self.__dict__['_attention_weights'] = thunk()
        return self.__dict__.get('_attention_weights')

    @attention_weights.setter
    def attention_weights(self, value):
        self.__dict__['_attention_thunk'] = None
        self.__dict__['_attention_weights'] = value

    def defer_attention_weights(self, thunk):
        self.__dict__['_attention_weights'] = None
        self.__dict__['_attention_thunk'] = thunk

    # ---- Runge-Kutta stage fusion (north_star (a)) ----------------------------------------------
    # The fixed-grid solvers (integrators._fixed_step) form the next stage's state  y + sum_j c_j k_j + c f(t, x)  in
    # the epilogue of the kernel that produces f instead of a separate pass over the state: they call forward_stage,
    # which parks a `gb_stage` descriptor, runs forward() and reports whether the route taken consumed it.  Only
    # without autograd (inference, and the forward solve of odeint_adjoint, which runs under no_grad).
    _pending_stage = None

    def forward_stage(self, t, x, y, prev, coeffs, c_self):
        """(f(t, x), y + sum_j coeffs[j] * prev[j] + c_self * f(t, x)); the second value is None when the evaluation
        could not fuse the combination (the caller then forms it itself)."""
        import ctypes as C
        from . import _lib
        ok = (not torch.is_grad_enabled() and self.partition is None and x.is_cuda and x.dtype == torch.float32
              and x.is_contiguous() and len(prev) <= 6 and x.dim() in (2, 3)
              and all(v.shape == x.shape and v.dtype == x.dtype and v.is_contiguous() for v in [y] + list(prev)))
        if not ok:
            return self.forward(t, x), None
        ynext = torch.empty_like(x)
        st = _lib.GbStage()
        st.y = y.data_ptr()
        for j, (k, c) in enumerate(zip(prev, coeffs)):
            st.k[j] = k.data_ptr()
            st.c[j] = float(c)
        st.c[len(prev)] = float(c_self)
        st.n_prev, st.write_f, st.ynext, st.ld = len(prev), 1, ynext.data_ptr(), x.shape[-1]
        self._pending_stage = st
        try:
            f = self.forward(t, x)
            fused = self._pending_stage is None
        finally:
            self._pending_stage = None
        return f, (ynext if fused else None)

    def _take_stage(self):
        """the parked stage descriptor (as the ctypes argument of the kernel entry points), at most once"""
        st = self._pending_stage
        if st is None:
            return None
        import ctypes as C
        self._pending_stage = None
        self._stage_keepalive = st               # the launch reads the struct by value; keep it until the next one
        return C.pointer(st)

    # ---- helpers shared by the two ODE functions ------------------------------------------------
    @staticmethod
    def _as_2d(x):
        if x.dim() == 3:
            if x.shape[0] != 1:
                raise NotImplementedError('batched states (B > 1) are evaluated one graph at a time; pass B == 1')
            return x[0]
        if x.dim() != 2:
            raise ValueError(f'state must be [N,d] or [1,N,d], got {tuple(x.shape)}')
        return x

    def _plan(self, num_nodes):
        if self.edge_index is None:
            raise RuntimeError('edge_index has not been set on the ODE function (ODEblock.reset_graph_data)')
        return _plan.get_plan(self.edge_index, num_nodes)

    def _check_nfe(self):
        from .utils import MaxNFEException
        if self.nfe > self.opt['max_nfe']:          # function_laplacian_diffusion.py:65-67
            raise MaxNFEException
        self.nfe += 1

    def __repr__(self):
        return self.__class__.__name__


class ODEblock(nn.Module):
    """base_classes.py:33-98"""

    def __init__(self, odefunc, regularization_fns, opt, device, t):
        super(ODEblock, self).__init__()
        self.opt = opt
        self.t = t
        self.device = device
        self.aug_dim = 2 if opt['augment'] else 1
        self.odefunc = odefunc(self.aug_dim * opt['hidden_dim'], self.aug_dim * opt['hidden_dim'], opt, device)
        self.nreg = len(regularization_fns)
        # the reference builds the function twice: `reg_odefunc` keeps the FIRST instance (base_classes.py:40-43) and
        # every concrete block then replaces `self.odefunc` with a second one (e.g. block_constant.py:11), so its
        # state_dict carries `reg_odefunc.odefunc.*` next to `odefunc.*`; same keys, same parameter count here
        self.reg_odefunc = RegularizedODEfunc(odefunc(self.aug_dim * opt['hidden_dim'], self.aug_dim * opt['hidden_dim'],
                                                      opt, device), regularization_fns)
        self.train_integrator = _integrators(opt['adjoint'])
        self.test_integrator = None
        self.set_tol()
        self._graph_cache = None
        self.partition = None

    def set_partition(self, part, halo):
        """split the graph by rows over the ranks of `halo` (partition.py): forward() then takes and returns
        this rank's [S, d] slab of the state; `graph_data` stays the GLOBAL graph on every rank."""
        self.partition = (part, halo)
        self.odefunc.partition = self.partition
        self._graph_cache = None

    def set_x0(self, x0):
        self.odefunc.x0 = x0.clone().detach()
        self.reg_odefunc.odefunc.x0 = self.odefunc.x0          # base_classes.py:53-55 (the reference clones twice)

    def set_tol(self):
        self.atol = self.opt['tol_scale'] * 1e-7
        self.rtol = self.opt['tol_scale'] * 1e-9
        if self.opt['adjoint']:
            self.atol_adjoint = self.opt['tol_scale_adjoint'] * 1e-7
            self.rtol_adjoint = self.opt['tol_scale_adjoint'] * 1e-9

    def reset_tol(self):
        self.atol = 1e-7
        self.rtol = 1e-9
        self.atol_adjoint = 1e-7
        self.rtol_adjoint = 1e-9

    def reset_graph_data(self, data, dtype, y=None):
        """base_classes.py:70-90; the normalised graph is reused while data.edge_index /
        data.edge_attr are the same tensors (identity + version)."""
        if data is not None:
            ei, ea = data.edge_index, data.edge_attr
            key = (id(ei), ei._version, id(ea), None if ea is None else ea._version, data.num_nodes,
                   self.opt['data_norm'], self.opt['self_loop_weight'])
            if self._graph_cache is None or self._graph_cache[0] != key or self._graph_cache[1] is not ei:
                self.num_nodes = data.num_nodes
                ei_d = ei.to(self.device)
                ea_d = None if ea is None else ea.to(self.device)
                if self.opt['data_norm'] == 'rw':
                    edge_index, edge_weight = get_rw_adj(ei_d, edge_weight=ea_d, norm_dim=1,
                                                         fill_value=self.opt['self_loop_weight'],
                                                         num_nodes=data.num_nodes, dtype=dtype)
                else:
                    edge_index, edge_weight = gcn_norm_fill_val(ei_d, edge_weight=ea_d,
                                                                fill_value=self.opt['self_loop_weight'],
                                                                num_nodes=data.num_nodes, dtype=dtype)
                if self.opt['self_loop_weight'] > 0:
                    edge_index, edge_weight = add_remaining_self_loops(
                        edge_index, edge_weight, fill_value=self.opt['self_loop_weight'], num_nodes=data.num_nodes)
                if self.partition is not None:               # keep the edges whose row this rank owns
                    batched = edge_index.dim() == 3
                    part, halo = self.partition
                    edge_index, edge_weight = part.local_edges(edge_index, edge_weight)
                    if hasattr(halo, 'localize'):             # CompactHalo: local ids, [own slab | referenced rows]
                        edge_index = halo.localize(edge_index)
                        self.odefunc.partition = (halo.view, halo)
                    if batched:
                        edge_index, edge_weight = edge_index[None], edge_weight[None]
                self._graph_cache = (key, ei, edge_index, edge_weight)
            _, _, edge_index, edge_weight = self._graph_cache
            self.data_edge_index = edge_index
            self.odefunc.edge_index = edge_index
            self.odefunc.edge_weight = edge_weight
        self.odefunc.y = y

    def set_time(self, time):
        self.t = torch.tensor([0, time]).to(self.device)

    def _integrate(self, x):
        """the integrator call shared by ConstantODEblock / AttODEblock (block_constant.py:34-58)"""
        integrator = self.train_integrator if self.training else self.test_integrator
        # the restated solvers read the time grid on the host: keep it a CPU tensor (no device sync per pass,
        # which also makes the pass capturable in a CUDA graph); torchdiffeq wants it next to the state
        ours = getattr(integrator, '__module__', '').endswith('integrators')
        if ours:
            host = getattr(self, '_t_host', None)
            if host is None or host[0] is not self.t:              # one device read per time grid, then cached
                host = self._t_host = (self.t, self.t.detach().cpu())
            t = host[1].to(x.dtype)
        else:
            t = self.t.type_as(x)
        func = self.odefunc
        opt = self.opt
        if opt['adjoint'] and self.training:
            state_dt = integrator(
                func, x, t, method=opt['method'],
                options=dict(step_size=opt['step_size'], max_iters=opt['max_iters']),
                adjoint_method=opt['adjoint_method'],
                adjoint_options=dict(step_size=opt['adjoint_step_size'], max_iters=opt['max_iters']),
                atol=self.atol, rtol=self.rtol, adjoint_atol=self.atol_adjoint, adjoint_rtol=self.rtol_adjoint)
        else:
            state_dt = integrator(
                func, x, t, method=opt['method'],
                options=dict(step_size=opt['step_size'], max_iters=opt['max_iters']),
                atol=self.atol, rtol=self.rtol)
        return state_dt[1]

    def __repr__(self):
        return self.__class__.__name__ + '( Time Interval ' + str(self.t[0].item()) + ' -> ' + str(self.t[1].item()) \
            + ")"


class RegularizedODEfunc(nn.Module):
    """regularized_ODE_function.py:8-35: holds its own instance of the ODE function (parameters registered under
    `reg_odefunc.odefunc.*`, as in the reference's checkpoints) and the regularisation functions.  Integrating through
    it needs create_graph=True through the function (double backward), which the B200 operators do not offer: the
    blocks refuse non-empty regularisation lists (`_no_regularisers`), so forward() is never reached on a valid
    configuration."""

    def __init__(self, odefunc, regularization_fns):
        super(RegularizedODEfunc, self).__init__()
        self.odefunc = odefunc
        self.regularization_fns = regularization_fns

    def forward(self, t, state):
        raise NotImplementedError('RNODE regularisers need double backward through the ODE function; the B200 '
                                  'operators implement first-order VJPs only')

    @property
    def _num_evals(self):
        return self.odefunc._num_evals


class GraphData(object):
    """base_classes.py:100-113"""

    def __init__(self):
        super(GraphData, self).__init__()
        self.edge_index = None
        self.edge_attr = None
        self.num_nodes = None

    def new_graph(self, edge_index, num_nodes, edge_attr=None):
        self.edge_index = edge_index
        self.edge_attr = edge_attr
        self.num_nodes = num_nodes

    def __repr__(self):
        return self.__class__.__name__


class BaseGNN(nn.Module):
    """base_classes.py:137-186 - encoder / decoder around the ODE block: `m1` (or `mx` + `mp` with opt['beltrami']),
    optional `m11` / `m12` MLP, `fc`, batch norms and the decoder `m2`.  Same parameter names and the same in-place
    edits of `opt['hidden_dim']` (beltrami, use_labels) as the reference, so its checkpoints load.  These are dense
    [N, F] x [F, d] layers executed once per step: `ops.Linear` (nn.Linear's parameters, the library's own GEMM kernels -
    K-sliced tcgen05 3xTF32 for the encoder, exact-fp32 tiles for the decoder)"""

    def __init__(self, opt, num_classes, num_features, device=torch.device('cpu')):
        super(BaseGNN, self).__init__()
        self.opt = opt
        self.T = opt['time']
        self.num_classes = num_classes
        self.num_features = num_features
        self.graph_data = GraphData()
        self.device = device
        self.fm = Meter()
        self.bm = Meter()
        if opt.get('beltrami', False):
            self.mx = ops.Linear(self.num_features, opt['feat_hidden_dim'])
            self.mp = ops.Linear(opt['pos_enc_dim'], opt['pos_enc_hidden_dim'])
            opt['hidden_dim'] = opt['feat_hidden_dim'] + opt['pos_enc_hidden_dim']
        else:
            self.m1 = ops.Linear(self.num_features, opt['hidden_dim'])
        if opt.get('use_mlp', False):
            self.m11 = ops.Linear(opt['hidden_dim'], opt['hidden_dim'])
            self.m12 = ops.Linear(opt['hidden_dim'], opt['hidden_dim'])
        if opt.get('use_labels', False):
            opt['hidden_dim'] = opt['hidden_dim'] + num_classes
        self.hidden_dim = opt['hidden_dim']
        if opt.get('fc_out', False):
            self.fc = ops.Linear(opt['hidden_dim'], opt['hidden_dim'])
        self.m2 = ops.Linear(opt['hidden_dim'], num_classes)
        if opt.get('batch_norm', False):
            self.bn_in = torch.nn.BatchNorm1d(opt['hidden_dim'])
            self.bn_out = torch.nn.BatchNorm1d(opt['hidden_dim'])
        self.regularization_fns, self.regularization_coeffs = create_regularization_fns(self.opt)

    def getNFE(self):
        return self.odeblock.odefunc.nfe + self.odeblock.reg_odefunc.odefunc.nfe

    def resetNFE(self):
        self.odeblock.odefunc.nfe = 0
        self.odeblock.reg_odefunc.odefunc.nfe = 0

    def reset(self):
        self.m1.reset_parameters()
        self.m2.reset_parameters()

    # ---- GNN.forward / GNNEarly.forward share everything but one line (GNN.py:17-74, GNN_early.py:42-105) ----------
    def _encode(self, x, pos_encoding):
        import torch.nn.functional as F
        opt = self.opt
        y = None
        if opt.get('use_labels', False):
            y = x[..., -self.num_classes:]
            x = x[..., :-self.num_classes]
        if opt.get('beltrami', False):
            x = self.mx(F.dropout(x, opt['input_dropout'], training=self.training))
            p = self.mp(F.dropout(pos_encoding, opt['input_dropout'], training=self.training))
            x = torch.cat([x, p], dim=-1)
        else:
            x = self.m1(F.dropout(x, opt['input_dropout'], training=self.training))
        if opt.get('use_mlp', False):
            x = F.dropout(x, opt['dropout'], training=self.training)
            x = F.dropout(x + self.m11(F.relu(x)), opt['dropout'], training=self.training)
            x = F.dropout(x + self.m12(F.relu(x)), opt['dropout'], training=self.training)
        if y is not None:
            x = torch.cat([x, y], dim=-1)
        if opt.get('batch_norm', False):
            x = self.bn_in(x.transpose(-1, -2)).transpose(-1, -2) if x.dim() == 3 else self.bn_in(x)
        if opt.get('augment', False):
            x = torch.cat([x, torch.zeros_like(x)], dim=-1)
        return x

    def _decode(self, z, width):
        import torch.nn.functional as F
        if self.opt.get('augment', False):
            z = torch.split(z, width // 2, dim=-1)[0]
        z = F.relu(z)
        if self.opt.get('fc_out', False):
            z = F.relu(self.fc(z))
        z = F.dropout(z, self.opt['dropout'], training=self.training)
        return self.m2(z)

    def _forward(self, x, graph_data, x2=None, pos_encoding=None):
        x = self._encode(x, pos_encoding)
        self.odeblock.set_x0(x)
        self.graph_data.new_graph(graph_data, x.shape[-2])
        if self.training and self.odeblock.nreg > 0:
            z, self.reg_states = self.odeblock(x, self.graph_data, x2)
        else:
            z = self.odeblock(x, self.graph_data, x2)
        return self._decode(z, x.shape[-1])

    def __repr__(self):
        return self.__class__.__name__
