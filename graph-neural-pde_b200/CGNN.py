"""`CGNN` - drop-in for the model classes of the reference's src/CGNN.py (ODEFunc :68-91, ODEblock :94-115,
CGNN :119-167, get_sym_adj / coo2tensor :269-284): Continuous GNN (Xhonneux et al. 2019) on the PyG data pipeline.

    f(t, x) = sigmoid(alpha_train)[:, None] * 0.5 * (A x - x) + x0          (A: sym-normalised adjacency, fixed)

The reference evaluates `torch.spmm(self.adj, x)` on a sparse COO tensor (src/CGNN.py:89).  Here the adjacency is
turned into a graph plan once and every evaluation is the fused CSR SpMM kernel (`gb_spmm_fwd`, epilogue `A x - x`;
VJP `gb_spmm_bwd`); the per-node alpha and the source term are two elementwise torch ops on the device - the alpha
of this model is a vector over NODES ([n], `adj.shape[1]`), not the scalar of the GRAND functions the kernel's
epilogue takes.  Encoder / decoder: `ops.Linear`.  Parameters (`alpha_train` [n], `w`, `d`, `m1`, `m2`) keep the
reference's names and shapes; `w` and `d` are never read, as upstream.  The training script of the file (ray tune,
argparse, dataset download) is out of scope.
"""
import torch
from torch import nn
import torch.nn.functional as F

from . import integrators, ops, plan as _plan
from .utils import Meter, gcn_norm_fill_val


def _adj_parts(adj):
    """(edge_index [2,E] int64, values [E] f32, n) of a torch sparse COO adjacency, or of an (edge_index, values, n)
    triple - duplicates are kept (spmm sums them, and so does the kernel)"""
    if isinstance(adj, (tuple, list)):
        ei, w, n = adj
        return ei.long(), w.float(), int(n)
    if not adj.is_sparse:
        raise TypeError('adj must be a torch sparse COO tensor (coo2tensor) or an (edge_index, values, n) triple')
    if adj.shape[0] != adj.shape[1]:
        raise ValueError(f'adjacency must be square, got {tuple(adj.shape)}')
    return adj._indices().long(), adj._values().float(), int(adj.shape[0])


class ODEFunc(nn.Module):
    """src/CGNN.py:68-91 (requires in_features == out_features, like the reference)"""

    def __init__(self, in_features, out_features, opt, adj):
        super(ODEFunc, self).__init__()
        self.opt = opt
        self.adj = adj
        self.x0 = None
        self.nfe = 0
        self.in_features = in_features
        self.out_features = out_features
        self.alpha = opt['alpha']
        ei, w, n = _adj_parts(adj)
        self.edge_index, self.edge_weight, self.num_nodes = ei, w, n
        self.alpha_train = nn.Parameter(self.alpha * torch.ones(n))
        self.w = nn.Parameter(torch.eye(opt['hidden_dim']))
        self.d = nn.Parameter(torch.zeros(opt['hidden_dim']) + 1)
        self._one = None

    def forward(self, t, x):
        self.nfe += 1
        if self.edge_index.device != x.device:
            self.edge_index, self.edge_weight = self.edge_index.to(x.device), self.edge_weight.to(x.device)
        plan = _plan.get_plan(self.edge_index, self.num_nodes)
        vals = _plan.cached_per_edge(plan, self.edge_weight, 'w', lambda w_: plan.to_csr_order(w_.detach()))
        if self._one is None or self._one.device != x.device:
            self._one = torch.ones(1, dtype=torch.float32, device=x.device)
        # alpha_sigmoid = False with alpha_train = 1, no source:  A x - x  in one kernel
        ax_x = ops.DiffusionFn.apply(x, vals, self._one, None, None, plan, False, None)
        alph = torch.sigmoid(self.alpha_train).unsqueeze(dim=1)
        return alph * 0.5 * ax_x + self.x0


class ODEblock(nn.Module):
    """src/CGNN.py:94-115: plain (non-adjoint) odeint with the solver defaults (dopri5, rtol 1e-7, atol 1e-9)"""

    def __init__(self, odefunc, t=torch.tensor([0, 1])):
        super(ODEblock, self).__init__()
        self.t = t
        self.odefunc = odefunc
        self.nfe = 0

    def set_x0(self, x0):
        self.odefunc.x0 = x0.clone().detach()

    def forward(self, x):
        t = self.t.type_as(x)
        z = integrators.odeint(self.odefunc, x, t)[1]
        self.nfe += self.odefunc.nfe
        self.odefunc.nfe = 0
        return z

    def __repr__(self):
        return self.__class__.__name__ + '( Time Interval ' + str(self.t[0].item()) + ' -> ' + str(self.t[1].item()) + ")"


class CGNN(nn.Module):
    """src/CGNN.py:119-167"""

    def __init__(self, opt, adj, time, device):
        super(CGNN, self).__init__()
        self.opt = opt
        self.adj = adj
        self.T = time
        self.fm = Meter()
        self.bm = Meter()
        self.device = device
        self.m1 = ops.Linear(opt['num_feature'], opt['hidden_dim'])
        self.odeblock = ODEblock(ODEFunc(2 * opt['hidden_dim'], 2 * opt['hidden_dim'], opt, adj),
                                 t=torch.tensor([0, self.T]))
        self.m2 = ops.Linear(opt['hidden_dim'], opt['num_class'])

    def getNFE(self):
        return self.odeblock.odefunc.nfe

    def resetNFE(self):
        self.odeblock.odefunc.nfe = 0

    def reset(self):
        self.m1.reset_parameters()
        self.m2.reset_parameters()

    def forward(self, x, pos_enc=None):
        x = F.dropout(x, self.opt['input_dropout'], training=self.training)
        x = self.m1(x)
        c_aux = torch.zeros(x.shape, device=x.device)            # augmentation: zeros of the same width
        x = torch.cat([x, c_aux], dim=1)
        self.odeblock.set_x0(x)
        z = self.odeblock(x)
        z = torch.split(z, x.shape[1] // 2, dim=1)[0]
        z = F.relu(z)
        z = F.dropout(z, self.opt['dropout'], training=self.training)
        return self.m2(z)


def get_sym_adj(data, opt, device):
    """src/CGNN.py:280-284: gcn-normalised adjacency with self loops of weight opt['self_loop_weight'], as the
    (edge_index, values, n) triple ODEFunc takes - the reference goes through scipy to a sparse COO tensor, which
    holds the same three things"""
    edge_index, edge_weight = gcn_norm_fill_val(data.edge_index, data.edge_attr, opt['self_loop_weight'], data.num_nodes,
                                                dtype=data.x.dtype)
    return edge_index.to(device), edge_weight.to(device), int(data.num_nodes)
