"""`ODEFuncAtt` / `SpGraphAttentionLayer` - drop-in for the reference's src/function_GAT_attention.py:9-149.

    f(t, x) = alpha * (A(x) x - x) + beta * x0,     A(x) = mean_h softmax_g(leaky_relu(a^T [W x_row || W x_col]))

The reference gathers both endpoints' [dk, H] projections for every edge, multiplies with the shared vector `a`
([1, 2 dk, 1, 1]) and sums over dk, then scatters the attention into a dense [N, N, H] tensor and multiplies.
The score is a SUM of a row term and a column term, so it factorises into per-node scalars:

    s_src[i,h] = <a[:dk], (x W)[i,h,:]> = x[i,:] . (W_h a[:dk])        s_dst[j,h] = x[j,:] . (W_h a[dk:])

one [N,d] x [d,2H] projection per evaluation (csrc/linear.cu), then ONE kernel (csrc/gat.cu) that gathers H floats
per edge, applies the leaky relu and the segment softmax and writes the head mean, then the fused SpMM + epilogue
of the Laplacian function.  Parameter names (`multihead_att_layer.W`, `.Wout`, `.a`) and shapes are the
reference's (there `nn.Parameter(...).to(device)` silently un-registers them on a GPU; here they are registered)."""
import torch
from torch import nn

from . import _lib
from . import plan as _plan
from .base_classes import ODEFunc
from .ops import QKProjFn, GatSoftmaxFn, DiffusionFn, ToCooOrderFn


class SpGraphAttentionLayer(nn.Module):
    """function_GAT_attention.py:73-149"""

    def __init__(self, in_features, out_features, opt, device, concat=True):
        super(SpGraphAttentionLayer, self).__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.alpha = opt['leaky_relu_slope']
        self.concat = concat
        self.device = device
        self.opt = opt
        self.h = int(opt['heads'])
        try:
            self.attention_dim = opt['attention_dim']
        except KeyError:
            self.attention_dim = out_features
        assert self.attention_dim % self.h == 0, "Number of heads must be a factor of the dimension size"
        self.d_k = self.attention_dim // self.h
        self.W = nn.Parameter(torch.zeros(size=(in_features, self.attention_dim)))
        nn.init.xavier_normal_(self.W.data, gain=1.414)
        self.Wout = nn.Parameter(torch.zeros(size=(self.attention_dim, self.in_features)))
        nn.init.xavier_normal_(self.Wout.data, gain=1.414)
        self.a = nn.Parameter(torch.zeros(size=(1, 2 * self.d_k, 1, 1)))
        nn.init.xavier_normal_(self.a.data, gain=1.414)
        if opt.get('multi_modal', False):
            raise NotImplementedError("opt['multi_modal'] is out of scope (SURVEY.md 2.1)")

    def score_weights(self):
        """[2H, d]: rows h < H map x to s_src[:, h] (W_h a[:dk]), rows H + h to s_dst[:, h] (W_h a[dk:])"""
        d = self.W.shape[0]
        w = self.W.view(d, self.h, self.d_k)
        a = self.a.reshape(2, self.d_k)
        return torch.cat([torch.einsum('dhc,c->hd', w, a[0]), torch.einsum('dhc,c->hd', w, a[1])], dim=0)

    def attention_csr(self, x2, plan):
        """(att [E,H], abar [E], prods [E,H]) in CSR-slot order for an un-batched state x2 [N,d]"""
        m = self.score_weights()
        zero = torch.zeros(self.h, dtype=m.dtype, device=m.device)
        s = QKProjFn.apply(x2, m[:self.h], zero, m[self.h:], zero)            # [N, 2H] = [s_src | s_dst]
        return GatSoftmaxFn.apply(s, plan, self.h, float(self.alpha), int(self.opt['attention_norm_idx']))

    def forward(self, x, edge, y=None):
        """returns (attention [B,E,H] in the caller's edge order, wx = x W) like the reference"""
        batched = x.dim() == 3
        x2 = ODEFunc._as_2d(x)
        plan = _plan.get_plan(edge, x2.shape[0])
        att, _abar, _prods = self.attention_csr(x2, plan)
        att = ToCooOrderFn.apply(att, plan)
        wx = torch.matmul(x, self.W)
        return (att[None] if batched else att), wx

    def __repr__(self):
        return self.__class__.__name__ + ' (' + str(self.in_features) + ' -> ' + str(self.out_features) + ')'


class ODEFuncAtt(ODEFunc):
    """function_GAT_attention.py:9-70"""

    def __init__(self, in_features, out_features, opt, device):
        super(ODEFuncAtt, self).__init__(opt, device)
        self.device = device
        self.in_features = in_features
        self.out_features = out_features
        self.multihead_att_layer = SpGraphAttentionLayer(in_features, out_features, opt, device).to(device)
        try:
            self.attention_dim = opt['attention_dim']
        except KeyError:
            self.attention_dim = out_features
        assert self.attention_dim % opt['heads'] == 0, "Number of heads must be a factor of the dimension size"
        self.d_k = self.attention_dim // opt['heads']
        self.y = None

    def forward(self, t, x):  # t is needed when called by the integrator
        self._check_nfe()
        if self.opt.get('mix_features', False):
            raise NotImplementedError("opt['mix_features'] (attention applied to x W, then Wout; "
                                      'function_GAT_attention.py:28-38) is not implemented')
        if self.partition is not None:
            raise NotImplementedError("function 'GAT' has no row-partitioned path")
        x2 = self._as_2d(x)
        plan = self._plan(x2.shape[0])
        _att, abar, _prods = self.multihead_att_layer.attention_csr(x2, plan)
        source = bool(self.opt['add_source'])
        x0 = self._as_2d(self.x0) if (source and self.x0 is not None) else None
        f = DiffusionFn.apply(x2, abar, self.alpha_train, self.beta_train if x0 is not None else None, x0, plan,
                              not self.opt['no_alpha_sigmoid'], None)
        return f.view_as(x)

    def __repr__(self):
        return self.__class__.__name__ + ' (' + str(self.in_features) + ' -> ' + str(self.out_features) + ')'
