"""`HardAttODEblock` - mirror of the reference's src/block_transformer_hard_attention.py:6-107: block-level
attention computed once at t = 0; while training only the edges whose head-mean attention lies above the
(1 - att_samp_pct) quantile are kept, their weights renormalised per group, and the ODE function integrates
over the SAMPLED graph (a new `edge_index` on the function -> a new GraphPlan, built by the CUDA plan builder
and cached by tensor identity); in eval mode the full graph with the head-mean attention is used.

The sampling (quantile, mask, renormalisation) is [E]-sized torch work under no_grad, once per block forward -
the reference does the same with torch.quantile and torch_scatter; the attention itself and every evaluation of
the integration run on the CUDA kernels.  RNODE regularisers (`nreg > 0`) need double backward and raise in
`create_regularization_fns`, so only the plain-state branch of the reference's forward exists here."""
import torch

from .base_classes import ODEblock, _integrators
from .function_transformer_attention import SpGraphTransAttentionLayer


def linear_quantile(values, q):
    """torch.quantile(values, q) (interpolation 'linear') without its 16 M-element input limit: sort, then
    interpolate between the two neighbouring order statistics"""
    flat = values.reshape(-1)
    if flat.numel() == 0:
        raise ValueError('quantile of an empty tensor')
    srt = torch.sort(flat).values
    pos = float(q) * (flat.numel() - 1)
    lo = int(pos)
    hi = min(lo + 1, flat.numel() - 1)
    return torch.lerp(srt[lo], srt[hi], pos - lo)


def hard_attention_mask(mean_att, att_samp_pct):
    """edges kept while training: head-mean attention strictly above the (1 - att_samp_pct) quantile
    (block_transformer_hard_attention.py:54-55)"""
    return mean_att > linear_quantile(mean_att, 1.0 - att_samp_pct)


def renormalise_attention(attention, index, num_nodes):
    """attention / (sum over the group + 1e-16) (block_transformer_hard_attention.py:33-36)"""
    sums = torch.zeros(num_nodes, dtype=attention.dtype, device=attention.device).index_add_(0, index, attention)
    return attention / (sums[index] + 1e-16)


class HardAttODEblock(ODEblock):
    def __init__(self, odefunc, regularization_fns, opt, device, t=torch.tensor([0, 1]), gamma=0.5):
        super(HardAttODEblock, self).__init__(odefunc, regularization_fns, opt, device, t)
        assert opt['att_samp_pct'] > 0 and opt['att_samp_pct'] <= 1, "attention sampling threshold must be in (0,1]"
        self.train_integrator = _integrators(opt['adjoint'])
        self.test_integrator = _integrators(opt['adjoint'])
        self.set_tol()
        if opt['function'] not in {'GAT', 'transformer'}:
            self.multihead_att_layer = SpGraphTransAttentionLayer(opt['hidden_dim'], opt['hidden_dim'], opt, device,
                                                                  edge_weights=self.odefunc.edge_weight).to(device)

    def get_attention_weights(self, x):
        if self.opt['function'] not in {'GAT', 'transformer'}:
            attention, values = self.multihead_att_layer(x, self.data_edge_index)
        else:
            attention, values = self.odefunc.multihead_att_layer(x, self.data_edge_index)
        return attention

    def renormalise_attention(self, attention):
        ei = self.odefunc.edge_index
        ei = ei[0] if ei.dim() == 3 else ei
        return renormalise_attention(attention.reshape(-1), ei[self.opt['attention_norm_idx']], self.num_nodes)

    def forward(self, x, graph_data, y=None):
        if self.partition is not None:
            raise NotImplementedError("block 'hard_attention' is not available on a row-partitioned graph")
        self.reset_graph_data(graph_data, x.dtype, y)
        attention_weights = self.get_attention_weights(x)
        batched = self.data_edge_index.dim() == 3
        if self.training:
            with torch.no_grad():
                mean_att = attention_weights.mean(dim=-1).reshape(-1)
                if self.opt.get('use_flux', False):
                    ei = self.data_edge_index[0] if batched else self.data_edge_index
                    x2 = x[0] if x.dim() == 3 else x
                    mean_att = mean_att * torch.linalg.norm(x2[ei[0]] - x2[ei[1]], dim=1)
                mask = hard_attention_mask(mean_att, self.opt['att_samp_pct'])
                self.odefunc.edge_index = self.data_edge_index[..., mask].contiguous()
                sampled = self.renormalise_attention(mean_att[mask])
                print('retaining {} of {} edges'.format(self.odefunc.edge_index.shape[-1], self.data_edge_index.shape[-1]))
                self.odefunc.attention_weights = sampled[None] if batched else sampled
        else:
            self.odefunc.edge_index = self.data_edge_index
            self.odefunc.attention_weights = attention_weights.mean(dim=-1)
        return self._integrate(x)
