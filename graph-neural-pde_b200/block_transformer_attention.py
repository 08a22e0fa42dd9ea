"""`AttODEblock` - mirror of the reference's src/block_transformer_attention.py:6-69: owns a second
SpGraphTransAttentionLayer, computes attention ONCE at t = 0 from the block input and parks it on
`odefunc.attention_weights` (read by LaplacianODEFunc when opt['block'] == 'attention'), then
integrates.  With opt['function'] == 'transformer' that block-level attention is never read on the path
(the reference computes and discards it); here it is evaluated lazily, on the first read of
`odefunc.attention_weights` (ODEFunc.defer_attention_weights)."""
import torch

from .base_classes import ODEblock, _integrators
from .function_transformer_attention import SpGraphTransAttentionLayer


class AttODEblock(ODEblock):
    def __init__(self, odefunc, regularization_fns, opt, device, t=torch.tensor([0, 1]), gamma=0.5):
        super(AttODEblock, self).__init__(odefunc, regularization_fns, opt, device, t)
        self.train_integrator = _integrators(opt['adjoint'])
        self.test_integrator = _integrators(opt['adjoint'])
        self.set_tol()
        self.multihead_att_layer = SpGraphTransAttentionLayer(opt['hidden_dim'], opt['hidden_dim'], opt, device,
                                                              edge_weights=self.odefunc.edge_weight).to(device)

    def get_attention_weights(self, x):
        attention, values = self.multihead_att_layer(x, self.odefunc.edge_index)
        return attention

    def forward(self, x, graph_data, y=None):
        self.reset_graph_data(graph_data, x.dtype, y)
        if self.partition is not None:
            # row-partitioned graph: the block-level attention is only read by LaplacianODEFunc; with
            # opt['function'] == 'transformer' the reference computes and discards it
            if self.opt['function'] != 'transformer':
                raise NotImplementedError("block 'attention' with a row-partitioned graph needs function 'transformer'")
            self.odefunc.attention_weights = None
            return self._integrate(x)
        if self.opt['function'] == 'transformer' and hasattr(self.odefunc, 'defer_attention_weights'):
            # ODEFuncTransformerAtt recomputes its own attention at every evaluation and never reads this one: it is
            # evaluated on the first read of `odefunc.attention_weights` instead (same values: same layer, same
            # block input, the graph of this forward), not once per block forward
            xin, edge_index = x.detach(), self.odefunc.edge_index
            self.odefunc.defer_attention_weights(lambda: self.multihead_att_layer(xin, edge_index)[0])
        else:
            self.odefunc.attention_weights = self.get_attention_weights(x)
        return self._integrate(x)
