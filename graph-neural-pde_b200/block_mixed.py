"""`MixedODEblock` - mirror of the reference's src/block_mixed.py:7-62: block-level attention computed once at
t = 0, mixed with the normalised edge weights by a learned gate, parked on `odefunc.attention_weights` (read
by LaplacianODEFunc when opt['block'] == 'mixed', function_laplacian_diffusion.py:50-53), then integrated.

    mixed = mean_h(attention) * (1 - sigmoid(gamma)) + edge_weight * sigmoid(gamma)        (block_mixed.py:29-33)

The mix is [E]-sized torch glue executed once per block forward (not per evaluation); the attention comes from
the CUDA edge-softmax kernels and the integration runs on the CUDA diffusion kernels, like AttODEblock."""
import torch
from torch import nn

from .base_classes import ODEblock, _integrators
from .function_transformer_attention import SpGraphTransAttentionLayer


def mix_attention(attention, edge_weight, gamma):
    """attention [..., E, H], edge_weight [..., E], gamma: the raw gate parameter -> mixed [..., E]"""
    g = torch.sigmoid(gamma)
    return attention.mean(dim=-1) * (1 - g) + edge_weight * g


class MixedODEblock(ODEblock):
    def __init__(self, odefunc, regularization_fns, opt, device, t=torch.tensor([0, 1]), gamma=0.):
        super(MixedODEblock, self).__init__(odefunc, regularization_fns, opt, device, t)
        self.train_integrator = _integrators(opt['adjoint'])
        self.test_integrator = _integrators(opt['adjoint'])
        self.set_tol()
        # parameter trading off between attention and the Laplacian (block_mixed.py:21-22)
        self.gamma = nn.Parameter(gamma * torch.ones(1))
        self.multihead_att_layer = SpGraphTransAttentionLayer(opt['hidden_dim'], opt['hidden_dim'], opt,
                                                              device).to(device)

    def get_attention_weights(self, x):
        attention, values = self.multihead_att_layer(x, self.odefunc.edge_index)
        return attention

    def get_mixed_attention(self, x):
        attention = self.get_attention_weights(x)
        ew = self.odefunc.edge_weight
        if attention.dim() == 3 and ew.dim() == 1:               # fork layout [1,E,H] against an un-batched weight
            ew = ew[None]
        return mix_attention(attention, ew, self.gamma)

    def forward(self, x, graph_data, y=None):
        if self.partition is not None:
            raise NotImplementedError("block 'mixed' is not available on a row-partitioned graph")
        self.reset_graph_data(graph_data, x.dtype, y)
        self.odefunc.attention_weights = self.get_mixed_attention(x)
        return self._integrate(x)
