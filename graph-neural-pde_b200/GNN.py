"""`GNN` - mirror of the reference's src/GNN.py:9-74: encoder -> ODE block -> relu -> decoder."""
import torch

from .base_classes import BaseGNN
from .model_configurations import set_block, set_function


class GNN(BaseGNN):
    def __init__(self, opt, num_classes, num_features, device=torch.device('cpu')):
        super(GNN, self).__init__(opt, num_classes, num_features, device)
        self.f = set_function(opt)
        block = set_block(opt)
        time_tensor = torch.tensor([0, self.T]).to(device)
        self.odeblock = block(self.f, self.regularization_fns, opt, device, t=time_tensor).to(device)

    def forward(self, x, graph_data, x2=None, pos_encoding=None):
        """x: [N, F] or the fork's [1, N, F]; graph_data: the edge_index tensor ([2, E] or [1, 2, E])"""
        return self._forward(x, graph_data, x2, pos_encoding)
