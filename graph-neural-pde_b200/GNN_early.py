"""`GNNEarly` - mirror of the reference's src/GNN_early.py:16-105: a GNN whose TEST-time integrator is EarlyStopInt
(per-step decode, best validation step remembered on `odeblock.test_integrator.solver`)."""
import torch

from .base_classes import BaseGNN
from .early_stop_solver import EarlyStopInt
from .model_configurations import set_block, set_function


class GNNEarly(BaseGNN):
    def __init__(self, opt, num_classes, num_features, device=torch.device('cpu')):
        super(GNNEarly, self).__init__(opt, num_classes, num_features, device)
        self.f = set_function(opt)
        block = set_block(opt)
        self.device = device
        time_tensor = torch.tensor([0, self.T]).to(device)
        self.odeblock = block(self.f, self.regularization_fns, opt, device, t=time_tensor).to(device)
        # overwrite the test integrator with this custom one (GNN_early.py:23-25)
        self.odeblock.test_integrator = EarlyStopInt(self.T, self.opt, self.device)

    def set_solver_m2(self):
        self.odeblock.test_integrator.m2_weight = self.m2.weight.data.detach().clone().to(self.device)
        self.odeblock.test_integrator.m2_bias = self.m2.bias.data.detach().clone().to(self.device)

    def set_solver_data(self, data):
        self.odeblock.test_integrator.data = data

    def forward(self, x, graph_data, x2=None, pos_encoding=None):
        with torch.no_grad():
            self.set_solver_m2()
        return self._forward(x, graph_data, x2, pos_encoding)
