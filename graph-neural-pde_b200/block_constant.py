"""`ConstantODEblock` - mirror of the reference's src/block_constant.py:5-63 (call pattern of the
hot path: reset graph -> integrate the ODE function over [0, T] -> return the end state)."""
import torch

from .base_classes import ODEblock, _integrators


class ConstantODEblock(ODEblock):
    def __init__(self, odefunc, regularization_fns, opt, device, t=torch.tensor([0, 1])):
        super(ConstantODEblock, self).__init__(odefunc, regularization_fns, opt, device, t)
        self.aug_dim = 2 if opt['augment'] else 1
        self.train_integrator = _integrators(opt['adjoint'])
        self.test_integrator = _integrators(opt['adjoint'])
        self.set_tol()

    def forward(self, x, graph_data, y=None):
        self.reset_graph_data(graph_data, x.dtype, y)
        return self._integrate(x)
