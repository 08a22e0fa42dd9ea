"""`EarlyStopInt` - mirror of the reference's src/early_stop_solver.py:236-317 (+ the two solver subclasses it
drives, :30-234): the TEST-time integrator of GNNEarly.  It integrates the ODE function over
[0, earlystopxT * T] with rk4 or dopri5, decodes the state after every step
(relu -> m2 linear -> arg-max, early_stop_solver.py:108-128, 210-226), and remembers the step with the best
validation accuracy (`solver.best_val / best_test / best_train / best_time`).

The reference subclasses torchdiffeq's private solver classes; here the restated solvers of integrators.py take a
per-step callback instead, so the ODE function (the CUDA hot path) is driven exactly as in training.  The decode
(relu -> [N, d] x [d, C] -> masked arg-max -> counts) is ONE kernel launch and one 24-byte read-back
per accepted step (`ops.decode_eval`, csrc/decode.cu) where the reference runs ~20 kernels and six `.item()` syncs.
"""
import torch

from . import integrators, ops


def _masks(data):
    """(name, mask) pairs like PyG's data('train_mask', 'val_mask', 'test_mask')"""
    return [(k, getattr(data, k)) for k in ('train_mask', 'val_mask', 'test_mask')]


class EarlyStopSolverState:
    """what the reference keeps on its EarlyStopDopri5 / EarlyStopRK4 instance (read by GNNEarly after the solve)"""

    def __init__(self, opt, data, m2_weight, m2_bias):
        self.opt, self.data, self.m2_weight, self.m2_bias = opt, data, m2_weight, m2_bias
        self.lf = torch.nn.CrossEntropyLoss()
        self.best_train = 0
        self.best_val = 0
        self.best_test = 0
        self.best_time = 0
        self.n_evaluations = 0

    def set_accs(self, train, val, test, time):
        self.best_train, self.best_val, self.best_test = train, val, test
        self.best_time = float(time)

    @torch.no_grad()
    def test(self, logits):
        accs = []
        for _, mask in _masks(self.data):
            pred = logits[..., mask, :].argmax(dim=-1) if logits.dim() == 3 else logits[mask].argmax(dim=-1)
            y = self.data.y[mask]
            accs.append(pred.reshape(-1).eq(y.reshape(-1)).sum().item() / int(mask.sum().item()))
        return accs

    @torch.no_grad()
    def evaluate(self, z, t0, t1):
        """relu -> m2 -> per-mask accuracy.  An augmented state (wider than m2) contributes its first
        m2_weight.shape[-1] columns (early_stop_solver.py:113-114): the kernel simply does not read the rest."""
        masks = [m for _, m in _masks(self.data)]
        counts, _, _ = ops.decode_eval(z, self.m2_weight, self.m2_bias, self.data.y, masks)
        c = counts.tolist()                            # the one host synchronisation of the step
        self.n_evaluations += 1
        return [c[i] / c[3 + i] for i in range(3)]

    def on_step(self, t0, t1, y1):
        y = y1[0] if isinstance(y1, (tuple, list)) else y1
        train_acc, val_acc, test_acc = self.evaluate(y, t0, t1)
        if val_acc > self.best_val:
            self.set_accs(train_acc, val_acc, test_acc, t1)


class EarlyStopInt(torch.nn.Module):
    def __init__(self, t, opt, device=None):
        super(EarlyStopInt, self).__init__()
        self.device = device
        self.solver = None
        self.data = None
        self.max_test_steps = opt['max_test_steps']
        self.m2_weight = None
        self.m2_bias = None
        self.opt = opt
        self.t = torch.tensor([0, opt['earlystopxT'] * t], dtype=torch.float).to(self.device)

    def __call__(self, func, y0, t, method=None, rtol=1e-7, atol=1e-9, adjoint_method='dopri5', adjoint_atol=1e-9,
                 adjoint_rtol=1e-7, options=None):
        """same signature as the reference; `t` is ignored in favour of self.t = [0, earlystopxT * T], like there"""
        method = self.opt['method']
        assert method in ['rk4', 'dopri5'], 'Only dopri5 and rk4 implemented with early stopping'
        if self.data is None or self.m2_weight is None:
            raise RuntimeError('EarlyStopInt needs .data (masks, y) and .m2_weight / .m2_bias before it is called')
        state = EarlyStopSolverState(self.opt, self.data, self.m2_weight, self.m2_bias)
        self.solver = state
        options = dict(options or {})
        options.pop('max_iters', None)
        return integrators.odeint(func, y0, self.t.detach().cpu().to(y0.dtype), rtol=rtol, atol=atol, method=method,
                                  options=options, _on_step=state.on_step,
                                  _max_attempts=self.max_test_steps if method == 'dopri5' else None)
