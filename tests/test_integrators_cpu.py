"""CPU: the product's restated solvers (grand_b200/integrators.py - host logic, plain torch) against the
oracle's restatement (oracle/integrate.py), closed-form solutions and autograd through the solver.
The ODE functions here are small dense torch modules: the solvers are generic, the CUDA operators are not
involved (they have no CPU path)."""
import math

import pytest
import torch

from oracle import integrate as oi

torch.manual_seed(0)


def product():
    from grand_b200 import integrators
    return integrators


class Linear(torch.nn.Module):
    """dy/dt = y A^T + b (autonomous) with an nfe counter like the reference's ODE functions"""

    def __init__(self, d, seed=0, scale=0.3):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.A = torch.nn.Parameter(scale * torch.randn(d, d, generator=g, dtype=torch.float64))
        self.b = torch.nn.Parameter(0.1 * torch.randn(d, generator=g, dtype=torch.float64))
        self.nfe = 0

    def forward(self, t, y):
        self.nfe += 1
        return y @ self.A.T + self.b


class TimeDependent(torch.nn.Module):
    """dy/dt = -y * w + sin(t): the time argument must reach the function with the right sign and value"""

    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.tensor(0.7, dtype=torch.float64))

    def forward(self, t, y):
        return -y * self.w + torch.sin(t)


def exact_linear(func, y0, t):
    """y(t) for dy/dt = A y + b via the augmented matrix exponential"""
    d = func.A.shape[0]
    m = torch.zeros(d + 1, d + 1, dtype=torch.float64)
    m[:d, :d] = func.A.detach()
    m[:d, d] = func.b.detach()
    e = torch.linalg.matrix_exp(m * t)
    aug = torch.cat([y0, torch.ones(y0.shape[0], 1, dtype=torch.float64)], dim=1)
    return (aug @ e.T)[:, :d]


@pytest.mark.parametrize('method', ['euler', 'midpoint', 'rk4'])
@pytest.mark.parametrize('t1,step', [(2.0, 1.0), (1.0, 0.3), (2.5, 1.0)])
def test_fixed_grid_solvers_match_the_oracle_bitwise(method, t1, step):
    """same grid (incl. the clamped last point), same stage formulas -> identical fp64 results up to the
    order of the stage sums"""
    P = product()
    func = Linear(5)
    y0 = torch.randn(7, 5, dtype=torch.float64)
    t = torch.tensor([0.0, t1], dtype=torch.float64)
    with torch.no_grad():
        got = P.odeint(func, y0, t, method=method, options=dict(step_size=step))
        nfe = func.nfe
        func.nfe = 0
        ref = oi.odeint(func, y0, [0.0, t1], method, step)
    assert got.shape == ref.shape == (2, 7, 5)
    assert torch.equal(got[0], y0)
    torch.testing.assert_close(got, ref, rtol=1e-13, atol=1e-13)
    assert nfe == func.nfe                                     # the reference counts one nfe per call


@pytest.mark.parametrize('method,order', [('euler', 1), ('midpoint', 2), ('rk4', 4)])
def test_fixed_grid_convergence_order(method, order):
    P = product()
    func = Linear(4, seed=3)
    y0 = torch.randn(3, 4, dtype=torch.float64)
    exact = exact_linear(func, y0, 1.0)
    errs = []
    with torch.no_grad():
        for h in (0.1, 0.05):
            y = P.odeint(func, y0, torch.tensor([0.0, 1.0]), method=method, options=dict(step_size=h))[1]
            errs.append(float((y - exact).abs().max()))
    rate = math.log2(errs[0] / errs[1])
    assert abs(rate - order) < 0.35, (errs, rate)


def test_rk4_is_the_three_eighths_rule():
    """torchdiffeq's fixed-grid rk4 is rk4_alt_step_func (3/8 rule), not the classic tableau: one step on
    dy/dt = t^3-free test function y' = y gives 1 + h + h^2/2 + h^3/6 + h^4/24 for both, so use y' = t*y"""
    P = product()
    f = lambda t, y: t * y
    y0 = torch.ones(1, dtype=torch.float64)
    h = 0.5
    got = P.odeint(f, y0, torch.tensor([0.0, h]), method='rk4', options=dict(step_size=h))[1]
    k1 = 0.0
    k2 = (h / 3) * (1 + h * k1 / 3)
    k3 = (2 * h / 3) * (1 + h * (k2 - k1 / 3))
    k4 = h * (1 + h * (k1 - k2 + k3))
    ref = 1 + h * (k1 + 3 * k2 + 3 * k3 + k4) / 8
    assert abs(float(got) - ref) < 1e-15
    classic = 1 + h * (0 + 2 * (h / 2) * 1 + 2 * (h / 2) * (1 + h / 2 * (h / 2)) + h * (1 + h * (h / 2) * (1 + h * h / 4))) / 6
    assert abs(float(got) - classic) > 1e-6


def test_time_argument_forward_and_reverse():
    P = product()
    func = TimeDependent()
    y0 = torch.tensor([0.3, -1.2], dtype=torch.float64)
    with torch.no_grad():
        fwd = P.odeint(func, y0, torch.tensor([0.0, 1.5]), method='rk4', options=dict(step_size=0.05))
        back = P.odeint(func, fwd[1], torch.tensor([1.5, 0.0]), method='rk4', options=dict(step_size=0.05))
        ref = oi.odeint(func, y0, [0.0, 1.5], 'rk4', 0.05)
    torch.testing.assert_close(fwd, ref, rtol=1e-13, atol=1e-13)
    torch.testing.assert_close(back[1], y0, rtol=0, atol=1e-6)   # reverse-time solve returns to the start


def test_intermediate_times_are_interpolated_on_the_fixed_grid():
    P = product()
    func = Linear(3, seed=5)
    y0 = torch.randn(2, 3, dtype=torch.float64)
    with torch.no_grad():
        sol = P.odeint(func, y0, torch.tensor([0.0, 0.25, 1.0]), method='euler', options=dict(step_size=0.5))
        one = P.odeint(func, y0, torch.tensor([0.0, 0.5]), method='euler', options=dict(step_size=0.5))[1]
    torch.testing.assert_close(sol[1], y0 + 0.5 * (one - y0), rtol=1e-14, atol=1e-14)
    assert sol.shape == (3, 2, 3)


# ---- adaptive solvers -----------------------------------------------------------------------------
@pytest.mark.parametrize('method', ['dopri5', 'adaptive_heun'])
def test_adaptive_solvers_match_the_oracle_and_the_exact_solution(method):
    P = product()
    func = Linear(4, seed=11)
    y0 = torch.randn(5, 4, dtype=torch.float64)
    rtol, atol = (1e-8, 1e-10) if method == 'dopri5' else (1e-5, 1e-7)
    with torch.no_grad():
        got = P.odeint(func, y0, torch.tensor([0.0, 2.0]), method=method, rtol=rtol, atol=atol)
        nfe = func.nfe
        func.nfe = 0
        ref = oi.odeint(func, y0, [0.0, 2.0], method, None, rtol, atol)
    assert nfe == func.nfe                                     # same accepted / rejected step sequence
    torch.testing.assert_close(got, ref, rtol=1e-12, atol=1e-12)
    exact = exact_linear(func, y0, 2.0)
    tol = 1e-6 if method == 'dopri5' else 1e-3
    assert float((got[1] - exact).abs().max()) < tol


def test_dopri5_lands_on_requested_times_by_dense_output():
    P = product()
    func = Linear(3, seed=2)
    y0 = torch.randn(2, 3, dtype=torch.float64)
    ts = [0.0, 0.37, 1.0, 1.9]
    with torch.no_grad():
        sol = P.odeint(func, y0, torch.tensor(ts), method='dopri5', rtol=1e-9, atol=1e-11)
    for i, tv in enumerate(ts):
        assert float((sol[i] - exact_linear(func, y0, tv)).abs().max()) < 1e-6


def test_unknown_method_and_bad_time_vectors_raise():
    P = product()
    f = lambda t, y: y
    y0 = torch.ones(2)
    with pytest.raises(NotImplementedError):
        P.odeint(f, y0, torch.tensor([0.0, 1.0]), method='implicit_adams')
    with pytest.raises(ValueError):
        P.odeint(f, y0, torch.tensor([0.0]), method='rk4')
    with pytest.raises(ValueError):
        P.odeint(f, y0, torch.tensor([0.0, 1.0, 0.5]), method='rk4')
    with pytest.raises(NotImplementedError):
        P.odeint_adjoint(f, (y0, y0), torch.tensor([0.0, 1.0]), method='rk4')


# ---- adjoint --------------------------------------------------------------------------------------
def _loss_and_grads(solver, func, y0, **kw):
    y0 = y0.clone().requires_grad_(True)
    for p in func.parameters():
        p.grad = None
    z = solver(func, y0)
    w = torch.linspace(-1, 1, z[-1].numel(), dtype=z.dtype).view_as(z[-1])
    loss = (z[-1] * w).sum()
    loss.backward()
    return float(loss.detach()), y0.grad.clone(), [p.grad.clone() for p in func.parameters()]


@pytest.mark.parametrize('method,step', [('rk4', 0.25), ('euler', 0.1), ('midpoint', 0.2)])
def test_adjoint_matches_the_oracle_adjoint(method, step):
    """same augmented reverse solve: identical numbers (sign conventions ride on stage coefficients here, on
    negated tensors in the oracle - algebraically the same)"""
    P = product()
    func = Linear(4, seed=7)
    y0 = torch.randn(6, 4, dtype=torch.float64)
    t = torch.tensor([0.0, 1.0], dtype=torch.float64)
    ours = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, method=method, options=dict(step_size=step),
                                                         adjoint_method=method,
                                                         adjoint_options=dict(step_size=step)), func, y0)
    ref = _loss_and_grads(lambda f, y: oi.odeint_adjoint(f, y, [0.0, 1.0], method, step), func, y0)
    assert abs(ours[0] - ref[0]) < 1e-13
    torch.testing.assert_close(ours[1], ref[1], rtol=1e-12, atol=1e-13)
    for a, b in zip(ours[2], ref[2]):
        torch.testing.assert_close(a, b, rtol=1e-12, atol=1e-13)


def test_adjoint_gradients_converge_to_backprop_through_the_solver():
    """continuous adjoint vs discretise-then-differentiate: equal up to the O(h^4) truncation of rk4"""
    P = product()
    func = Linear(3, seed=9)
    y0 = torch.randn(4, 3, dtype=torch.float64)
    t = torch.tensor([0.0, 1.0], dtype=torch.float64)
    h = 0.05
    adj = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, method='rk4', options=dict(step_size=h)), func, y0)
    bp = _loss_and_grads(lambda f, y: P.odeint(f, y, t, method='rk4', options=dict(step_size=h)), func, y0)
    assert abs(adj[0] - bp[0]) < 1e-14
    torch.testing.assert_close(adj[1], bp[1], rtol=1e-6, atol=1e-8)
    for a, b in zip(adj[2], bp[2]):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-8)


def test_adjoint_with_time_dependent_function_and_dopri5():
    P = product()
    func = TimeDependent()
    y0 = torch.tensor([0.4, -0.9, 1.3], dtype=torch.float64)
    t = torch.tensor([0.0, 1.2], dtype=torch.float64)
    kw = dict(method='dopri5', rtol=1e-9, atol=1e-11)
    adj = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), func, y0)
    bp = _loss_and_grads(lambda f, y: P.odeint(f, y, t, **kw), func, y0)
    torch.testing.assert_close(adj[1], bp[1], rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(adj[2][0], bp[2][0], rtol=1e-6, atol=1e-8)


class FusedLinear(Linear):
    """exposes the fused f + VJP hook of the CUDA functions (ODEFuncTransformerAtt.eval_vjp) with closed-form
    products, and counts how often each route was taken"""

    def __init__(self, *a, decline=False, **kw):
        super().__init__(*a, **kw)
        self.fused_calls, self.decline = 0, decline

    def eval_vjp(self, t, y, cot, params):
        if self.decline:
            return None
        self.fused_calls += 1
        self.nfe += 1
        f = y @ self.A.T + self.b
        grads = {id(self.A): cot.T @ y, id(self.b): cot.sum(0)}
        return f, cot @ self.A, [grads.get(id(p)) for p in params]


@pytest.mark.parametrize('method', ['rk4', 'dopri5'])
def test_eval_vjp_hook_gives_the_autograd_numbers(method):
    P = product()
    y0 = torch.randn(5, 4, dtype=torch.float64)
    t = torch.tensor([0.0, 1.0], dtype=torch.float64)
    kw = dict(method=method, options=dict(step_size=0.25), rtol=1e-8, atol=1e-10)
    plain, fused, declined = Linear(4, seed=1), FusedLinear(4, seed=1), FusedLinear(4, seed=1, decline=True)
    ref = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), plain, y0)
    got = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), fused, y0)
    dec = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), declined, y0)
    assert fused.fused_calls > 0 and declined.fused_calls == 0
    assert fused.nfe == plain.nfe == declined.nfe              # one nfe per augmented-dynamics call either way
    for a, b in ((got, ref), (dec, ref)):
        torch.testing.assert_close(a[1], b[1], rtol=1e-12, atol=1e-13)
        for u, v in zip(a[2], b[2]):
            torch.testing.assert_close(u, v, rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize('method,step', [('rk4', 0.25), ('euler', 0.1), ('midpoint', 0.2)])
@pytest.mark.parametrize('fused', [False, True])
def test_packed_parameter_integrals_are_bit_identical_and_cheaper(method, step, fused, monkeypatch):
    """fixed-grid adjoint: all parameter integrals in ONE flat buffer that is only combined at the end of a step (and
    no time integral) - the same bits as one tensor per parameter with every stage value formed, a fraction of the
    small-tensor operations"""
    P = product()
    from grand_b200 import integrators as I
    y0 = torch.randn(30, 4, dtype=torch.float64)                 # >= 100 elements: only parameter-sized tensors are counted
    t = torch.tensor([0.0, 0.5, 1.0], dtype=torch.float64)
    solve = lambda f, y: P.odeint_adjoint(f, y, t, method=method, options=dict(step_size=step), adjoint_method=method,
                                          adjoint_options=dict(step_size=step))
    counts = {}
    real = I._combine_one

    def counting(yi, terms):
        counts[key] = counts.get(key, 0) + (1 if yi.numel() < 100 and any(c != 0 for _, c in terms) else 0)
        return real(yi, terms)

    monkeypatch.setattr(I, '_combine_one', counting)
    out = {}
    for key in (True, False):
        monkeypatch.setattr(I, '_PACK_PARAMS', key)
        func = (FusedLinear if fused else Linear)(4, seed=3)
        out[key] = _loss_and_grads(solve, func, y0)
    assert out[True][0] == out[False][0] and torch.equal(out[True][1], out[False][1])
    for a, b in zip(out[True][2], out[False][2]):
        assert torch.equal(a, b) and a.shape == b.shape
    n_stage = {'euler': 1, 'midpoint': 2, 'rk4': 4}[method]
    steps = 2 * math.ceil(0.5 / step - 1e-9)                     # two output intervals, each on its own grid
    assert counts[True] == steps                                 # one end-of-step combination of the flat buffer
    assert counts[False] == steps * n_stage * 3                  # (time, A, b) x every stage and end-of-step combination


def test_adjoint_params_subset_and_frozen_parameters():
    P = product()
    func = Linear(3, seed=4)
    func.b.requires_grad_(False)
    y0 = torch.randn(2, 3, dtype=torch.float64, requires_grad=True)
    z = P.odeint_adjoint(func, y0, torch.tensor([0.0, 1.0]), method='rk4', options=dict(step_size=0.5))
    z[-1].sum().backward()
    assert func.A.grad is not None and func.b.grad is None and y0.grad is not None


# ---- two-phase adjoint protocol (row-partitioned graphs: the next stage's exchange overlaps the column pass) -------
class PhasedLinear(Linear):
    """mock of ODEFuncTransformerAtt's two-phase protocol with closed-form products; records the order of issue"""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.log = []

    def phased_ok(self, y):
        return True

    def prefetch_state(self, y):
        self.log.append('prefetch')
        return ('pending', y)

    def eval_vjp_phased(self, t, y, cot, params, prefetched=None):
        assert prefetched is None or prefetched[1] is y           # the solver hands back exactly what it prefetched
        self.log.append('rows' if prefetched is None else 'rows(prefetched)')
        self.nfe += 1
        f = y @ self.A.T + self.b

        def finish():
            self.log.append('cols')
            grads = {id(self.A): cot.T @ y, id(self.b): cot.sum(0)}
            return cot @ self.A, [grads.get(id(p)) for p in params]
        return f, finish


@pytest.mark.parametrize('method,step,t1', [('rk4', 0.5, 2.0), ('rk4', 1.0, 2.5), ('midpoint', 0.25, 1.0), ('euler', 0.2, 1.0)])
def test_phased_adjoint_solve_is_the_plain_adjoint_solve(method, step, t1):
    P = product()
    y0 = torch.randn(5, 4, dtype=torch.float64)
    t = torch.tensor([0.0, t1], dtype=torch.float64)
    kw = dict(method=method, options=dict(step_size=step))
    plain, phased = Linear(4, seed=1), PhasedLinear(4, seed=1)
    ref = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), plain, y0)
    got = _loss_and_grads(lambda f, y: P.odeint_adjoint(f, y, t, **kw), phased, y0)
    assert phased.nfe == plain.nfe
    assert torch.equal(got[1], ref[1])                             # same arithmetic in the same order: bit-identical
    for a, b in zip(got[2], ref[2]):
        assert torch.equal(a, b)
    log = phased.log
    stages = {'rk4': 4, 'midpoint': 2, 'euler': 1}[method]
    n_evals = log.count('rows') + log.count('rows(prefetched)')
    assert n_evals == log.count('cols')                            # every first half is finished
    assert log.count('rows') == 1                                  # only the very first stage has nothing prefetched
    assert log.count('prefetch') == n_evals - 1                    # ... and nothing is prefetched past the last stage
    assert n_evals % stages == 0
    for i, ev in enumerate(log[:-1]):                              # order of issue: rows, prefetch (next state), cols
        if ev.startswith('rows') and i + 2 < len(log):
            assert log[i + 1] == 'prefetch' and log[i + 2] == 'cols'


def test_phased_adjoint_with_intermediate_output_times():
    """several requested times: the y component is reset to the stored forward value at every boundary, so the gather
    prefetched across a boundary is never used"""
    P = product()
    y0 = torch.randn(3, 4, dtype=torch.float64)
    t = torch.tensor([0.0, 1.0, 2.0], dtype=torch.float64)
    kw = dict(method='rk4', options=dict(step_size=0.5))

    def run(func):
        y = y0.clone().requires_grad_(True)
        z = P.odeint_adjoint(func, y, t, **kw)
        (z[1].sum() + 2 * z[2].sum()).backward()
        return y.grad, func.A.grad
    ref, got = run(Linear(4, seed=2)), run(PhasedLinear(4, seed=2))
    assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1])


def test_dormand_prince_tableau_is_pinned_to_scipy():
    """torchdiffeq is absent, so the adaptive solver's numbers cannot be pinned to it; its Butcher tableau can be pinned
    to an independent implementation of the same published method: scipy's RK45.  Nodes, stage matrix and 5th-order
    weights are identical; the error weights are Shampine's variant (the one torchdiffeq ships), which is -2/3 of the
    original Dormand-Prince embedded difference that scipy uses - the same error direction, a different scale."""
    import numpy as np
    from scipy.integrate._ivp.rk import RK45
    for tab in (product()._DP, dict(alpha=oi._DP_ALPHA, beta=oi._DP_BETA, c_sol=oi._DP_C_SOL, c_err=oi._DP_C_ERR)):
        assert np.allclose(RK45.C[1:], tab['alpha'][:5], rtol=0, atol=1e-15) and tab['alpha'][5] == 1.0
        for i, row in enumerate(tab['beta'][:5]):
            assert np.allclose(RK45.A[i + 1, :len(row)], row, rtol=0, atol=1e-15)
        assert np.allclose(RK45.B, tab['beta'][5], rtol=0, atol=1e-15)          # FSAL: last stage row = solution weights
        assert np.allclose(RK45.B, tab['c_sol'][:6], rtol=0, atol=1e-15) and tab['c_sol'][6] == 0
        assert np.allclose(np.asarray(tab['c_err']), -2.0 / 3.0 * RK45.E, rtol=0, atol=1e-15)


def test_dopri5_agrees_with_scipy_rk45_on_a_linear_system():
    """different step controllers, same method and tolerances: the two solutions agree to the requested accuracy"""
    import numpy as np
    from scipy.integrate import solve_ivp
    P = product()
    func = Linear(4, seed=6)
    y0 = torch.randn(1, 4, dtype=torch.float64)
    A, b = func.A.detach().numpy(), func.b.detach().numpy()
    ref = solve_ivp(lambda t, y: A @ y + b, (0.0, 2.0), y0[0].numpy(), method='RK45', rtol=1e-9, atol=1e-11).y[:, -1]
    with torch.no_grad():
        got = P.odeint(func, y0, torch.tensor([0.0, 2.0]), method='dopri5', rtol=1e-9, atol=1e-11)[1, 0].numpy()
    assert np.allclose(got, ref, rtol=1e-7, atol=1e-9)


def test_one_pass_stage_combination_is_differentiable(emulated_abi, monkeypatch):
    """above integrators._FUSE_MIN the stage combination is one kernel: under autograd it must stay on the graph
    (plain odeint differentiated through the solver: adjoint=False training, CGNN)"""
    from grand_b200 import integrators
    calls = []

    def raw(yi, ks, cs):
        calls.append(len(ks))
        out = yi.clone()
        for k, c in zip(ks, cs):
            out = out + c * k
        return out.detach()

    monkeypatch.setattr(integrators, '_stage_combine_raw', raw)
    monkeypatch.setattr(integrators, '_FUSE_MIN', 1)
    real_is_cuda = torch.Tensor.is_cuda
    monkeypatch.setattr(torch.Tensor, 'is_cuda', property(lambda self: True))
    try:
        a = torch.tensor(0.7, requires_grad=True)
        y0 = torch.randn(6, 3, requires_grad=True)
        z = integrators.odeint(lambda t, y: -a * y, y0, torch.tensor([0., 1.]), method='rk4', options=dict(step_size=0.5))[1]
        gy, ga = torch.autograd.grad(z.sum(), (y0, a))
    finally:
        monkeypatch.setattr(torch.Tensor, 'is_cuda', real_is_cuda)
    assert calls                                                   # the one-pass form was taken
    y0d, ad = y0.detach().double().requires_grad_(True), torch.tensor(0.7, dtype=torch.float64, requires_grad=True)
    acc = y0d
    for _ in range(2):                                             # two rk4 steps of y' = -a y by hand
        h = 0.5
        k1 = -ad * acc; k2 = -ad * (acc + h / 2 * k1); k3 = -ad * (acc + h / 2 * k2); k4 = -ad * (acc + h * k3)
        acc = acc + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
    ry, ra = torch.autograd.grad(acc.sum(), (y0d, ad))
    assert torch.allclose(gy.double(), ry, rtol=1e-5) and torch.allclose(ga.double(), ra, rtol=1e-5)


@pytest.mark.parametrize('method', ['dopri5', 'adaptive_heun'])
def test_adaptive_adjoint_with_packed_parameter_integrals_is_the_unpacked_solve(monkeypatch, method):
    """the adaptive adjoint carries all parameter integrals in one flat component (4 components instead of 3 + one per
    parameter tensor: ~8x fewer launch-sized kernels per attempted dopri5 step); the error norm still takes the RMS
    of every parameter tensor on its own (torchdiffeq's mixed norm), so step sequence, NFE and gradients are unchanged
    (up to the rounding of a segmented mean in the step-size controller)"""
    from grand_b200 import integrators

    class Func(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.A = torch.nn.Parameter(0.5 * torch.randn(5, 5, dtype=torch.float64))
            self.b = torch.nn.Parameter(0.1 * torch.randn(5, dtype=torch.float64))
            self.c = torch.nn.Parameter(torch.tensor(0.3, dtype=torch.float64))
            self.unused = torch.nn.Parameter(torch.ones(3, dtype=torch.float64))
            self.nfe = 0

        def forward(self, t, y):
            self.nfe += 1
            return torch.tanh(y @ self.A.T + self.b) * self.c - 0.1 * y

    out = {}
    for pack in (True, False):
        monkeypatch.setattr(integrators, '_PACK_PARAMS', pack)
        torch.manual_seed(0)
        f = Func()
        y0 = torch.randn(7, 5, dtype=torch.float64, requires_grad=True)
        z = integrators.odeint_adjoint(f, y0, torch.tensor([0., 2.]), method=method, rtol=1e-6, atol=1e-8,
                                       adjoint_params=tuple(f.parameters()))[1]
        f.nfe = 0
        grads = torch.autograd.grad(z.square().sum(), [y0] + list(f.parameters()), allow_unused=True)
        out[pack] = (f.nfe, grads)
    assert out[True][0] == out[False][0] > 0
    for a, b in zip(out[True][1], out[False][1]):            # the per-tensor RMS comes from a segmented sum: the step
        assert torch.allclose(a, b, rtol=1e-9, atol=1e-13)   # controller sees the same ratio up to its last bits
    assert float(out[True][1][4].abs().max()) == 0.0          # the parameter the dynamics never read
