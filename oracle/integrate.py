"""Oracle: the integrators the reference calls (TEST INFRASTRUCTURE ONLY).  **PARITY UNPINNED.**

The reference delegates integration to torchdiffeq (`src/base_classes.py:45-48`,
`src/block_constant.py:34-51`, `src/block_transformer_attention.py:39-64`; README pin
`torchdiffeq==0.2.1`), which is neither vendored under /root/reference nor installable in this
image.  This file restates the published torchdiffeq 0.2.x algorithm from its documentation:

* fixed grid: t_k = t0 + k*step_size, niters = ceil((t1-t0)/step_size + 1), last point clamped to
  t1, linear interpolation onto requested times; `euler`, `midpoint`, `rk4` (the 3/8-rule variant,
  `rk4_alt_step_func`).
* adaptive: `dopri5` (Dormand-Prince 5(4), FSAL, Hairer initial step, rms error norm, safety 0.9,
  ifactor 10, dfactor 0.2, 4th-order dense output used to land exactly on t1) and
  `adaptive_heun` (order 2).
* `odeint_adjoint`: forward solve under no_grad; backward = reverse-time solve of the augmented
  state (vjp_t, y, a_y, a_params) whose dynamics evaluate func under enable_grad and take
  autograd.grad(f, (t, y, *params), -a_y, allow_unused=True).  A decreasing time vector is
  handled as torchdiffeq does: negate t and integrate  -func(-t, y)  forward.

No reference test pins numeric results at this boundary (only output shapes,
`test/test_early_stop.py:48-115`), hence "parity unpinned".  What CAN be pinned without torchdiffeq is
pinned in tests/test_integrators_cpu.py: the Dormand-Prince nodes / stage matrix / 5th-order weights equal
scipy's RK45 tableau (the error weights are Shampine's variant, -2/3 of scipy's embedded difference), dopri5
agrees with scipy's RK45 at equal tolerances, the fixed-grid methods show their orders 1 / 2 / 4 against a
closed-form solution, and the adjoint gradients converge to backprop through the solver.
Written for clarity, not speed.
"""
import math
import torch

# Dormand-Prince 5(4) (Shampine's coefficients, as torchdiffeq's _DORMAND_PRINCE_SHAMPINE_TABLEAU)
_DP_ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1., 1.]
_DP_BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
_DP_C_SOL = [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0]
_DP_C_ERR = [35 / 384 - 1951 / 21600, 0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
             -2187 / 6784 - -12231 / 42400, 11 / 84 - 649 / 6300, -1. / 60.]
_DP_C_MID = [6025192743 / 30085553152 / 2, 0, 51252292925 / 65400821598 / 2,
             -2691868925 / 45128329728 / 2, 187940372067 / 1594534317056 / 2,
             -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2]

_AH_ALPHA = [1.]
_AH_BETA = [[1.]]
_AH_C_SOL = [0.5, 0.5]
_AH_C_ERR = [0.5, -0.5]


def _rms(x):
    return x.pow(2).mean().sqrt()


def _norm(state):
    """torchdiffeq `_mixed_norm`: max over the tuple components of the rms norm."""
    return max(_rms(s) for s in state if s.numel() > 0)


def _axpy(state, dt, ks, coeffs):
    out = []
    for i, y in enumerate(state):
        acc = y
        for k, c in zip(ks, coeffs):
            if c != 0:
                acc = acc + (dt * c) * k[i]
        out.append(acc)
    return tuple(out)


def _grid(t0, t1, step_size):
    n = int(math.ceil((t1 - t0) / step_size + 1))
    g = [t0 + k * step_size for k in range(n)]
    g[-1] = t1
    return g


def _fixed_step(func, method, t0, dt, t1, y):
    f = lambda t, s: tuple(func(t, s))
    if method == 'euler':
        k1 = f(t0, y)
        return _axpy(y, dt, [k1], [1.0])
    if method == 'midpoint':
        k1 = f(t0, y)
        k2 = f(t0 + 0.5 * dt, _axpy(y, dt, [k1], [0.5]))
        return _axpy(y, dt, [k2], [1.0])
    if method == 'rk4':
        k1 = f(t0, y)
        k2 = f(t0 + dt / 3, _axpy(y, dt, [k1], [1 / 3]))
        k3 = f(t0 + dt * 2 / 3, _axpy(y, dt, [k1, k2], [-1 / 3, 1.0]))
        k4 = f(t1, _axpy(y, dt, [k1, k2, k3], [1.0, -1.0, 1.0]))
        return _axpy(y, dt, [k1, k2, k3, k4], [0.125, 0.375, 0.375, 0.125])
    raise NotImplementedError(method)


def _initial_step(func, t0, y0, order, rtol, atol, f0):
    scale = tuple(atol + y.abs() * rtol for y in y0)
    d0 = _norm(tuple(y / s for y, s in zip(y0, scale)))
    d1 = _norm(tuple(f / s for f, s in zip(f0, scale)))
    h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else float(0.01 * d0 / d1)
    y1 = tuple(y + h0 * f for y, f in zip(y0, f0))
    f1 = tuple(func(t0 + h0, y1))
    d2 = float(_norm(tuple((a - b) / s for a, b, s in zip(f1, f0, scale))) / h0)
    d1 = float(d1)
    if d1 <= 1e-15 and d2 <= 1e-15:
        h1 = max(1e-6, h0 * 1e-3)
    else:
        h1 = (0.01 / max(d1, d2)) ** (1.0 / float(order + 1))
    return min(100 * h0, h1)


def _adaptive(func, method, y0, t0, t1, rtol, atol, max_steps=100000):
    """Integrate from t0 to t1 > t0; returns y(t1)."""
    f = lambda t, s: tuple(func(t, s))
    if method == 'dopri5':
        alpha, beta, c_sol, c_err, order = _DP_ALPHA, _DP_BETA, _DP_C_SOL, _DP_C_ERR, 5
    elif method == 'adaptive_heun':
        alpha, beta, c_sol, c_err, order = _AH_ALPHA, _AH_BETA, _AH_C_SOL, _AH_C_ERR, 2
    else:
        raise NotImplementedError(method)
    f0 = f(t0, y0)
    dt = _initial_step(func, t0, y0, order - 1, rtol, atol, f0)
    t, y = t0, y0
    for _ in range(max_steps):
        if t >= t1:
            break
        # one trial step
        ks = [f0]
        for a_i, b_i in zip(alpha, beta):
            ks.append(f(t + a_i * dt, _axpy(y, dt, ks, b_i)))
        y1 = _axpy(y, dt, ks, c_sol)
        err = tuple(sum((dt * c) * k[i] for k, c in zip(ks, c_err) if c != 0) for i in range(len(y)))
        tol = tuple(atol + rtol * torch.max(a.abs(), b.abs()) for a, b in zip(y, y1))
        ratio = float(_norm(tuple(e / s for e, s in zip(err, tol))))
        accept = ratio <= 1
        if accept:
            t_new = t + dt
            if t_new > t1:
                # dense output: land exactly on t1 (dopri5: 4th order fit; heun: linear)
                if method == 'dopri5':
                    y_mid = _axpy(y, dt, ks, _DP_C_MID)
                    y = _interp4(y, y1, y_mid, f0, ks[-1], dt, (t1 - t) / dt)
                else:
                    x = (t1 - t) / dt
                    y = tuple(a + x * (b - a) for a, b in zip(y, y1))
                return y
            # torchdiffeq (0.1.x and 0.2.x rk_common._runge_kutta_step): f1 = k[..., -1] for every tableau
            t, y, f0 = t_new, y1, ks[-1]
        # step-size controller
        if ratio == 0:
            factor = 10.0
        else:
            dfactor = 1.0 if ratio < 1 else 0.2
            factor = min(10.0, max(0.9 / ratio ** (1.0 / order), dfactor))
        dt = dt * factor
    return y


def _interp4(y0, y1, y_mid, f0, f1, dt, x):
    """torchdiffeq `_interp_fit` + `_interp_evaluate`: quartic through (y0,f0),(y_mid),(y1,f1)."""
    out = []
    for a0, a1, am, g0, g1 in zip(y0, y1, y_mid, f0, f1):
        a = 2 * dt * (g1 - g0) - 8 * (a1 + a0) + 16 * am
        b = dt * (5 * g0 - 3 * g1) + 18 * a0 + 14 * a1 - 32 * am
        c = dt * (g1 - 4 * g0) - 11 * a0 - 5 * a1 + 16 * am
        d = dt * g0
        out.append(a0 + x * (d + x * (c + x * (b + x * a))))
    return tuple(out)


def odeint(func, y0, t, method='rk4', step_size=1.0, rtol=1e-7, atol=1e-9):
    """y0: tensor or tuple of tensors; t: list/1-d tensor of >= 2 times (monotone).  Returns the
    solution stacked over t like torchdiffeq ([len(t), *y.shape], or a tuple of those)."""
    is_tuple = isinstance(y0, (tuple, list))
    state = tuple(y0) if is_tuple else (y0,)
    f_user = (lambda tt, s: func(tt, s)) if is_tuple else (lambda tt, s: (func(tt, s[0]),))
    ts = [float(v) for v in t]
    if ts[-1] < ts[0]:
        ts = [-v for v in ts]
        f_dir = lambda tt, s: tuple(-k for k in f_user(_tt(-tt, state), s))
    else:
        f_dir = lambda tt, s: tuple(f_user(_tt(tt, state), s))
    sols = [state]
    for a, b in zip(ts[:-1], ts[1:]):
        if method in ('euler', 'midpoint', 'rk4'):
            g = _grid(a, b, step_size)
            for g0, g1 in zip(g[:-1], g[1:]):
                state = _fixed_step(f_dir, method, g0, g1 - g0, g1, state)
        else:
            state = _adaptive(f_dir, method, state, a, b, rtol, atol)
        sols.append(state)
    out = tuple(torch.stack([s[i] for s in sols]) for i in range(len(state)))
    return out if is_tuple else out[0]


def _tt(tval, state):
    return torch.tensor(tval, dtype=state[0].dtype, device=state[0].device)


class _Adjoint(torch.autograd.Function):
    @staticmethod
    def forward(ctx, func, t, method, step_size, rtol, atol, adj_method, adj_step, adj_rtol, adj_atol, y0, *params):
        with torch.no_grad():
            ys = odeint(func, y0, t, method, step_size, rtol, atol)
        ctx.func, ctx.t, ctx.params = func, [float(v) for v in t], params
        ctx.cfg = (adj_method, adj_step, adj_rtol, adj_atol)
        ctx.save_for_backward(ys)
        return ys

    @staticmethod
    def backward(ctx, grad_ys):
        func, t, params = ctx.func, ctx.t, ctx.params
        adj_method, adj_step, adj_rtol, adj_atol = ctx.cfg
        ys, = ctx.saved_tensors

        def aug_dynamics(tt, s):
            y, a_y = s[1], s[2]
            with torch.enable_grad():
                tt_ = tt.detach().requires_grad_(True)
                y_ = y.detach().requires_grad_(True)
                f = func(tt_, y_)
                grads = torch.autograd.grad(f, (tt_, y_) + tuple(params), -a_y, allow_unused=True)
            vt = grads[0] if grads[0] is not None else torch.zeros_like(tt)
            vy = grads[1] if grads[1] is not None else torch.zeros_like(y)
            vp = [g if g is not None else torch.zeros_like(p) for g, p in zip(grads[2:], params)]
            return (vt, f.detach(), vy, *vp)

        with torch.no_grad():
            state = [torch.zeros((), dtype=ys.dtype, device=ys.device), ys[-1], grad_ys[-1]] + \
                    [torch.zeros_like(p) for p in params]
            for i in range(len(t) - 1, 0, -1):
                sol = odeint(aug_dynamics, tuple(state), [t[i], t[i - 1]], adj_method, adj_step, adj_rtol, adj_atol)
                state = [s[1] for s in sol]
                state[1] = ys[i - 1]
                state[2] = state[2] + grad_ys[i - 1]
        return (None,) * 10 + (state[2],) + tuple(state[3:])


def odeint_adjoint(func, y0, t, method='rk4', step_size=1.0, rtol=1e-7, atol=1e-9,
                   adjoint_method=None, adjoint_step_size=None, adjoint_rtol=None, adjoint_atol=None,
                   adjoint_params=None):
    params = tuple(adjoint_params) if adjoint_params is not None else tuple(func.parameters())
    params = tuple(p for p in params if p.requires_grad)
    return _Adjoint.apply(func, t, method, step_size, rtol, atol,
                          adjoint_method or method, adjoint_step_size or step_size,
                          adjoint_rtol if adjoint_rtol is not None else rtol,
                          adjoint_atol if adjoint_atol is not None else atol, y0, *params)
