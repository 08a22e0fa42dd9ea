"""torchdiffeq-compatible `odeint` / `odeint_adjoint` (the subset the reference calls).

The reference integrates with torchdiffeq (src/base_classes.py:45-48, src/block_constant.py:34-51,
README pin 0.2.1), which is not installable in this image; ODEblock uses the real package when it
imports and these solvers otherwise.  Same call signature:

    odeint(func, y0, t, rtol=, atol=, method=, options=dict(step_size=...))
    odeint_adjoint(func, y0, t, ..., adjoint_method=, adjoint_options=, adjoint_rtol=, adjoint_atol=,
                   adjoint_params=)

Methods: euler, midpoint, rk4 (3/8 rule = torchdiffeq's rk4_alt_step_func), dopri5, adaptive_heun.
Fixed grid: t0 + k*step_size, ceil((t1-t0)/step_size + 1) points, last clamped to t1.  Adaptive:
Hairer initial step, rms norm (max over tuple components), safety 0.9 / ifactor 10 / dfactor 0.2,
dense output to land on requested times.  Adjoint: reverse-time solve of (vjp_t, y, a_y, a_params)
with f evaluated under enable_grad and autograd.grad(f, (t, y, *params), -a_y, allow_unused=True);
decreasing time is handled by negating t and f, as torchdiffeq does.  NFE is whatever `func`
counts: one per call.

Stage combinations y + dt*sum(b_j k_j) are ONE pass over the state (csrc/stage.cu, gb_stage_combine) for
fp32 CUDA tensors instead of torchdiffeq's chain of axpy kernels, and the reverse-time / adjoint minus
signs ride on the stage coefficients instead of negating [N,d] tensors.  A function may expose
`eval_vjp(t, y, cot, params)` (fused f + VJP); the adjoint driver uses it when present.
"""
import ctypes as C
import math

import torch

from . import _lib

_FIXED = ('euler', 'midpoint', 'rk4')
_ADAPTIVE = ('dopri5', 'adaptive_heun')

# Dormand-Prince 5(4), Shampine coefficients
_DP = dict(
    alpha=[1 / 5, 3 / 10, 4 / 5, 8 / 9, 1., 1.],
    beta=[[1 / 5], [3 / 40, 9 / 40], [44 / 45, -56 / 15, 32 / 9],
          [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
          [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
          [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84]],
    c_sol=[35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0],
    c_err=[35 / 384 - 1951 / 21600, 0, 500 / 1113 - 22642 / 50085, 125 / 192 - 451 / 720,
           -2187 / 6784 - -12231 / 42400, 11 / 84 - 649 / 6300, -1. / 60.],
    c_mid=[6025192743 / 30085553152 / 2, 0, 51252292925 / 65400821598 / 2, -2691868925 / 45128329728 / 2,
           187940372067 / 1594534317056 / 2, -1776094331 / 19743644256 / 2, 11237099 / 235043384 / 2],
    order=5, fsal=True)
_AH = dict(alpha=[1.], beta=[[1.]], c_sol=[0.5, 0.5], c_err=[0.5, -0.5], c_mid=None, order=2, fsal=False)


_FUSE_MIN = 1 << 14        # below this many elements the launch overhead of either form is the same
_FUSE_STAGES = True        # stage combinations in the epilogue of the function's kernel (tests switch it off to compare)


_KPTR = [None] + [C.c_void_p * n for n in range(1, 8)]       # ctypes array types, built once
_CVAL = [None] + [C.c_float * n for n in range(1, 8)]


def _stage_combine_raw(yi, ks, cs):
    # launch-bound solves (adaptive adjoint on small graphs) call this ~600 times per step: keep the host side lean
    n = len(ks)
    out = torch.empty_like(yi)
    rc = _lib.load().gb_stage_combine(yi.data_ptr(), _KPTR[n](*[k.data_ptr() for k in ks]), _CVAL[n](*cs), n,
                                      yi.numel(), out.data_ptr(), _lib.stream())
    if rc:
        _lib.check(rc, 'gb_stage_combine')
    return out


class _StageCombineFn(torch.autograd.Function):
    """the one-pass combination under autograd (plain `odeint` differentiated through the solver, e.g. adjoint=False
    training and CGNN): d/dyi = g, d/dk_j = c_j g"""

    @staticmethod
    def forward(ctx, cs, yi, *ks):
        ctx.cs = cs
        return _stage_combine_raw(yi, ks, cs)

    @staticmethod
    def backward(ctx, g):
        need = ctx.needs_input_grad
        return (None, g if need[1] else None) + tuple((g * c) if need[2 + j] else None for j, c in enumerate(ctx.cs))


def _combine_one(yi, terms):
    """yi + sum_j c_j * k_j: ONE pass (csrc/stage.cu) for fp32 CUDA tensors, else torchdiffeq's axpy chain"""
    terms = [(k, c) for k, c in terms if c != 0]
    if not terms:
        return yi
    if (yi.is_cuda and yi.dtype == torch.float32 and yi.numel() >= _FUSE_MIN and len(terms) <= 7 and yi.is_contiguous()
            and all(k.shape == yi.shape and k.dtype == yi.dtype and k.is_contiguous() for k, _ in terms)):
        ks, cs = [k for k, _ in terms], tuple(float(c) for _, c in terms)
        if torch.is_grad_enabled() and (yi.requires_grad or any(k.requires_grad for k in ks)):
            return _StageCombineFn.apply(cs, yi, *ks)
        return _stage_combine_raw(yi, ks, cs)
    acc = yi
    for k, c in terms:
        acc = torch.add(acc, k, alpha=c)
    return acc


def _combine(y, dt, ks, coeffs, signs=None, skip=()):
    """y + dt * sum_j coeffs[j] * ks[j], per tuple component; signs[i] (+-1) multiplies component i's
    derivatives (reverse-time / adjoint sign conventions folded into the coefficients instead of
    negating whole tensors).  Components listed in `skip` are passed through unchanged."""
    out = []
    for i, yi in enumerate(y):
        if i in skip:
            out.append(yi)
            continue
        sg = 1.0 if signs is None else signs[i]
        out.append(_combine_one(yi, [(k[i], dt * c * sg) for k, c in zip(ks, coeffs)]))
    return tuple(out)


_TIME_CACHE = {}


def _time_tensor(value, dtype, device):
    """0-d tensor holding a time point.  A fixed grid asks for the same few values pass after pass and every
    torch.full on a GPU is a kernel launch: the constants are kept (never written to; not while a CUDA graph is being
    captured, whose allocations belong to the graph's pool)."""
    key = (float(value), dtype, str(device))
    hit = _TIME_CACHE.get(key)
    if hit is not None:
        return hit
    t = torch.full((), float(value), dtype=dtype, device=device)
    if device.type != 'cuda' or not torch.cuda.is_current_stream_capturing():
        if len(_TIME_CACHE) >= 4096:
            _TIME_CACHE.clear()
        _TIME_CACHE[key] = t
    return t


def _rms(x):
    return x.pow(2).mean().sqrt()


def _segment_rms(v, lengths):
    """RMS of each consecutive segment of the flat tensor v (lengths: int64 tensor on v's device) - the packed
    parameter integrals of the adjoint are judged per parameter tensor, exactly like the unpacked components"""
    sq = v.detach().reshape(-1).pow(2)
    return torch.segment_reduce(sq, 'mean', lengths=lengths).sqrt()


def _mixed_norm(state, segments=None):
    """max over the components of their RMS (torchdiffeq's mixed norm) - ONE device read for all of them.
    `segments` = {component index: lengths tensor}: that component is a concatenation whose pieces each count as a
    component of their own."""
    if segments:
        vals = []
        for i, s in enumerate(state):
            if s.numel() == 0:
                continue
            if i in segments:
                vals.extend(_segment_rms(s, segments[i]).unbind(0))
            else:
                vals.append(_rms(s.detach()))
        return float(torch.stack(vals).max())
    vals = [_rms(s.detach()) for s in state if s.numel() > 0]
    if len(vals) == 1 or len({(v.device, v.dtype) for v in vals}) > 1:
        return max(float(v) for v in vals)
    return float(torch.stack(vals).max())


class _Problem:
    """normalises (func, y0, t) to tuple states and forward time"""

    def __init__(self, func, y0, t):
        self.is_tuple = isinstance(y0, (tuple, list))
        self.y0 = tuple(y0) if self.is_tuple else (y0,)
        ref = self.y0[0]
        self.times = [float(v) for v in (t.tolist() if torch.is_tensor(t) else t)]
        if len(self.times) < 2:
            raise ValueError('t must hold at least two time points')
        diffs = [b - a for a, b in zip(self.times[:-1], self.times[1:])]
        if not (all(d > 0 for d in diffs) or all(d < 0 for d in diffs)):
            raise ValueError('t must be strictly increasing or decreasing')
        self.sign = 1.0 if diffs[0] > 0 else -1.0
        self.times = [self.sign * v for v in self.times]
        self.dtype, self.device = ref.dtype, ref.device
        self._func = func
        self.comp_signs = None           # set by the adjoint driver: func returns raw VJPs (see _OdeintAdjoint)
        # components the dynamics never READ (the adjoint's parameter / time integrals): their intermediate stage
        # values are never formed, only the end-of-step combination (`passive`), or not even that (`dead`: the
        # caller drops the component).  Same end-of-step numbers as forming everything, fewer tiny kernels.
        self.passive, self.dead = (), ()
        self.segments = None             # adaptive solvers: see _mixed_norm

    def tensor_time(self, s):
        return _time_tensor(self.sign * s, self.dtype, self.device)

    def f_raw(self, s, state):
        """func at internal time s, as returned (no sign applied)"""
        t = self.tensor_time(s)
        out = self._func(t, state) if self.is_tuple else (self._func(t, state[0]),)
        return tuple(out)

    def f(self, s, state):
        """derivative in the internal (forward) time s"""
        out = self.f_raw(s, state)
        if self.sign < 0:
            out = tuple(-o for o in out)
        return out

    def signs(self, n):
        """per-component factor turning f_raw into the internal-time derivative"""
        comp = self.comp_signs if self.comp_signs is not None else (1.0,) * n
        return tuple(self.sign * c for c in comp)


def _fixed_grid(t0, t1, step_size):
    n = int(math.ceil((t1 - t0) / step_size + 1))
    grid = [t0 + k * step_size for k in range(n)]
    grid[-1] = t1
    return grid


def _fused_stages_ok(prob, y):
    """can the function form the next stage's state in the epilogue of its own kernel? (ODEFunc.forward_stage:
    plain tensor state, no autograd - inference and the forward solve of odeint_adjoint)"""
    return (_FUSE_STAGES and not prob.is_tuple and len(y) == 1 and not torch.is_grad_enabled()
            and getattr(prob._func, 'forward_stage', None) is not None and y[0].numel() >= _FUSE_MIN)


def _fixed_step_fused(prob, method, t0, dt, t1, y):
    """_fixed_step for a single-tensor state with every stage combination written by the kernel that produces the
    stage's f (same operations in the same order as gb_stage_combine: bit-identical states)."""
    fs, sg, y0 = prob._func.forward_stage, prob.signs(1)[0], y[0]
    A, b, c = _TABLEAU[method]
    rows = list(A) + [b]                                  # combination that FOLLOWS stage i: rows[i] (last: the step)
    ks, x = [], y0
    for i, row in enumerate(rows):
        s = t1 if c[i] == 1.0 else t0 + c[i] * dt
        coeffs = [dt * cf * sg for cf in row]
        prev = [(k, cf) for k, cf in zip(ks, coeffs[:-1]) if cf != 0]
        f, xn = fs(prob.tensor_time(s), x, y0, [k for k, _ in prev], [cf for _, cf in prev], coeffs[-1])
        ks.append(f)
        if xn is None:                                    # the route taken could not fuse: one separate pass
            xn = _combine_one(y0, list(zip(ks, coeffs)))
        x = xn
    return (x,)


def _fixed_step(prob, method, t0, dt, t1, y):
    if _fused_stages_ok(prob, y):
        return _fixed_step_fused(prob, method, t0, dt, t1, y)
    f = prob.f_raw                      # signs (reverse time, adjoint conventions) ride on the coefficients
    sg = prob.signs(len(y))
    mid, end = tuple(prob.passive) + tuple(prob.dead), tuple(prob.dead)
    if method == 'euler':
        return _combine(y, dt, [f(t0, y)], [1.0], sg, end)
    if method == 'midpoint':
        k1 = f(t0, y)
        k2 = f(t0 + 0.5 * dt, _combine(y, dt, [k1], [0.5], sg, mid))
        return _combine(y, dt, [k2], [1.0], sg, end)
    k1 = f(t0, y)
    k2 = f(t0 + dt / 3, _combine(y, dt, [k1], [1 / 3], sg, mid))
    k3 = f(t0 + dt * 2 / 3, _combine(y, dt, [k1, k2], [-1 / 3, 1.0], sg, mid))
    k4 = f(t1, _combine(y, dt, [k1, k2, k3], [1.0, -1.0, 1.0], sg, mid))
    return _combine(y, dt, [k1, k2, k3, k4], [0.125, 0.375, 0.375, 0.125], sg, end)


def _solve_fixed(prob, method, step_size, on_step=None):
    times = prob.times
    if step_size is None:
        grid = list(times)
    else:
        grid = _fixed_grid(times[0], times[-1], float(step_size))
    sols = [prob.y0]
    y = prob.y0
    j = 1
    for g0, g1 in zip(grid[:-1], grid[1:]):
        y1 = _fixed_step(prob, method, g0, g1 - g0, g1, y)
        if on_step is not None:
            on_step(prob.sign * g0, prob.sign * g1, y1)
        while j < len(times) and g1 >= times[j]:
            if times[j] == g1:
                sols.append(y1)
            else:                                   # linear interpolation onto a requested time
                w = (times[j] - g0) / (g1 - g0)
                sols.append(tuple(a + w * (b - a) for a, b in zip(y, y1)))
            j += 1
        y = y1
    return sols


def _initial_step(prob, t0, y0, order, rtol, atol, f0):
    scale = tuple(atol + y.abs() * rtol for y in y0)
    d0 = _mixed_norm(tuple(y / s for y, s in zip(y0, scale)), prob.segments)
    d1 = _mixed_norm(tuple(f / s for f, s in zip(f0, scale)), prob.segments)
    h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
    y1 = _combine(y0, h0, [f0], [1.0])
    f1 = prob.f(t0 + h0, y1)
    d2 = _mixed_norm(tuple((a - b) / s for a, b, s in zip(f1, f0, scale)), prob.segments) / h0
    if d1 <= 1e-15 and d2 <= 1e-15:
        h1 = max(1e-6, h0 * 1e-3)
    else:
        h1 = (0.01 / max(d1, d2)) ** (1.0 / float(order + 1))
    return min(100 * h0, h1)


def _interp4(y0, y1, y_mid, f0, f1, dt, x):
    out = []
    for a0, a1, am, g0, g1 in zip(y0, y1, y_mid, f0, f1):
        a = 2 * dt * (g1 - g0) - 8 * (a1 + a0) + 16 * am
        b = dt * (5 * g0 - 3 * g1) + 18 * a0 + 14 * a1 - 32 * am
        c = dt * (g1 - 4 * g0) - 11 * a0 - 5 * a1 + 16 * am
        d = dt * g0
        out.append(a0 + x * (d + x * (c + x * (b + x * a))))
    return tuple(out)


def _solve_adaptive(prob, method, rtol, atol, max_num_steps=2 ** 31 - 1, on_step=None, max_attempts=None):
    """on_step(t0, t1, y1) is called after every ATTEMPTED step with the last accepted state (what the reference's
    EarlyStopDopri5.advance evaluates, early_stop_solver.py:82-95); max_attempts bounds the attempts per requested
    time: when it is hit the solution at that time is the last accepted state instead of the dense output."""
    tab = _DP if method == 'dopri5' else _AH
    times = prob.times
    t, y = times[0], prob.y0
    f0 = prob.f(t, y)
    dt = _initial_step(prob, t, y, tab['order'] - 1, rtol, atol, f0)
    sols = [y]
    # interpolation window of the last accepted step
    win = None
    steps = 0
    for target in times[1:]:
        attempts = 0
        while t < target:
            if max_attempts is not None and attempts >= max_attempts:
                break
            attempts += 1
            if steps >= max_num_steps:
                raise RuntimeError('max_num_steps exceeded')
            steps += 1
            ks = [f0]
            for a_i, b_i in zip(tab['alpha'], tab['beta']):
                ks.append(prob.f(t + a_i * dt, _combine(y, dt, ks, b_i)))
            y1 = _combine(y, dt, ks, tab['c_sol'])
            err = _combine(tuple(torch.zeros_like(v) for v in y), dt, ks, tab['c_err'])
            tol = tuple(atol + rtol * torch.max(a.abs(), b.abs()) for a, b in zip(y, y1))
            ratio = _mixed_norm(tuple(e / s for e, s in zip(err, tol)), prob.segments)
            if ratio <= 1:
                # torchdiffeq's _runge_kutta_step returns f1 = k[..., -1] for EVERY tableau (exact for the FSAL
                # Dormand-Prince pair; for adaptive_heun it is f(t1, y + dt k1), no extra evaluation)
                f1 = ks[-1]
                y_mid = _combine(y, dt, ks, tab['c_mid']) if tab['c_mid'] else None
                win = (t, t + dt, y, y1, y_mid, f0, f1)
                t, y, f0 = t + dt, y1, f1
            if ratio == 0:
                factor = 10.0
            else:
                dfactor = 1.0 if ratio < 1 else 0.2
                factor = min(10.0, max(0.9 / ratio ** (1.0 / tab['order']), dfactor))
            dt = dt * factor
            if on_step is not None and win is not None:
                on_step(prob.sign * win[0], prob.sign * win[1], win[3])
        if win is None:                               # no step was accepted within the attempt budget
            sols.append(y)
            continue
        # t >= target: evaluate the dense output of the last accepted step at `target`
        w0, w1, ya, yb, ym, fa, fb = win
        if t < target:                                # attempt budget hit: the last accepted state stands in
            sols.append(yb)
        elif target == w1:
            sols.append(yb)
        elif ym is not None:
            sols.append(_interp4(ya, yb, ym, fa, fb, w1 - w0, (target - w0) / (w1 - w0)))
        else:
            x = (target - w0) / (w1 - w0)
            sols.append(tuple(a + x * (b - a) for a, b in zip(ya, yb)))
    return sols


def odeint(func, y0, t, **kwargs):
    """torchdiffeq.odeint(func, y0, t, rtol=, atol=, method=, options=) - see _odeint_impl"""
    from . import _lib
    with _lib.nvtx_range('grand_b200.odeint' if '_comp_signs' not in kwargs else 'grand_b200.adjoint_solve'):
        return _odeint_impl(func, y0, t, **kwargs)


def _odeint_impl(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, event_fn=None, _comp_signs=None,
                 _on_step=None, _max_attempts=None, _passive=(), _dead=(), _segments=None):
    if event_fn is not None:
        raise NotImplementedError('event handling is not part of the reference call sites')
    method = method or 'dopri5'
    options = dict(options or {})
    prob = _Problem(func, y0, t)
    if _comp_signs is not None:
        if method not in _FIXED:
            raise ValueError('_comp_signs is only meaningful for the fixed-grid solvers')
        prob.comp_signs = tuple(_comp_signs)
    if _passive or _dead:
        if method not in _FIXED:
            raise ValueError('_passive / _dead components are only supported by the fixed-grid solvers')
        prob.passive, prob.dead = tuple(_passive), tuple(_dead)
    prob.segments = _segments
    if method in _FIXED:
        sols = _solve_fixed(prob, method, options.get('step_size'), on_step=_on_step)
    elif method in _ADAPTIVE:
        sols = _solve_adaptive(prob, method, float(rtol), float(atol), on_step=_on_step, max_attempts=_max_attempts)
    else:
        raise NotImplementedError(f"method '{method}' (supported: {_FIXED + _ADAPTIVE})")
    out = tuple(torch.stack([s[i] for s in sols]) for i in range(len(prob.y0)))
    return out if prob.is_tuple else out[0]


# Butcher tableaus of the fixed-grid methods as (A rows, b, c): stage i > 0 is evaluated at y + dt * sum_j A[i-1][j] k_j
_TABLEAU = {
    'euler': ([], [1.0], [0.0]),
    'midpoint': ([[0.5]], [0.0, 1.0], [0.0, 0.5]),
    'rk4': ([[1 / 3], [-1 / 3, 1.0], [1.0, -1.0, 1.0]], [0.125, 0.375, 0.375, 0.125], [0.0, 1 / 3, 2 / 3, 1.0]),
}


def _adjoint_fixed_phased(func, params, state, t_hi, t_lo, method, step_size):
    """The fixed-grid reverse solve of the augmented adjoint state (vjp_t, y, a_y, *a_params) from real time t_hi down
    to t_lo for a function with the two-phase protocol (`eval_vjp_phased` / `prefetch_state`,
    ODEFuncTransformerAtt on a row-partitioned graph).  Same arithmetic, in the same order, as
    odeint(aug_dynamics, ..., _comp_signs=(-1, 1, -1, -1...)) over _fixed_step - only the ORDER OF ISSUE differs: as soon
    as the row pass of stage i is queued (f known) the y component of stage i+1 is formed and its all-gather started,
    and only then the column pass / dense tail / reduce-scatter of stage i are queued, so the gather overlaps them."""
    A, b, c = _TABLEAU[method]
    ref = state[1]
    s0, s1 = -float(t_hi), -float(t_lo)                              # internal (increasing) time, as in _Problem
    grid = [s0, s1] if step_size is None else _fixed_grid(s0, s1, float(step_size))
    n_state = len(state)
    sg = (1.0, -1.0) + (1.0,) * (n_state - 2)                       # prob.sign (-1) x comp_signs (-1, 1, -1, -1, ...)
    state = tuple(state)
    pre = None
    for step, (g0, g1) in enumerate(zip(grid[:-1], grid[1:])):
        dt = g1 - g0
        last_step = step == len(grid) - 2
        ks = []
        y, a_y = state[1], state[2]
        y_stage, a_stage = y, a_y
        for i in range(len(b)):
            s_i = g1 if (method == 'rk4' and i == 3) else g0 + c[i] * dt
            tt = _time_tensor(-s_i, ref.dtype, ref.device)
            f, finish = func.eval_vjp_phased(tt, y_stage, a_stage, params, prefetched=pre)
            fs = [k[1] for k in ks] + [f]
            coeffs = A[i] if i + 1 < len(b) else b
            y_next = _combine_one(y, [(fk, dt * cf * sg[1]) for fk, cf in zip(fs, coeffs)])
            pre = None if (i + 1 == len(b) and last_step) else func.prefetch_state(y_next)
            vy, vps = finish()
            vp = [g if g is not None else torch.zeros_like(p) for g, p in zip(vps, params)]
            ks.append((torch.zeros_like(tt), f, vy, *vp))
            if i + 1 < len(b):
                y_stage = y_next
                a_stage = _combine_one(a_y, [(k[2], dt * cf * sg[2]) for k, cf in zip(ks, A[i])])
        new = list(_combine(state, dt, ks, b, sg))
        new[1] = y_next                                              # the same sum, already formed for the prefetch
        state = tuple(new)
    return list(state)


_PACK_PARAMS = True      # adjoint: one flat buffer for all parameter integrals (see _OdeintAdjoint.backward)


class _OdeintAdjoint(torch.autograd.Function):

    @staticmethod
    def forward(ctx, func, t, cfg, n_params, y0, *params):
        with torch.no_grad():
            ys = odeint(func, y0, t, rtol=cfg['rtol'], atol=cfg['atol'], method=cfg['method'], options=cfg['options'])
        ctx.func, ctx.cfg, ctx.params = func, cfg, params
        ctx.times = [float(v) for v in (t.tolist() if torch.is_tensor(t) else t)]
        ctx.save_for_backward(ys)
        return ys

    @staticmethod
    def backward(ctx, grad_ys):
        func, cfg, params = ctx.func, ctx.cfg, ctx.params
        times = ctx.times
        ys, = ctx.saved_tensors

        fused_vjp = getattr(func, 'eval_vjp', None)
        fixed = cfg['adjoint_method'] in _FIXED
        # Real-time augmented dynamics (torchdiffeq): d/dt (vjp_t, y, a_y, a_p) = (-a.df/dt, f, -a.df/dy, -a.df/dp).
        # Fixed-grid solves take the RAW products (cotangent +a_y) and carry the minus signs on the stage
        # coefficients (comp_signs) - no negated copies of [N,d] tensors; adaptive solves get signed values.
        comp_signs = (-1.0, 1.0, -1.0) + (-1.0,) * len(params)
        # Fixed grid: the parameter integrals ride in ONE flat buffer (the VJPs of an evaluation are packed with one
        # cat), are only combined at the end of a step (nothing reads their stage values) and the time integral is
        # dropped altogether - the reference pays ~10 axpy kernels per parameter tensor per step for the same numbers.
        # Adaptive solves pack as well (a dopri5 step forms 8 combinations of every component: 12 parameter tensors
        # of the transformer function = ~650 launch-sized kernels per attempted step, 4 components = ~80); their error
        # norm still judges every parameter tensor on its own (`_segments`), as torchdiffeq's mixed norm does.
        packed = _PACK_PARAMS and len(params) > 0 and all(p.numel() > 0 for p in params) and \
            all(p.dtype == ys.dtype and p.device == ys.device for p in params)
        sizes = [p.numel() for p in params]
        segments = None
        # the ODE functions that evaluate their own VJP (eval_vjp) are autonomous: the time integral is identically
        # zero, never the maximum of the mixed norm, and is not returned - an adaptive packed solve leaves it out
        no_t = packed and not fixed and fused_vjp is not None
        off = 0 if no_t else 1                                     # index of y in the augmented state
        if packed and not fixed:
            segments = {off + 2: torch.tensor(sizes, dtype=torch.int64, device=ys.device)}
        zero_t = torch.zeros((), dtype=ys.dtype, device=ys.device)
        zero_piece = {}

        def pack(vps):
            parts = []
            for i, (g, p) in enumerate(zip(vps, params)):
                if g is None:
                    if i not in zero_piece:
                        zero_piece[i] = torch.zeros(sizes[i], dtype=ys.dtype, device=ys.device)
                    g = zero_piece[i]
                parts.append(g.reshape(-1))
            return torch.cat(parts)

        def aug_dynamics(tt, state):
            y, a_y = state[off], state[off + 1]
            cot = a_y if fixed else -a_y
            if fused_vjp is not None:
                # the function evaluates f and cot^T df/d(y, params) in one fused pass (same numbers as
                # func(t, y) followed by autograd.grad; df/dt == 0 for the autonomous ODE functions)
                res = fused_vjp(tt, y, cot, params)
                if res is not None:
                    f, vy, vps = res
                    if no_t:
                        return (f, vy, pack(vps))
                    if packed:
                        return (zero_t, f, vy, pack(vps))              # df/dt == 0: the time integral stays zero
                    vp = [g if g is not None else torch.zeros_like(p) for g, p in zip(vps, params)]
                    return (torch.zeros_like(tt), f, vy, *vp)
            with torch.enable_grad():
                tt_ = tt.detach().requires_grad_(True)
                y_ = y.detach().requires_grad_(True)
                f = func(tt_, y_)
                grads = torch.autograd.grad(f, (tt_, y_) + tuple(params), cot, allow_unused=True)
            vy = grads[1] if grads[1] is not None else torch.zeros_like(y)
            if no_t:
                return (f.detach(), vy, pack(grads[2:]))
            if packed:
                vt = zero_t if (fixed or grads[0] is None) else grads[0]
                return (vt, f.detach(), vy, pack(grads[2:]))
            vt = grads[0] if grads[0] is not None else torch.zeros_like(tt)
            vp = [g if g is not None else torch.zeros_like(p) for g, p in zip(grads[2:], params)]
            return (vt, f.detach(), vy, *vp)

        with torch.no_grad():
            if packed:
                state = [zero_t, ys[-1], grad_ys[-1], torch.zeros(sum(sizes), dtype=ys.dtype, device=ys.device)]
                comp_signs = (-1.0, 1.0, -1.0, -1.0)
                if no_t:
                    state = state[1:]
            else:
                state = [torch.zeros((), dtype=ys.dtype, device=ys.device), ys[-1], grad_ys[-1]] + \
                        [torch.zeros_like(p) for p in params]
            phased = fixed and getattr(func, 'phased_ok', None) is not None and func.phased_ok(ys[-1])
            for i in range(len(times) - 1, 0, -1):
                if phased:
                    if packed:                                     # the phased solver keeps one tensor per parameter
                        state = state[:3] + [v.reshape(p.shape) for v, p in zip(state[3].split(sizes), params)]
                    state = _adjoint_fixed_phased(func, params, state, times[i], times[i - 1], cfg['adjoint_method'],
                                                  cfg['adjoint_options'].get('step_size'))
                    if packed:
                        state = state[:3] + [torch.cat([v.reshape(-1) for v in state[3:]])]
                    state[1] = ys[i - 1]
                    state[2] = state[2] + grad_ys[i - 1]
                    continue
                sol = odeint(aug_dynamics, tuple(state), [times[i], times[i - 1]], rtol=cfg['adjoint_rtol'],
                             atol=cfg['adjoint_atol'], method=cfg['adjoint_method'], options=cfg['adjoint_options'],
                             _comp_signs=comp_signs if fixed else None, _passive=(3,) if (packed and fixed) else (),
                             _dead=(0,) if (packed and fixed) else (), _segments=segments)
                state = [s[1] for s in sol]
                state[off] = ys[i - 1]
                state[off + 1] = state[off + 1] + grad_ys[i - 1]
            if no_t:
                state = [zero_t] + state
            if packed:
                state = state[:3] + [v.reshape(p.shape) for v, p in zip(state[3].split(sizes), params)]
        return (None, None, None, None, state[2]) + tuple(state[3:])


def odeint_adjoint(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, event_fn=None,
                   adjoint_rtol=None, adjoint_atol=None, adjoint_method=None, adjoint_options=None,
                   adjoint_params=None):
    if event_fn is not None:
        raise NotImplementedError('event handling is not part of the reference call sites')
    if isinstance(y0, (tuple, list)):
        raise NotImplementedError('odeint_adjoint: tuple states (regularised ODE functions) are not supported')
    if adjoint_params is None:
        adjoint_params = tuple(func.parameters()) if isinstance(func, torch.nn.Module) else ()
    params = tuple(p for p in adjoint_params if p.requires_grad)
    cfg = dict(rtol=rtol, atol=atol, method=method or 'dopri5', options=dict(options or {}),
               adjoint_rtol=rtol if adjoint_rtol is None else adjoint_rtol,
               adjoint_atol=atol if adjoint_atol is None else adjoint_atol,
               adjoint_method=adjoint_method or method or 'dopri5',
               adjoint_options=dict(adjoint_options if adjoint_options is not None else (options or {})))
    return _OdeintAdjoint.apply(func, t, cfg, len(params), y0, *params)
