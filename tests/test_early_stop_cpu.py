"""CPU (C ABI emulated): EarlyStopInt (mirror of the reference's src/early_stop_solver.py) over the restated solvers: the per-step
decode + best-validation bookkeeping against a hand-rolled loop, the returned solution against plain odeint, the
attempt budget of the adaptive solver."""
import types

import pytest
import torch
import torch.nn.functional as F


def setup(n=60, d=6, classes=3, seed=0):
    g = torch.Generator().manual_seed(seed)
    A = 0.8 * torch.randn(d, d, generator=g, dtype=torch.float32)
    func = lambda t, y: y @ A.T
    y0 = torch.randn(n, d, generator=g, dtype=torch.float32)
    w = torch.randn(classes, d, generator=g, dtype=torch.float32)
    b = torch.randn(classes, generator=g, dtype=torch.float32)
    labels = torch.randint(0, classes, (n,), generator=g)
    idx = torch.randperm(n, generator=g)
    masks = [torch.zeros(n, dtype=torch.bool) for _ in range(3)]
    for m, sl in zip(masks, (idx[:20], idx[20:40], idx[40:])):
        m[sl] = True
    data = types.SimpleNamespace(y=labels, train_mask=masks[0], val_mask=masks[1], test_mask=masks[2])
    return func, y0, w, b, data


def accs(y, w, b, data):
    logits = F.linear(F.relu(y), w, b)
    return [float((logits[m].argmax(1) == data.y[m]).sum()) / int(m.sum()) for m in (data.train_mask, data.val_mask, data.test_mask)]


def make(opt_over, T=2.0):
    from grand_b200.early_stop_solver import EarlyStopInt
    opt = dict(max_test_steps=100, earlystopxT=3.0, method='rk4', step_size=0.5)
    opt.update(opt_over)
    return EarlyStopInt(T, opt, device=torch.device('cpu')), opt


def test_rk4_tracks_the_best_validation_step_and_returns_the_plain_solution(emulated_abi):
    from grand_b200 import integrators
    func, y0, w, b, data = setup()
    es, opt = make({})
    es.data, es.m2_weight, es.m2_bias = data, w, b
    sol = es(func, y0, torch.tensor([0.0, 1.0]), method='rk4', options=dict(step_size=0.5, max_iters=100))
    assert torch.equal(es.t, torch.tensor([0.0, 6.0]))                       # [0, earlystopxT * T], the `t` argument is ignored
    ref = integrators.odeint(func, y0, torch.tensor([0.0, 6.0]), method='rk4', options=dict(step_size=0.5))
    assert torch.equal(sol, ref)
    best, y = (0, 0, 0, 0), y0
    for k in range(12):                                                      # hand-rolled: one rk4 step at a time
        y = integrators.odeint(func, y, torch.tensor([0.5 * k, 0.5 * (k + 1)]), method='rk4', options=dict(step_size=0.5))[1]
        tr, va, te = accs(y, w, b, data)
        if va > best[1]:
            best = (tr, va, te, 0.5 * (k + 1))
    s = es.solver
    assert s.n_evaluations == 12
    assert (s.best_train, s.best_val, s.best_test) == pytest.approx(best[:3]) and s.best_time == pytest.approx(best[3])


def test_batched_fork_layout_and_augmented_state(emulated_abi):
    func, y0, w, b, data = setup()
    es, _ = make({})
    es.data, es.m2_weight, es.m2_bias = data, w, b
    aug = torch.cat([y0, torch.zeros_like(y0)], dim=1)[None]                 # [1, N, 2d]: decode reads the first half
    func2 = lambda t, y: torch.cat([func(t, y[..., :6]), torch.zeros_like(y[..., 6:])], dim=-1)
    sol = es(func2, aug, None, options=dict(step_size=0.5))
    assert sol.shape == (2, 1, 60, 12)
    s1 = es.solver
    es(func, y0, None, options=dict(step_size=0.5))
    assert (s1.best_val, s1.best_test, s1.best_time) == (es.solver.best_val, es.solver.best_test, es.solver.best_time)


def test_dopri5_evaluates_every_attempt_and_honours_the_attempt_budget(emulated_abi):
    from grand_b200 import integrators
    func, y0, w, b, data = setup(seed=3)
    es, _ = make(dict(method='dopri5', earlystopxT=1.0))
    es.data, es.m2_weight, es.m2_bias = data, w, b
    sol = es(func, y0, None, rtol=1e-6, atol=1e-8)
    ref = integrators.odeint(func, y0, torch.tensor([0.0, 2.0]), method='dopri5', rtol=1e-6, atol=1e-8)
    assert torch.equal(sol, ref) and es.solver.n_evaluations >= 3 and 0 < es.solver.best_time <= 2.0 + 1e-9
    n_full = es.solver.n_evaluations
    es2, _ = make(dict(method='dopri5', earlystopxT=1.0, max_test_steps=2))
    es2.data, es2.m2_weight, es2.m2_bias = data, w, b
    short = es2(func, y0, None, rtol=1e-6, atol=1e-8)
    assert es2.solver.n_evaluations == 2 < n_full
    assert not torch.equal(short[1], ref[1])                                 # stopped early: last accepted state, not y(T)
    assert es2.solver.best_time < 2.0


def test_requires_data_and_only_rk4_or_dopri5(emulated_abi):
    func, y0, w, b, data = setup()
    es, _ = make({})
    with pytest.raises(RuntimeError):
        es(func, y0, None)
    es3, _ = make(dict(method='euler'))
    es3.data, es3.m2_weight, es3.m2_bias = data, w, b
    with pytest.raises(AssertionError):
        es3(func, y0, None)
