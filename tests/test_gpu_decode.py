"""GPU: SURVEY.md 8f-2 - the encoder / decoder layers on the library's own GEMMs (ops.Linear: K-sliced tcgen05 3xTF32
for the [N, F] x [F, d] encoder of src/GNN.py:30-38) and the per-step evaluation of the early-stopping integrator in
one launch (gb_decode_eval, csrc/decode.cu <- src/early_stop_solver.py:108-128, :88-98), against the fp64 oracle
(oracle/funcs.py decode_eval) and through EarlyStopInt itself."""
import types

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
RTOL = 2e-6          # of the result scale: fp32-class accuracy (plain TF32 would be ~1e-3)


def _close(got, ref, what, rtol=RTOL):
    scale = float(ref.abs().max())
    err = float((got.double().cpu() - ref.double().cpu()).abs().max())
    assert err <= rtol * scale, f'{what}: max err {err:.3e} at scale {scale:.3e}'


def _problem(n, d, classes, seed, ldz=None, device='cuda'):
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(n, ldz or d, generator=g)
    w = torch.randn(classes, d, generator=g) / d ** 0.5
    b = 0.1 * torch.randn(classes, generator=g)
    y = torch.randint(0, classes, (n,), generator=g)
    r = torch.rand(n, generator=g)
    masks = [r < 0.2, (r >= 0.2) & (r < 0.5), r >= 0.5]
    return z, w, b, y, masks


# (n, d, classes, ldz): Cora-like, augmented state (ldz = 2 d), ragged d (scalar staging), widest supported, C4-like
SHAPES = [(2708, 64, 7, None), (1000, 80, 47, 160), (33, 6, 3, None), (5000, 128, 64, None), (70001, 128, 47, None),
          (257, 100, 10, 100), (1, 64, 2, None)]


@pytest.mark.parametrize('n,d,classes,ldz', SHAPES)
def test_decode_eval_matches_the_oracle(n, d, classes, ldz):
    from grand_b200 import ops
    from oracle import funcs
    z, w, b, y, masks = _problem(n, d, classes, n + classes, ldz)
    logits_ref, pred_ref, _ = funcs.decode_eval(z.double(), w.double(), b.double(), y, masks)
    top2 = logits_ref.topk(min(2, classes), dim=-1).values
    clear = (top2[:, 0] - top2[:, -1]) > 1e-4 if classes > 1 else torch.ones(n, dtype=torch.bool)
    masks = [m & clear for m in masks]                        # rows whose arg-max an fp32 rounding could flip: not counted
    _, _, cnt_ref = funcs.decode_eval(z.double(), w.double(), b.double(), y, masks)
    dev = lambda t: t.cuda()
    counts, logits, pred = ops.decode_eval(dev(z), dev(w), dev(b), dev(y), [dev(m) for m in masks], want_logits=True,
                                           want_pred=True)
    _close(logits, logits_ref, 'logits', rtol=4e-6)
    assert torch.equal(pred.cpu()[clear], pred_ref[clear])
    assert counts.tolist() == [c for c, _ in cnt_ref] + [t for _, t in cnt_ref]
    assert counts.tolist()[3:] == [int(m.sum()) for m in masks]
    counts2, _, _ = ops.decode_eval(dev(z), dev(w), dev(b), dev(y), [dev(m) for m in masks])
    assert torch.equal(counts, counts2)                       # integer atomics: independent of arrival order


def test_decode_eval_without_bias_masks_or_counts():
    from grand_b200 import _lib, ops
    from oracle import funcs
    z, w, _, y, masks = _problem(500, 32, 5, 3)
    counts, logits, _ = ops.decode_eval(z.cuda(), w.cuda(), None, y.cuda(), [masks[0].cuda(), None, masks[2].cuda()],
                                        want_logits=True)
    logits_ref, _, cnt = funcs.decode_eval(z.double(), w.double(), None, y, [masks[0], None, masks[2]])
    _close(logits, logits_ref, 'logits without bias', rtol=4e-6)
    assert counts.tolist()[3:] == [int(masks[0].sum()), 0, int(masks[2].sum())]
    assert counts.tolist()[1] == 0
    lib = _lib.load()                                         # labels == NULL: prediction only
    zc, wc = z.cuda(), w.cuda()
    pred = torch.full((500,), -1, dtype=torch.int64, device='cuda')
    _lib.check(lib.gb_decode_eval(zc.data_ptr(), 32, 500, 32, wc.data_ptr(), 32, None, 5, None, None, None, None, None,
                                  None, 0, pred.data_ptr(), _lib.stream()), 'gb_decode_eval')
    assert torch.equal(pred.cpu(), logits.cpu().argmax(-1))


def test_decode_eval_ties_and_nan_follow_torch():
    """first maximum on exact ties; a NaN state row gives NaN logits, whose arg-max torch puts at the first NaN"""
    from grand_b200 import ops
    n, d, classes = 64, 16, 6
    z = torch.randn(n, d)
    z[:8] = -1.0                                              # relu -> 0: logits = bias
    z[8, 3] = float('nan')
    w = torch.randn(classes, d)
    b = torch.tensor([0.5, 2.0, 2.0, -1.0, 2.0, 0.0])
    y = torch.zeros(n, dtype=torch.int64)
    m = torch.ones(n, dtype=torch.bool)
    _, logits, pred = ops.decode_eval(z.cuda(), w.cuda(), b.cuda(), y.cuda(), [m.cuda(), None, None], want_logits=True,
                                      want_pred=True)
    ref = F.linear(F.relu(z.cuda()), w.cuda(), b.cuda())
    assert pred[:8].tolist() == [1] * 8
    assert torch.isnan(logits[8]).all() and int(pred[8]) == int(ref[8:9].argmax(-1)) == 0
    assert torch.equal(pred[9:], ref[9:].argmax(-1))


def test_decode_eval_refuses_more_classes_than_registers():
    from grand_b200 import _lib, ops
    assert _lib.load().gb_decode_eval_max_classes() == 64
    z, w, b, y, masks = _problem(100, 32, 65, 1)
    with pytest.raises(RuntimeError, match='classes'):
        ops.decode_eval(z.cuda(), w.cuda(), b.cuda(), y.cuda(), [m.cuda() for m in masks])


def test_decode_eval_full_size_properties():
    """C4 size (ogbn-products: 2 449 029 nodes, d = 128, 47 classes): totals equal the mask sizes exactly, the
    predictions agree with torch's own relu -> linear -> argmax except on fp32 near-ties, counts reproduce"""
    from grand_b200 import ops
    n, d, classes = 2_449_029, 128, 47
    g = torch.Generator(device='cuda').manual_seed(5)
    z = torch.randn(n, d, device='cuda', generator=g)
    w = torch.randn(classes, d, device='cuda', generator=g) / d ** 0.5
    b = torch.randn(classes, device='cuda', generator=g)
    y = torch.randint(0, classes, (n,), device='cuda', generator=g)
    r = torch.rand(n, device='cuda', generator=g)
    masks = [r < 0.08, (r >= 0.08) & (r < 0.1), r >= 0.1]
    counts, _, pred = ops.decode_eval(z, w, b, y, masks, want_pred=True)
    c = counts.tolist()
    assert c[3:] == [int(m.sum()) for m in masks] and sum(c[3:]) == n
    ref = F.linear(F.relu(z), w, b).argmax(-1)
    assert float((pred != ref).float().mean()) < 1e-5
    assert c[:3] == [int((pred[m] == y[m]).sum()) for m in masks]
    assert ops.decode_eval(z, w, b, y, masks)[0].tolist() == c


# ---- encoder / decoder layers -----------------------------------------------------------------------------------
# (rows, F, nout): Cora encoder (ragged K, unaligned rows), Pubmed, ogbn-products, ogbn-arxiv, decoder (CUDA-core tiles),
# few rows (tiles path), K of exactly one slice
LINEAR = [(2708, 1433, 64), (13000, 1433, 80), (19717, 500, 64), (50000, 100, 128), (40000, 128, 64), (2708, 64, 7),
          (100, 300, 64), (1000, 256, 16), (12500, 33, 32)]


@pytest.mark.parametrize('rows,k,nout', LINEAR)
@pytest.mark.parametrize('bias', [True, False])
def test_linear_layer_forward_and_backward(rows, k, nout, bias):
    from grand_b200 import ops
    torch.manual_seed(rows + k + nout)
    lin = ops.Linear(k, nout, bias=bias).cuda()
    x = torch.randn(rows, k, device='cuda', requires_grad=True)
    gy = torch.randn(rows, nout, device='cuda')
    y = lin(x)
    y.backward(gy)
    xd, wd, gd = x.detach().double(), lin.weight.detach().double(), gy.double()
    _close(y, xd @ wd.t() + (lin.bias.detach().double() if bias else 0), 'x W^T + b', rtol=4e-6)
    _close(x.grad, gd @ wd, 'dx', rtol=4e-6)
    _close(lin.weight.grad, gd.t() @ xd, 'dW', rtol=4e-6)
    if bias:
        _close(lin.bias.grad, gd.sum(0), 'db', rtol=4e-6)
    y2 = lin(x)
    assert torch.equal(y, y2)                                 # deterministic


def test_linear_layer_takes_the_sliced_tensor_core_path_for_wide_reductions_on_large_graphs():
    """[20000, 1433] x [1433, 64]: 12 launches of the tcgen05 kernel (K slices of 128); the Cora-sized encoder
    (2708 rows: 22 row tiles, launch-bound per slice) stays on the CUDA-core tiles - one launch"""
    from grand_b200 import _lib, ops
    lin = ops.Linear(1433, 64).cuda()
    for rows, launches in ((20000, 12), (2708, 1)):
        x = torch.randn(rows, 1433, device='cuda')
        before = _lib.launch_count()
        lin(x)
        assert _lib.launch_count() - before == launches


def test_linear_layer_on_the_batched_fork_layout_and_empty_input():
    from grand_b200 import ops
    lin = ops.Linear(40, 16).cuda()
    x = torch.randn(1, 300, 40, device='cuda')
    _close(lin(x), F.linear(x.double(), lin.weight.double(), lin.bias.double()), 'batched', rtol=4e-6)
    assert lin(x).shape == (1, 300, 16)
    assert lin(torch.empty(0, 40, device='cuda')).shape == (0, 16)


# ---- EarlyStopInt end to end --------------------------------------------------------------------------------------
def test_early_stop_integrator_on_the_device_matches_a_hand_rolled_loop():
    from grand_b200 import integrators
    from grand_b200.early_stop_solver import EarlyStopInt
    n, d, classes = 3000, 64, 7
    z, w, b, y, masks = _problem(n, d, classes, 11)
    A = (0.3 * torch.randn(d, d)).cuda()
    func = lambda t, v: v @ A.T
    data = types.SimpleNamespace(y=y.cuda(), train_mask=masks[0].cuda(), val_mask=masks[1].cuda(), test_mask=masks[2].cuda())
    opt = dict(max_test_steps=100, earlystopxT=2.0, method='rk4', step_size=0.5)
    es = EarlyStopInt(2.0, opt, device=torch.device('cuda'))
    es.data, es.m2_weight, es.m2_bias = data, w.cuda(), b.cuda()
    es(func, z.cuda(), torch.tensor([0.0, 1.0]), method='rk4', options=dict(step_size=0.5, max_iters=100))
    best, v = (0, 0, 0, 0), z.cuda()
    for k in range(8):
        v = integrators.odeint(func, v, torch.tensor([0.5 * k, 0.5 * (k + 1)]), method='rk4', options=dict(step_size=0.5))[1]
        logits = F.linear(F.relu(v), w.cuda(), b.cuda())
        acc = [float((logits[m].argmax(1) == data.y[m]).sum()) / int(m.sum()) for m in (data.train_mask, data.val_mask, data.test_mask)]
        if acc[1] > best[1]:
            best = (*acc, 0.5 * (k + 1))
    s = es.solver
    assert s.n_evaluations == 8
    assert (s.best_train, s.best_val, s.best_test) == pytest.approx(best[:3], abs=2.0 / 600) and s.best_time == pytest.approx(best[3])
