"""shared helpers of the GPU parity tests (the oracle is the checker, never the thing under test)"""
import numpy as np
import torch

from oracle import graph as og


def rand_graph(n, e, seed, self_loops=True, hub=None):
    """random directed multigraph [2,E] int64 (duplicates allowed); optional hub = (node, degree)"""
    rng = np.random.default_rng(seed)
    row = rng.integers(0, n, e)
    col = rng.integers(0, n, e)
    if hub is not None:
        node, deg = hub
        row = np.concatenate([row, np.full(deg, node)])
        col = np.concatenate([col, rng.integers(0, n, deg)])
        row = np.concatenate([row, rng.integers(0, n, deg)])
        col = np.concatenate([col, np.full(deg, node)])
    if self_loops:
        row = np.concatenate([row, np.arange(n)])
        col = np.concatenate([col, np.arange(n)])
    perm = rng.permutation(row.size)
    return np.stack([row[perm], col[perm]]).astype(np.int64)


def normalized_graph(n, e, seed):
    """shaped graph + self loops + rw normalisation via the oracle: (edge_index int64 [2,E'], w f32 [E'])"""
    ei = og.shaped_graph(n, e, seed)
    return og.reset_graph_data(ei, None, n, 'rw', 1.0)


def assert_close(got, ref, rtol, what='', floor=0.0):
    """|got - ref| <= rtol * (|ref| + max(max|ref|, floor)): elementwise relative with a floor at the tensor's
    scale; `floor` is for quantities that are mathematically zero (e.g. the K-bias gradient under row softmax)"""
    got = torch.as_tensor(got).detach().double().cpu()
    ref = torch.as_tensor(ref).detach().double().cpu()
    assert got.shape == ref.shape, f'{what}: shape {tuple(got.shape)} vs {tuple(ref.shape)}'
    scale = max(float(ref.abs().max()) if ref.numel() else 0.0, floor)
    err = (got - ref).abs()
    bound = rtol * (ref.abs() + scale) + 1e-30
    worst = float((err / bound).max()) if ref.numel() else 0.0
    assert worst <= 1.0, f'{what}: max err {float(err.max()):.3e} at scale {scale:.3e} exceeds rtol {rtol} by x{worst:.2f}'


def error_report(got, ref):
    """the three error figures the parity gates are read with (SURVEY.md 8d "parity gates", hard part 1):
    normwise  max|d| / max|ref|                                   (what assert_close bounds, up to its |ref| term)
    per_row   max_i ( max_j |d_ij| / max(max_j |ref_ij|, 1e-6 max|ref|) )   rows judged on their own scale
    element   max_ij |d_ij| / max(|ref_ij|, 1e-6 max|ref|)        the literal element-wise reading of the survey gate;
              in fp32 it is dominated by entries that are small through cancellation (|ref_ij| << row scale)"""
    got = torch.as_tensor(got).detach().double().cpu().reshape(got.shape[0], -1)
    ref = torch.as_tensor(ref).detach().double().cpu().reshape(ref.shape[0], -1)
    d = (got - ref).abs()
    top = float(ref.abs().max())
    floor = 1e-6 * top
    return {'normwise': float(d.max()) / max(top, 1e-300),
            'per_row': float((d.max(dim=1).values / ref.abs().max(dim=1).values.clamp_min(floor)).max()),
            'element': float((d / ref.abs().clamp_min(floor)).max())}


def assert_rows_close(got, ref, rtol, what=''):
    """per-row gate: every row's worst absolute error is within rtol of that ROW's largest entry (a scaling error on a
    low-magnitude row passes the normwise gate; it does not pass this one).  Returns the error_report."""
    rep = error_report(got, ref)
    assert rep['per_row'] <= rtol, f'{what}: per-row relative error {rep["per_row"]:.3e} exceeds {rtol} ({rep})'
    return rep


DEVICE = 'cuda'      # tests/test_host_plumbing_cpu.py sets 'cpu' while the emulated ABI is installed


def device():
    return torch.device(DEVICE)


def dev(a, dtype=None):
    t = torch.as_tensor(np.asarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(DEVICE)
