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


DEVICE = 'cuda'      # tests/test_host_plumbing_cpu.py sets 'cpu' while the emulated ABI is installed


def device():
    return torch.device(DEVICE)


def dev(a, dtype=None):
    t = torch.as_tensor(np.asarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(DEVICE)
