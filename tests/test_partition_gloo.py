"""CPU, world_size 2 over gloo: the host side of the row-partitioned path (RowPartition bookkeeping, the Halo
collectives, and the partition algebra - who adds the self term, what is reduce-scattered, what is
all-reduced).  The per-rank device compute is stood in for by the oracle restricted to the rank's own edges;
the CUDA kernels themselves are checked rank by rank in test_gpu_partition.py."""
import os
import socket
import sys

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, d, heads, adim, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import torch.distributed as dist
    from grand_b200.partition import RowPartition, Halo
    from oracle import funcs, graph as og
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                                     # same global problem on every rank
        ei = torch.tensor(og.reset_graph_data(og.shaped_graph(n, 6 * n // 2 * 2, 4), None, n, 'rw', 1.0)[0])
        f64 = dict(dtype=torch.float64)
        x, x0, g = (torch.randn(n, d, **f64) for _ in range(3))
        wq, wk = (torch.randn(adim, d, **f64) / d ** 0.5 for _ in range(2))
        bq, bk = (0.1 * torch.randn(adim, **f64) for _ in range(2))
        alpha, beta = torch.tensor(0.3, **f64), torch.tensor(0.1, **f64)
        part = RowPartition(n, world, rank)
        halo = Halo(part)
        assert part.n_pad == part.slab * world and part.bounds(rank) == (part.lo, part.hi)
        ei_loc, _ = part.local_edges(ei)
        assert bool(((ei_loc[0] >= part.lo) & (ei_loc[0] < part.hi)).all())
        counts = torch.tensor([ei_loc.shape[1]])
        dist.all_reduce(counts)
        assert int(counts) == ei.shape[1]                        # the slabs tile the edge set

        # ---- forward: gather the slabs, evaluate the rank's rows --------------------------------------
        x_full = halo.all_gather_rows(part.to_slab(x)).requires_grad_(True)
        assert torch.equal(x_full[:n].detach(), x) and float(x_full[n:].abs().sum()) == 0.0
        leaves = [t.clone().requires_grad_(True) for t in (wq, bq, wk, bk, alpha, beta)]
        x0_full = torch.cat([x0, torch.zeros(part.n_pad - n, d, **f64)])
        f_all = funcs.transformer_f(x_full, ei_loc, leaves[0], leaves[1], leaves[2], leaves[3], heads, leaves[4],
                                    leaves[5], x0_full)
        f_slab = part.to_slab(f_all)                              # rows outside [lo, hi) have no edges here
        # ---- backward: share for all columns -> reduce-scatter; parameter shares -> all-reduce ----------
        g_slab = part.to_slab(g)
        grads = torch.autograd.grad(f_slab, [x_full] + leaves, g_slab)
        dx_slab = halo.reduce_scatter_rows(grads[0])
        flat = torch.cat([t.reshape(-1) for t in grads[1:]])
        halo.all_reduce_(flat)

        # ---- single-process truth ----------------------------------------------------------------------
        xr = x.clone().requires_grad_(True)
        leaves_r = [t.clone().requires_grad_(True) for t in (wq, bq, wk, bk, alpha, beta)]
        f_ref = funcs.transformer_f(xr, ei, *leaves_r[:4], heads, leaves_r[4], leaves_r[5], x0)
        ref = torch.autograd.grad(f_ref, [xr] + leaves_r, g)
        torch.testing.assert_close(f_slab, part.to_slab(f_ref.detach()), rtol=1e-10, atol=1e-12)
        torch.testing.assert_close(dx_slab, part.to_slab(ref[0]), rtol=1e-9, atol=1e-11)
        torch.testing.assert_close(flat, torch.cat([t.reshape(-1) for t in ref[1:]]), rtol=1e-9, atol=1e-11)

        # ---- GRAND-l with fixed weights: same algebra through laplacian_f -------------------------------
        ei2, w2 = og.reset_graph_data(og.shaped_graph(n, 4 * n // 2 * 2, 7), None, n, 'rw', 1.0)
        ei2, w2 = torch.tensor(ei2), torch.tensor(w2, **f64)
        e_loc, w_loc = part.local_edges(ei2, w2)
        xf = halo.all_gather_rows(part.to_slab(x)).requires_grad_(True)
        a2, b2 = alpha.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        fl = part.to_slab(funcs.laplacian_f(xf, e_loc, w_loc, a2, b2, x0_full))
        gl = torch.autograd.grad(fl, [xf, a2, b2], g_slab)
        dxl = halo.reduce_scatter_rows(gl[0])
        pl = halo.all_reduce_(torch.stack([gl[1], gl[2]]))
        xr2, a3, b3 = x.clone().requires_grad_(True), alpha.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        fr = funcs.laplacian_f(xr2, ei2, w2, a3, b3, x0)
        gr = torch.autograd.grad(fr, [xr2, a3, b3], g)
        torch.testing.assert_close(fl, part.to_slab(fr.detach()), rtol=1e-10, atol=1e-12)
        torch.testing.assert_close(dxl, part.to_slab(gr[0]), rtol=1e-9, atol=1e-11)
        torch.testing.assert_close(pl, torch.stack([gr[1], gr[2]]), rtol=1e-9, atol=1e-11)
        assert halo.bytes_in > 0
        ret[rank] = 'ok'
    finally:
        dist.destroy_process_group()


def test_two_rank_partition_algebra_over_gloo():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), 157, 12, 2, 8, ret), nprocs=world, join=True)
    assert dict(ret) == {0: 'ok', 1: 'ok'}


def test_row_partition_bookkeeping():
    sys.path.insert(0, ROOT)
    from grand_b200.partition import RowPartition, LoopbackHalo
    for n, world in ((10, 4), (2449029, 8), (16, 8), (3, 8), (1 << 24, 8)):
        covered = 0
        for r in range(world):
            p = RowPartition(n, world, r)
            assert p.n_pad >= n and p.n_pad - n < world * 1 + p.slab and 0 <= p.hi - p.lo <= p.slab
            assert p.lo == min(r * p.slab, n)
            covered += p.hi - p.lo
        assert covered == n
    p = RowPartition(7, 1, 0)
    h = LoopbackHalo(p)
    t = torch.arange(14.).reshape(7, 2)
    assert torch.equal(h.all_gather_rows(t), t) and torch.equal(h.reduce_scatter_rows(t), t)
    ei = torch.tensor([[0, 3, 6, 6, 2], [1, 1, 0, 6, 5]])
    p1 = RowPartition(7, 2, 1)                                   # rows [4, 7)
    e_loc, w_loc = p1.local_edges(ei, torch.arange(5.))
    assert e_loc.tolist() == [[6, 6], [0, 6]] and w_loc.tolist() == [2., 3.]
    assert np.allclose(p1.to_slab(t).numpy(), np.vstack([t[4:7].numpy(), np.zeros((1, 2))]))
