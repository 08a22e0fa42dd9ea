"""CPU: the single-gather route (NlTransformerFn / eval_vjp) and the row-partitioned path of the product's
Python host side through the emulated C ABI (tests/abi_emulator.py): rank-by-rank sums (ports of
tests/test_gpu_partition.py, verified on the B200), the block API with a 1-rank partition, and the REAL
block path of two processes over gloo (ODEblock.set_partition + Halo + odeint_adjoint) against the
single-process result."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import helpers
from helpers import assert_close, rand_graph
from oracle import funcs, graph as og

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU = torch.device('cpu')
T32 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32)
T64 = lambda a: torch.tensor(np.asarray(a.detach() if torch.is_tensor(a) else a), dtype=torch.float64)

OPT = {'hidden_dim': 32, 'block': 'attention', 'function': 'transformer', 'max_nfe': 10000, 'no_alpha_sigmoid': False,
       'add_source': True, 'augment': False, 'adjoint': True, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0,
       'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': 'rk4', 'step_size': 1.0, 'max_iters': 100,
       'adjoint_method': 'rk4', 'adjoint_step_size': 1.0, 'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 16,
       'attention_norm_idx': 0, 'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot',
       'square_plus': False, 'mix_features': False, 'multi_modal': False, 'kinetic_energy': None,
       'jacobian_norm2': None, 'total_deriv': None, 'directional_penalty': None}


def install_nl_route(emu, monkeypatch_setattr):
    """let the host pick the single-gather entry points although the tensors are not CUDA tensors"""
    from grand_b200 import function_transformer_attention as fta, ops
    emu.nl_enabled = True

    def nl_ok(plan, x, heads, adim, att_type, norm_idx, edge_w):
        return (norm_idx == 0 and edge_w is None and att_type == 0 and adim % heads == 0 and x.dim() == 2
                and bool(emu.gb_nl_supported(heads, adim // heads, x.shape[1])))
    monkeypatch_setattr(fta, 'nl_supported', nl_ok)
    monkeypatch_setattr(ops, 'nl_supported', nl_ok)


@pytest.fixture
def host(emulated_abi, monkeypatch):
    monkeypatch.setattr(helpers, 'DEVICE', 'cpu')
    install_nl_route(emulated_abi, monkeypatch.setattr)
    return emulated_abi


def make_func(d, heads, adim, seed=1):
    import grand_b200
    opt = dict(OPT, heads=heads, attention_dim=adim, hidden_dim=d)
    func = grand_b200.ODEFuncTransformerAtt(d, d, opt, CPU)
    g = torch.Generator().manual_seed(seed)
    L = func.multihead_att_layer
    with torch.no_grad():
        for lin in (L.Q, L.K):
            lin.weight.copy_(torch.randn(lin.weight.shape, generator=g) / d ** 0.5)
            lin.bias.copy_(0.1 * torch.randn(lin.bias.shape, generator=g))
        func.alpha_train.fill_(0.2)
        func.beta_train.fill_(-0.1)
    return func


@pytest.mark.parametrize('impl', ['2', '3'])
def test_single_gather_route_forward_gradients_and_fused_vjp(host, monkeypatch, impl):
    """impl 2: single-gather entry points (gb_nl_*); 3 (default): k-gather entry points (gb_nlk_*), d > attention_dim"""
    monkeypatch.setenv('GRANDB200_NL_IMPL', impl)
    n, e, d, heads, adim = 300, 2400, 32, 4, 16
    ei = rand_graph(n, e, 12, hub=(7, 300))
    rng = np.random.default_rng(5)
    x, x0, gout = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    func = make_func(d, heads, adim)
    L = func.multihead_att_layer
    func.edge_index, func.x0 = torch.tensor(ei), T32(x0)
    params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]
    xg = T32(x).requires_grad_(True)
    out = func(0.0, xg)
    grads = torch.autograd.grad(out, [xg] + params, T32(gout))
    fwd, adj = ('gb_nl_fwd', 'gb_nl_adjoint') if impl == '2' else ('gb_nlk_fwd', 'gb_nlk_adjoint')
    assert fwd in host.calls and adj in host.calls and 'gb_edge_softmax_fwd' not in host.calls
    assert not ({'gb_nl_fwd', 'gb_nl_adjoint', 'gb_nlk_fwd', 'gb_nlk_adjoint'} - {fwd, adj}) & set(host.calls)
    leaves = [T64(v).requires_grad_(True) for v in [x] + params]
    ref = funcs.transformer_f(leaves[0], torch.tensor(ei), leaves[1], leaves[2], leaves[3], leaves[4], heads, leaves[5],
                              leaves[6], T64(x0))
    rg = torch.autograd.grad(ref, leaves, T64(gout))
    assert_close(out, ref, 1e-5, 'f')
    bias_scale = max(float(rg[2].abs().max()), float(rg[4].abs().max()))
    for g, r, name in zip(grads, rg, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
        assert_close(g, r, 1e-4, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)
    f2, vx, vps = func.eval_vjp(0.0, T32(x), T32(gout), params)              # what the adjoint solver calls
    assert_close(f2, out, 1e-6, 'f of eval_vjp')
    for a, b, name in zip([vx] + vps, grads, ('gx', 'gwq', 'gbq', 'gwk', 'gbk', 'galpha', 'gbeta')):
        assert_close(a, b, 1e-6, name + ' of eval_vjp', floor=bias_scale if name in ('gbq', 'gbk') else 0.0)


@pytest.mark.parametrize('world', [1, 2, 3])
def test_transformer_ranks_sum_to_the_single_process_result(host, world):
    """port of tests/test_gpu_partition.py (rank emulation): every rank's local compute, summed"""
    import grand_b200
    from grand_b200 import partition as P
    n, d, heads, adim = 203, 32, 4, 16
    ei = rand_graph(n, 6 * n, 3, hub=(11, 280))
    rng = np.random.default_rng(1)
    x, x0, g = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    func = make_func(d, heads, adim)
    L = func.multihead_att_layer
    func.edge_index, func.x0 = torch.tensor(ei), T32(x0)
    xg = T32(x).requires_grad_(True)
    f_ref = func(0.0, xg)
    params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias, func.alpha_train, func.beta_train]
    ref = torch.autograd.grad(f_ref, [xg] + params, T32(g))
    part0 = P.RowPartition(n, world, 0)
    pad = lambda a: torch.cat([T32(a), torch.zeros(part0.n_pad - n, d)])
    x_full, x0_full, g_full = pad(x), pad(x0), pad(g)
    f_parts, dx_share, grads = [], 0, 0
    for r in range(world):
        part = P.RowPartition(n, world, r)
        ei_loc, _ = part.local_edges(torch.tensor(ei))
        plan = grand_b200.GraphPlan(ei_loc, part.n_pad)
        sl = slice(r * part.slab, (r + 1) * part.slab)
        x_slab, x0_slab, g_slab = x_full[sl].contiguous(), x0_full[sl].contiguous(), g_full[sl].contiguous()
        out, q = P.nl_local_forward(part, plan, x_full, x_slab, L.Q.weight.detach(), L.Q.bias.detach(),
                                    L.K.weight.detach(), func.alpha_train.detach(), func.beta_train.detach(), x0_slab,
                                    heads, True)
        f2, share, gr = P.nl_local_vjp(part, plan, x_full, q, L.Q.weight.detach(), L.K.weight.detach(),
                                       func.alpha_train.detach(), func.beta_train.detach(), x0_slab, heads, True, g_slab,
                                       True, True)
        assert_close(f2, out, 1e-6, 'f of the adjoint entry point vs the forward one')
        f_parts.append(out)
        dx_share, grads = dx_share + share, grads + gr
    assert_close(torch.cat(f_parts)[:n], f_ref, 2e-6, 'f')
    assert_close(dx_share[:n], ref[0], 1e-5, 'dx')
    dwq, dwk, dbq, dbk, dalpha, dbeta = P.unpack_nl_grads(grads, adim, d)
    bias_scale = float(ref[2].abs().max())
    for got, r, name in ((dwq, ref[1], 'dWq'), (dbq, ref[2], 'dbq'), (dwk, ref[3], 'dWk'), (dbk, ref[4], 'dbk'),
                         (dalpha, ref[5], 'dalpha'), (dbeta, ref[6], 'dbeta')):
        assert_close(got, r, 2e-5, name, floor=bias_scale if name == 'dbk' else 0.0)


@pytest.mark.parametrize('world', [1, 2, 5])
def test_laplacian_ranks_sum_to_the_single_process_result(host, world):
    import grand_b200
    from grand_b200 import partition as P
    n, d = 303, 16
    ei, w = og.reset_graph_data(rand_graph(n, 7 * n, 5, self_loops=False, hub=(11, 280)), None, n, 'rw', 1.0)
    rng = np.random.default_rng(1)
    x, x0, g = (rng.standard_normal((n, d)).astype(np.float32) for _ in range(3))
    opt = {'hidden_dim': d, 'block': 'constant', 'max_nfe': 100, 'no_alpha_sigmoid': False, 'add_source': True}
    func = grand_b200.LaplacianODEFunc(d, d, opt, CPU)
    with torch.no_grad():
        func.alpha_train.fill_(0.4)
        func.beta_train.fill_(-0.2)
    func.edge_index, func.edge_weight, func.x0 = torch.tensor(ei), T32(w), T32(x0)
    xg = T32(x).requires_grad_(True)
    f_ref = func(0.0, xg)
    dx_ref, da_ref, db_ref = torch.autograd.grad(f_ref, [xg, func.alpha_train, func.beta_train], T32(g))
    part0 = P.RowPartition(n, world, 0)
    pad = lambda a: torch.cat([T32(a), torch.zeros(part0.n_pad - n, d)])
    x_full, x0_full, g_full = pad(x), pad(x0), pad(g)
    f_parts, dx, dparams = [], 0, 0
    for r in range(world):
        part = P.RowPartition(n, world, r)
        ei_loc, w_loc = part.local_edges(torch.tensor(ei), T32(w))
        plan = grand_b200.GraphPlan(ei_loc, part.n_pad)
        vals = plan.to_csr_order(w_loc)
        sl = slice(r * part.slab, (r + 1) * part.slab)
        x0_slab, g_slab = x0_full[sl].contiguous(), g_full[sl].contiguous()
        f_parts.append(P.diffusion_local_forward(part, plan, vals, x_full, x0_slab, func.alpha_train.detach(),
                                                 func.beta_train.detach(), True))
        share, dp = P.diffusion_local_vjp(part, plan, vals, x_full, x0_slab, func.alpha_train.detach(), True, g_slab)
        dx, dparams = dx + share, dparams + dp
    assert_close(torch.cat(f_parts)[:n], f_ref, 2e-6, 'f')
    assert_close(dx[:n], dx_ref, 1e-5, 'dx')
    assert_close(dparams[0], da_ref, 2e-5, 'dalpha')
    assert_close(dparams[1], db_ref, 2e-5, 'dbeta')


# ---- two processes over gloo: the product's partitioned BLOCK path, end to end ----------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _train_pass(block, data, x, g):
    block.odefunc.nfe = 0
    xg = x.detach().requires_grad_(True)
    block.set_x0(xg)
    z = block(xg, data)
    params = [block.odefunc.alpha_train, block.odefunc.beta_train]
    L = getattr(block.odefunc, 'multihead_att_layer', None)
    if L is not None:
        params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias] + params
    grads = torch.autograd.grad((z * g).sum(), [xg] + params)
    return z.detach(), grads, block.odefunc.nfe


def _make_block(function='transformer', adjoint=True):
    import grand_b200
    opt = dict(OPT, function=function, block='attention' if function == 'transformer' else 'constant', adjoint=adjoint)
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, CPU, t=torch.tensor([0., 2.]))
    block.train()
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        if function == 'transformer':
            for lin in (block.odefunc.multihead_att_layer.Q, block.odefunc.multihead_att_layer.K):
                lin.weight.copy_(torch.randn(lin.weight.shape, generator=g) / 6)
                lin.bias.copy_(0.1 * torch.randn(lin.bias.shape, generator=g))
        block.odefunc.alpha_train.fill_(0.3)
        block.odefunc.beta_train.fill_(0.1)
    return block


def _gloo_worker(rank, world, port, n, d, ret, overlap=False, function='transformer', sparse=False, tiny=False,
                 adjoint=True):
    os.environ['GRANDB200_OVERLAP'] = '1' if overlap else '0'      # on by default since round 2; the serial form is pinned too
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import contextlib
    import torch.distributed as dist
    import grand_b200
    from grand_b200 import _lib
    from abi_emulator import Emulator
    emu = Emulator()
    _lib.load, _lib.require_cuda, _lib.stream = (lambda: emu), (lambda *a: None), (lambda: 0)
    torch.cuda.device = lambda dev: contextlib.nullcontext()
    install_nl_route(emu, setattr)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        # sparse halo: a graph thin enough that every slab references only part of the others
        ei = torch.tensor(og.shaped_graph(n, (n if sparse else 6 * n) // 2 * 2, 9))
        gen = torch.Generator().manual_seed(0)
        x, gz = torch.randn(n, d, generator=gen), torch.randn(n, d, generator=gen)
        data = grand_b200.GraphData()
        data.new_graph(ei, n)
        # single-process truth (same emulated ABI, un-partitioned block)
        z_ref, g_ref, nfe_ref = _train_pass(_make_block(function, adjoint), data, x, gz)
        # partitioned: this rank's slab in, this rank's slab out
        part = grand_b200.RowPartition(n, world, rank)
        halo = {False: grand_b200.Halo, True: grand_b200.SparseHalo, 'compact': grand_b200.CompactHalo}[sparse](part)
        block = _make_block(function, adjoint)
        block.set_partition(part, halo)
        del emu.calls[:]
        z, grads, nfe = _train_pass(block, data, part.to_slab(x), part.to_slab(gz))
        assert nfe == nfe_ref == (16 if adjoint else 8) and z.shape == (part.slab, d)
        assert halo.n_gathers == nfe and halo.n_scatters == 8         # one gather per evaluation, one scatter per VJP
        if sparse and not tiny:                                       # only referenced rows travelled
            assert 0 < halo.referenced_fraction < 0.9
            dense_bytes = 24 * (world - 1) * part.slab * d * 4
            assert halo.bytes_in < 0.9 * dense_bytes
        if sparse == 'compact':
            # the rank's kernels ran on the compact problem: S own rows + the referenced remote rows, not N_pad
            view = block.odefunc.partition[0]
            assert view.n_pad == part.slab + sum(halo.recv_splits) < part.n_pad and (view.lo, view.hi) == (0, part.hi - part.lo)
            ei_loc = block.odefunc.edge_index
            assert ei_loc.numel() == 0 or (int(ei_loc.max()) < view.n_pad and int(ei_loc[0].max()) < view.hi)
            # relabelling round trip: local ids map back to the global edges of this rank, in their original order
            ref_block = _make_block(function)
            ref_block.set_partition(part, grand_b200.Halo(part))
            ref_block.reset_graph_data(data, torch.float32)
            ei_glob = ref_block.odefunc.edge_index
            remote_ids = torch.cat([halo.recv_idx, halo.recv_idx.new_zeros(1)])     # never empty; the pad is not read
            cols = torch.where(ei_loc[1] < part.slab, ei_loc[1] + part.rank * part.slab,
                               remote_ids[(ei_loc[1] - part.slab).clamp(min=0)])
            assert torch.equal(ei_loc[0] + part.lo, ei_glob[0]) and torch.equal(cols, ei_glob[1])
            assert torch.equal(block.odefunc.edge_weight, ref_block.odefunc.edge_weight)
        if overlap:
            # two-phase adjoint: 7 of the 8 adjoint evaluations find their state already gathered, and every row pass
            # is followed by the next gather BEFORE its own column pass is issued
            assert getattr(halo, 'n_prefetched', 0) == 7
            if function == 'transformer':
                calls = [c for c in emu.calls if c.startswith('gb_nl_adjoint')]
                assert calls == ['gb_nl_adjoint_rows', 'gb_nl_adjoint_cols'] * 8
            else:
                calls = [c for c in emu.calls if c.startswith('gb_spmm_')][8:]            # after the forward integration
                assert calls == ['gb_spmm_fwd_part', 'gb_spmm_bwd_part'] * 8
        else:
            assert getattr(halo, 'n_prefetched', 0) == 0 and 'gb_nl_adjoint_rows' not in emu.calls
        assert_close(z, part.to_slab(z_ref), 2e-5, 'z slab')
        assert_close(grads[0], part.to_slab(g_ref[0]), 1e-4, 'dL/dx slab')
        if function == 'transformer':
            bias_scale = max(float(g_ref[2].abs().max()), float(g_ref[4].abs().max()))
            names = ('dWq', 'dbq', 'dWk', 'dbk', 'dalpha', 'dbeta')
        else:
            bias_scale, names = 0.0, ('dalpha', 'dbeta')
        for got, r, name in zip(grads[1:], g_ref[1:], names):
            assert_close(got, r, 1e-4, name, floor=bias_scale if name in ('dbq', 'dbk') else 0.0)   # all-reduced: full value
        ret[rank] = 'ok'
    finally:
        dist.destroy_process_group()


def test_two_rank_block_path_over_gloo_matches_the_single_process_block():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret), nprocs=world, join=True)
    assert dict(ret) == {0: 'ok', 1: 'ok'}


def test_two_rank_block_path_with_the_overlapped_exchange():
    """GRANDB200_OVERLAP=1: the phased adjoint solve (integrators._adjoint_fixed_phased) drives eval_vjp_phased /
    prefetch_state over real async gloo collectives; same numbers as the single-process block"""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret, True), nprocs=world, join=True)
    assert dict(ret) == {0: 'ok', 1: 'ok'}


@pytest.mark.parametrize('overlap', [False, True])
def test_two_rank_laplacian_block_path(overlap):
    """GRAND-l (constant block) on two gloo ranks, serial and overlapped exchange"""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret, overlap, 'laplacian'), nprocs=world, join=True)
    assert dict(ret) == {0: 'ok', 1: 'ok'}


@pytest.mark.parametrize('world,overlap,function', [(2, False, 'transformer'), (3, True, 'transformer'),
                                                     (3, False, 'laplacian'), (2, True, 'laplacian')])
def test_sparse_halo_block_path(world, overlap, function):
    """SparseHalo (only referenced rows exchanged, all_to_all_v) under the real block path: same numbers as the
    single-process block, fewer bytes than the dense exchange; serial and overlapped, GRAND-nl and GRAND-l"""
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret, overlap, function, True), nprocs=world, join=True)
    assert dict(ret) == {r: 'ok' for r in range(world)}


@pytest.mark.parametrize('world,overlap,function', [(2, False, 'transformer'), (3, True, 'transformer'),
                                                     (3, False, 'laplacian'), (2, True, 'laplacian')])
def test_compact_halo_block_path(world, overlap, function):
    """CompactHalo: the rank's graph relabelled to [own slab | referenced remote rows]; every buffer and kernel of the
    rank runs over S + R rows instead of N_pad, the all_to_all_v lands in / leaves from the tail of the local matrices.
    Same numbers as the single-process block."""
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret, overlap, function, 'compact'), nprocs=world, join=True)
    assert dict(ret) == {r: 'ok' for r in range(world)}


@pytest.mark.parametrize('halo', [False, True, 'compact'])
def test_rank_without_rows(halo):
    """9 nodes over 4 ranks: slabs of 3, the last rank owns no row at all (and the thin graph leaves ranks that
    reference nothing remote); dense, sparse and compact halos"""
    world = 4
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 9, 32, ret, False, 'transformer', halo, True), nprocs=world,
             join=True)
    assert dict(ret) == {r: 'ok' for r in range(world)}


def test_push_halo_world_of_one(host, monkeypatch, tmp_path):
    """PushHalo's host logic on a 1-rank gloo group with local memory standing in for the symmetric allocation: the
    gather is one gb_stage_combine_multi call (no stage terms, one destination per rank) + a barrier, buffers come
    from a ring of 3, autograd gets private copies; the block path equals the un-partitioned block"""
    import torch.distributed as dist
    import grand_b200
    n, d = 150, 32
    ei = torch.tensor(og.shaped_graph(n, 900, 9))
    gen = torch.Generator().manual_seed(0)
    x, gz = torch.randn(n, d, generator=gen), torch.randn(n, d, generator=gen)
    data = grand_b200.GraphData()
    data.new_graph(ei, n)
    base = _train_pass(_make_block(), data, x, gz)
    dist.init_process_group('gloo', init_method=f'file://{tmp_path}/rdzv', rank=0, world_size=1)
    try:
        part = grand_b200.RowPartition(n, 1, 0)
        halo = grand_b200.PushHalo(part)
        allocs, barriers = [], []

        def fake_symmetric(rows, width, dtype, device):
            t = torch.full((rows, width), float('nan'), dtype=dtype)
            allocs.append(t)
            return t, [t.data_ptr()], lambda: barriers.append(len(allocs))

        monkeypatch.setattr(halo, '_symmetric', fake_symmetric)
        block = _make_block()
        block.set_partition(part, halo)
        del host.calls[:]
        other = _train_pass(block, data, x, gz)
        assert host.calls.count('gb_stage_combine_multi') == halo.n_gathers == 16
        # the 8 VJPs formed their dL/dx share in symmetric buffers and were reduced by a pull (no collective)
        assert host.calls.count('gb_stage_combine') == halo.n_scatters == 8 and len(barriers) == 24
        assert len(allocs) == 6                                        # two rings (gathered state, dL/dx share) of 3
        assert_close(other[0], base[0], 2e-5, 'z')
        for a, b in zip(other[1], base[1]):
            assert_close(a, b, 2e-5, 'grad', floor=4e-3)   # bias gradients are rounding noise (sum of ds = 0); the un-partitioned base takes the k-gather route
        # backprop through the solver: 8 evaluations each keep their gathered state, more than the ring holds
        base_bp = _train_pass(_make_block(adjoint=False), data, x, gz)
        block_bp = _make_block(adjoint=False)
        block_bp.set_partition(part, halo)
        other_bp = _train_pass(block_bp, data, x, gz)
        assert_close(other_bp[0], base_bp[0], 2e-5, 'z (backprop)')
        for a, b in zip(other_bp[1], base_bp[1]):
            assert_close(a, b, 2e-5, 'grad (backprop)', floor=4e-3)
        with torch.no_grad():                                          # no autograd: the ring buffer itself is handed out
            full = halo.all_gather_rows(x)
        assert any(full is t for t in allocs) and torch.equal(full, x)
        with torch.enable_grad():
            assert not any(halo.all_gather_rows(x) is t for t in allocs)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('halo,function', [(False, 'transformer'), (True, 'transformer'), ('compact', 'transformer'),
                                           (True, 'laplacian')])
def test_backprop_through_the_solver_keeps_every_gathered_state(halo, function):
    """adjoint=False: autograd runs back through all 8 evaluations, each of which saved ITS gathered state - a halo that
    recycles gather buffers must not hand a recycled one to an evaluation that will be differentiated"""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, _free_port(), 157, 32, ret, False, function, halo, False, False), nprocs=world,
             join=True)
    assert dict(ret) == {r: 'ok' for r in range(world)}


def test_block_api_world_of_one_matches_unpartitioned(host):
    """port of tests/test_gpu_partition.py::test_block_api_world_of_one_matches_unpartitioned"""
    import grand_b200
    n, d = 150, 32
    ei = torch.tensor(og.shaped_graph(n, 900, 9))
    gen = torch.Generator().manual_seed(0)
    x, gz = torch.randn(n, d, generator=gen), torch.randn(n, d, generator=gen)
    data = grand_b200.GraphData()
    data.new_graph(ei, n)
    base = _train_pass(_make_block(), data, x, gz)
    block = _make_block()
    part = grand_b200.RowPartition(n, 1, 0)
    block.set_partition(part, grand_b200.LoopbackHalo(part))
    other = _train_pass(block, data, x, gz)
    assert_close(other[0], base[0], 2e-5, 'z')
    for a, b in zip(other[1], base[1]):
        assert_close(a, b, 2e-5, 'grad', floor=4e-3)   # bias gradients are rounding noise (sum of ds = 0); the un-partitioned base takes the k-gather route


def test_two_phase_adjoint_equals_the_single_call(host, monkeypatch):
    """host plumbing of tests/test_gpu_variants.py::test_two_phase_adjoint_is_bit_identical_to_the_single_call"""
    import test_gpu_variants as zz
    zz.test_two_phase_adjoint_is_bit_identical_to_the_single_call(monkeypatch)
    calls = [c for c in host.calls if c.startswith('gb_nl_adjoint')]
    assert calls == ['gb_nl_adjoint'] * 8 + ['gb_nl_adjoint_rows', 'gb_nl_adjoint_cols'] * 8


@pytest.mark.parametrize('adjoint', [False, True])
def test_attention_block_transformer_function_vs_the_oracle(host, adjoint):
    """port of tests/test_gpu_blocks.py::test_attention_block_transformer_function (the headline configuration:
    AttODEblock + ODEFuncTransformerAtt, rk4, odeint_adjoint with the fused eval_vjp) through the single-gather route"""
    import grand_b200
    from oracle import integrate as oint
    n, e, d, T = 200, 1400, 16, 2.0
    opt = dict(OPT, hidden_dim=d, adjoint=adjoint)
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, CPU, t=torch.tensor([0., T]))
    ei = og.shaped_graph(n, e, 2)
    data = grand_b200.GraphData()
    data.new_graph(torch.tensor(ei), n)
    block.train()
    func = block.odefunc
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for layer in (func.multihead_att_layer, block.multihead_att_layer):
            for lin in (layer.Q, layer.K):
                lin.weight.copy_(torch.randn(lin.weight.shape, generator=g) / 4.0)
                lin.bias.copy_(0.1 * torch.randn(lin.bias.shape, generator=g))
        func.alpha_train.fill_(-0.3)
        func.beta_train.fill_(0.5)
    x = torch.randn(n, d, generator=g).requires_grad_(True)
    gout = torch.randn(n, d, generator=g)
    block.set_x0(x)
    z = block(x, data)
    L = func.multihead_att_layer
    params = [x, func.alpha_train, func.beta_train, L.Q.weight, L.Q.bias, L.K.weight, L.K.bias]
    grads = torch.autograd.grad(z, params, gout)
    assert 'gb_nl_adjoint' in host.calls or 'gb_nlk_adjoint' in host.calls   # both modes differentiate through a fused entry point
    ei_n, _ = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    leaves = [T64(p).requires_grad_(True) for p in params]
    f = lambda t, y: funcs.transformer_f(y, torch.tensor(ei_n), leaves[3], leaves[4], leaves[5], leaves[6], 4, leaves[1],
                                         leaves[2], T64(x))
    if adjoint:
        zo = oint.odeint_adjoint(f, leaves[0], [0., T], 'rk4', 1.0, adjoint_params=tuple(leaves[1:]))[1]
    else:
        zo = oint.odeint(f, leaves[0], [0., T], 'rk4', 1.0)[1]
    ro = torch.autograd.grad(zo, leaves, T64(gout))
    assert_close(z, zo, 1e-4, 'z(T)')
    bias_scale = max(float(ro[4].abs().max()), float(ro[6].abs().max()))
    for got, r, name in zip(grads, ro, ('gx', 'galpha', 'gbeta', 'gwq', 'gbq', 'gwk', 'gbk')):
        assert_close(got, r, 1e-4, name, floor=bias_scale if name in ('gbq', 'gbk') else 0.0)
