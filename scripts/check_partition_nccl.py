#!/usr/bin/env python
"""Parity of the row-partitioned block path over NCCL, every halo kind, serial and overlapped exchange.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
        scripts/check_partition_nccl.py [--nodes 20000] [--rmat 14]

Every rank integrates the SAME graph un-partitioned on its own GPU (the hardware-verified single-GPU path), then takes
part in the partitioned solve (ODEblock.set_partition) with Halo / SparseHalo / CompactHalo, GRANDB200_OVERLAP off and
on, GRAND-nl (attention block) and GRAND-l (constant block), and compares its slab of z(T) and of dL/dx and the
all-reduced parameter gradients with the un-partitioned numbers.  The CPU suite runs the same protocol over gloo through
the emulated ABI (tests/test_host_partition_cpu.py); this is the same check on the real library and NCCL.
Prints one line per case and exits non-zero on the first mismatch (rank 0 reports)."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import grand_b200  # noqa: E402  (the shim at the repo root loads graph-neural-pde_b200/)
from grand_b200 import synth  # noqa: E402

OPT = {'hidden_dim': 64, 'max_nfe': 10 ** 6, 'no_alpha_sigmoid': False, 'add_source': True, 'augment': False,
       'adjoint': True, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0, 'data_norm': 'rw', 'self_loop_weight': 1.0,
       'method': 'rk4', 'step_size': 1.0, 'max_iters': 100, 'adjoint_method': 'rk4', 'adjoint_step_size': 1.0,
       'leaky_relu_slope': 0.2, 'heads': 4, 'attention_dim': 64, 'attention_norm_idx': 0, 'reweight_attention': False,
       'beltrami': False, 'attention_type': 'scaled_dot', 'square_plus': False, 'mix_features': False,
       'multi_modal': False, 'kinetic_energy': None, 'jacobian_norm2': None, 'total_deriv': None,
       'directional_penalty': None}


def make_block(function, device, adjoint=True):
    opt = dict(OPT, function=function, block='attention' if function == 'transformer' else 'constant', adjoint=adjoint)
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device, t=torch.tensor([0., 2.])).to(device)
    block.train()
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        if function == 'transformer':
            for lin in (block.odefunc.multihead_att_layer.Q, block.odefunc.multihead_att_layer.K):
                lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / 8).to(device))
                lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=g)).to(device))
        block.odefunc.alpha_train.fill_(0.3)
        block.odefunc.beta_train.fill_(0.1)
    return block


def train_pass(block, data, x, g):
    block.odefunc.nfe = 0
    xg = x.detach().requires_grad_(True)
    block.set_x0(xg)
    z = block(xg, data)
    params = [block.odefunc.alpha_train, block.odefunc.beta_train]
    L = getattr(block.odefunc, 'multihead_att_layer', None)
    if L is not None:
        params = [L.Q.weight, L.Q.bias, L.K.weight, L.K.bias] + params
    # plain backprop leaves K.bias out of the graph (a per-row constant of the scores: the row softmax cancels it)
    grads = torch.autograd.grad((z * g).sum(), [xg] + params, allow_unused=True)
    return z.detach(), [torch.zeros_like(p) if t is None else t.detach() for t, p in zip(grads, [xg] + params)]


def rel(a, b, floor=0.0):
    """max abs error over max(|b|, floor): the bias gradients of a softmax layer are sums that cancel to rounding
    noise, they are judged on the scale of the weight gradients"""
    return float((a - b).abs().max() / b.abs().max().clamp_min(max(floor, 1e-30)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--nodes', type=int, default=20000, help='nodes of the shaped graph (average degree 12)')
    ap.add_argument('--rmat', type=int, default=14, help='also an R-MAT graph of this scale (0: skip)')
    ap.add_argument('--tol', type=float, default=2e-4)
    ap.add_argument('--emulate', action='store_true',
                    help='self-test of THIS SCRIPT without a GPU: gloo + the host-memory ABI emulator of tests/')
    args = ap.parse_args()
    rank, world, local = (int(os.environ[k]) for k in ('RANK', 'WORLD_SIZE', 'LOCAL_RANK'))
    if args.emulate:
        import contextlib
        sys.path.insert(0, os.path.join(ROOT, 'tests'))
        from abi_emulator import Emulator
        from test_host_partition_cpu import install_nl_route
        from grand_b200 import _lib
        emu = Emulator()
        _lib.load, _lib.require_cuda, _lib.stream = (lambda: emu), (lambda *a: None), (lambda: 0)
        torch.cuda.device = lambda dev: contextlib.nullcontext()
        install_nl_route(emu, setattr)
        device = torch.device('cpu')
        dist.init_process_group('gloo')
    else:
        torch.cuda.set_device(local)
        device = torch.device('cuda', local)
        dist.init_process_group('nccl', device_id=device)
    graphs = [('shaped', *_shaped(args.nodes, device))]
    if args.rmat:
        graphs.append((f'rmat{args.rmat}', synth.rmat_graph(args.rmat, 16, 77, device), 1 << args.rmat))
    bad = 0
    for gname, ei, n in graphs:
        data = grand_b200.GraphData()
        data.new_graph(ei, n)
        d = OPT['hidden_dim']
        gen = torch.Generator(device=device).manual_seed(3)               # same seed on every rank: same tensors
        x, gz = torch.randn(n, d, generator=gen, device=device), torch.randn(n, d, generator=gen, device=device)
        part = grand_b200.RowPartition(n, world, rank)
        for function, adjoint in (('transformer', True), ('laplacian', True), ('transformer', False)):
            os.environ['GRANDB200_OVERLAP'] = '0'
            z_ref, g_ref = train_pass(make_block(function, device, adjoint), data, x, gz)
            kinds = [('dense', grand_b200.Halo), ('sparse', grand_b200.SparseHalo), ('compact', grand_b200.CompactHalo)]
            if not args.emulate:                                   # symmetric memory needs real GPUs
                kinds.append(('push', grand_b200.PushHalo))
                kinds.append(('copy', grand_b200.CopyHalo))
                kinds.append(('cpush', grand_b200.CompactPushHalo))
            for kind, cls in kinds:
                for overlap in (('0', '1') if adjoint else ('0',)):
                    os.environ['GRANDB200_OVERLAP'] = overlap
                    halo = cls(part)
                    block = make_block(function, device, adjoint)
                    block.set_partition(part, halo)
                    z, grads = train_pass(block, data, part.to_slab(x), part.to_slab(gz))
                    errs = [rel(z, part.to_slab(z_ref)), rel(grads[0], part.to_slab(g_ref[0]), float(g_ref[0].abs().max()))]
                    scale = max(float(t.abs().max()) for t in g_ref[1:])
                    errs += [rel(a, b, 1e-2 * scale) for a, b in zip(grads[1:], g_ref[1:])]
                    worst = torch.tensor([max(errs)], device=device)
                    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
                    ok = float(worst) <= args.tol
                    bad += 0 if ok else 1
                    if rank == 0:
                        frac = getattr(halo, 'referenced_fraction', None)
                        print(f"{gname:8s} {function:11s} {'adjoint ' if adjoint else 'backprop'} halo={kind:7s} overlap={overlap} world={world} "
                              f"max rel err {float(worst):.2e} referenced={frac if frac is None else round(frac, 3)} "
                              f"{'ok' if ok else 'MISMATCH'}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)


def _shaped(n, device):
    ei = synth.shaped_graph(n, 12 * n // 2 * 2, 1234, device)
    return ei, n


if __name__ == '__main__':
    main()
