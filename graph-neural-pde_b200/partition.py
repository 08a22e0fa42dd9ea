"""1-D row partition of the diffusion operator over the GPUs of one node.

The reference has no model-parallel mode (its only multi-GPU path is nn.DataParallel replicas); graphs
too large for one GPU are split here by contiguous row slabs, one process per GPU:

    rank p owns nodes [p*S, min(N, (p+1)*S)),  S = ceil(N / P);  states / gradients live as [S, d] slabs
    (zero padded on the last rank), the plan of rank p holds the edges whose ROW it owns, with global
    column ids, over N_pad = P*S nodes.

One exchange per evaluation, none inside the kernels:
    forward   x_full  = all_gather(x_slab)                       (NCCL over NVLink, (P-1)/P * 4 N d bytes in)
              f_slab  = rows [lo, hi) of the fused kernels on (plan_p, x_full)
    backward  every rank forms its share of dx for ALL columns (transpose pass over its own edges; only the
              owner adds the "- g_j" self term), dx_slab = reduce_scatter(dx_share); parameter gradients
              (tiny) are all-reduced.  g is never gathered: a rank's edges only touch its own rows of g.

`RowPartition` and `Halo` are plain torch / torch.distributed and run on any backend (the world_size-2
gloo tests exercise them on CPU); the compute entry points below call the CUDA library and need a GPU.
"""
import os

import torch

from . import _lib


def _nvtx(name):
    """NVTX range around a halo exchange (GRANDB200_NVTX=1)"""
    def deco(fn):
        import functools

        @functools.wraps(fn)
        def wrapped(*a, **k):
            with _lib.nvtx_range(name):
                return fn(*a, **k)
        return wrapped
    return deco


class RowPartition:
    """contiguous equal-height row slabs of an n-node graph over `world` ranks"""

    def __init__(self, n, world, rank):
        if not (0 <= rank < world):
            raise ValueError(f'rank {rank} outside world of {world}')
        self.n, self.world, self.rank = int(n), int(world), int(rank)
        self.slab = -(-self.n // self.world)                 # S = ceil(N / P)
        self.n_pad = self.slab * self.world
        self.lo = min(self.rank * self.slab, self.n)
        self.hi = min(self.lo + self.slab, self.n)

    def bounds(self, rank):
        lo = min(rank * self.slab, self.n)
        return lo, min(lo + self.slab, self.n)

    def local_edges(self, edge_index, edge_weight=None):
        """edges whose row this rank owns (global ids kept), in their original relative order"""
        ei = edge_index[0] if edge_index.dim() == 3 else edge_index
        keep = (ei[0] >= self.lo) & (ei[0] < self.hi)
        ei_loc = ei[:, keep].contiguous()
        ew_loc = None
        if edge_weight is not None:
            ew_loc = edge_weight.reshape(-1)[keep].contiguous()
        return ei_loc, ew_loc

    def to_slab(self, full):
        """[N, d] (or [N_pad, d]) global matrix -> this rank's zero-padded [S, d] slab"""
        out = full.new_zeros((self.slab,) + tuple(full.shape[1:]))
        out[:self.hi - self.lo] = full[self.lo:self.hi]
        return out

    def __repr__(self):
        return f'RowPartition(n={self.n}, world={self.world}, rank={self.rank}, rows=[{self.lo}, {self.hi}))'


class Halo:
    """the collectives of the partitioned path over a torch.distributed process group"""

    def __init__(self, part, group=None):
        import torch.distributed as dist
        self.part, self.group, self.dist = part, group, dist
        if dist.get_world_size(group) != part.world or dist.get_rank(group) != part.rank:
            raise ValueError('process group does not match the partition')
        self.backend = dist.get_backend(group)
        self.bytes_in = 0                                     # bytes received so far (bench accounting)
        self.n_gathers = self.n_scatters = 0

    def prepare(self, plan):
        """hook for halos that need the graph (SparseHalo); the dense exchange needs nothing"""

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        """[S, d] slab of every rank -> [N_pad, d] on every rank.  keep=True: the caller will hold on to the result
        (saved for a backward pass) - halos that recycle their gather buffers hand out a private one then."""
        slab = slab.contiguous()
        full = slab.new_empty((self.part.n_pad,) + tuple(slab.shape[1:]))
        self.dist.all_gather_into_tensor(full, slab, group=self.group)
        self.bytes_in += (self.part.world - 1) * slab.numel() * slab.element_size()
        self.n_gathers += 1
        return full

    def all_gather_rows_async(self, slab):
        """start the all-gather of `slab` and return at once: a PendingGather whose wait() hands out the [N_pad, d]
        matrix.  On NCCL the collective runs on the communicator's own stream (ordered after the work already queued
        on the current stream), so kernels launched before wait() overlap it."""
        slab_c = slab.contiguous()
        full = slab_c.new_empty((self.part.n_pad,) + tuple(slab_c.shape[1:]))
        work = self.dist.all_gather_into_tensor(full, slab_c, group=self.group, async_op=True)
        self.bytes_in += (self.part.world - 1) * slab_c.numel() * slab_c.element_size()
        self.n_gathers += 1
        self.n_prefetched = getattr(self, 'n_prefetched', 0) + 1
        return PendingGather(slab, slab_c, full, work)

    @_nvtx('halo.reduce_scatter')
    def reduce_scatter_rows(self, full):
        """sum over ranks of [N_pad, d], rank p keeps rows [p*S, (p+1)*S)"""
        full = full.contiguous()
        out = full.new_empty((self.part.slab,) + tuple(full.shape[1:]))
        if self.backend == 'gloo':                            # gloo has no reduce_scatter: all_reduce + slice
            self.dist.all_reduce(full, group=self.group)
            out.copy_(full[self.part.rank * self.part.slab:(self.part.rank + 1) * self.part.slab])
        else:
            self.dist.reduce_scatter_tensor(out, full, group=self.group)
        self.bytes_in += (self.part.world - 1) * out.numel() * out.element_size()
        self.n_scatters += 1
        return out

    def reduce_scatter_rows_async(self, full):
        """start the reduce-scatter and return a pending handle (wait() -> this rank's [S, d] rows of the sum)"""
        full = full.contiguous()
        out = full.new_empty((self.part.slab,) + tuple(full.shape[1:]))
        self.bytes_in += (self.part.world - 1) * out.numel() * out.element_size()
        self.n_scatters += 1
        if self.backend == 'gloo':                            # gloo has no reduce_scatter: all_reduce + slice on wait()
            work = self.dist.all_reduce(full, group=self.group, async_op=True)
            lo = self.part.rank * self.part.slab
            return PendingScatter(full, work, lambda: out.copy_(full[lo:lo + self.part.slab]))
        work = self.dist.reduce_scatter_tensor(out, full, group=self.group, async_op=True)
        return PendingScatter(full, work, lambda: out)

    def all_reduce_(self, t):
        self.dist.all_reduce(t, group=self.group)
        return t


class SparseHalo(Halo):
    """Halo that exchanges only the REFERENCED rows (SURVEY.md 8e "refinement").

    Rank p's kernels read x[j] only for the columns j its own edges point at, and its share of dL/dx is zero on every
    other column it does not own.  On graphs whose slabs reference a small part of the other slabs (R-MAT: ~25 % at
    P = 8, 38 % of the vertices are isolated) the all-gather / reduce-scatter of whole [N_pad, d] matrices moves
    mostly rows nobody reads.  Here every pair of ranks exchanges index lists once per graph (`prepare`), and then

        gather   send x_slab[rows peer q needs]  ->  all_to_all_v  ->  scatter into a zero [N_pad, d] matrix
        scatter  send share[columns of peer q I touched]  ->  all_to_all_v  ->  out = share[own rows] + sum over peers,
                 added peer after peer in rank order (indices unique per peer: deterministic, no float atomics)

    Rows that are neither owned nor referenced stay exactly zero in the gathered matrix (the column pass and the
    GEMMs run over all N_pad rows, but multiply them with zero coefficients)."""

    def __init__(self, part, group=None):
        super().__init__(part, group)
        self._key = None

    def prepare(self, plan):
        """build the per-peer index lists for the graph of `plan` (this rank's edges, global column ids)"""
        if self._key is plan:
            return
        part = self.part
        cols = torch.unique(plan.colidx.long())                       # sorted
        owner = torch.div(cols, part.slab, rounding_mode='floor')
        self._exchange_lists(cols[owner != part.rank], owner[owner != part.rank], plan.colidx.device)
        self.__dict__.pop('_ring', None)                              # other rows are referenced now: fresh zero buffers
        self._key = plan

    def _exchange_lists(self, remote, owner, dev):
        """remote: sorted global ids of the rows of other slabs this rank's edges reference, owner: their ranks.
        Every pair of ranks swaps counts, then ids (two all_to_all calls, once per graph)."""
        part, dist = self.part, self.dist
        need_counts = torch.bincount(owner, minlength=part.world)      # rows I need from each peer
        give_counts = torch.empty_like(need_counts)
        dist.all_to_all_single(give_counts, need_counts, group=self.group)
        self.recv_splits = [int(v) for v in need_counts.tolist()]
        self.send_splits = [int(v) for v in give_counts.tolist()]
        wanted = torch.empty(sum(self.send_splits), dtype=torch.int64, device=dev)
        dist.all_to_all_single(wanted, remote.contiguous(), self.send_splits, self.recv_splits, group=self.group)
        self.recv_idx = remote.contiguous()                            # global positions of the rows I receive (peer order)
        self.send_idx = (wanted - part.rank * part.slab).contiguous()  # my local rows each peer wants (peer order)
        if self.send_idx.numel() and (int(self.send_idx.min()) < 0 or int(self.send_idx.max()) >= part.slab):
            raise RuntimeError('SparseHalo: a peer asked for rows outside this slab')
        self.referenced_fraction = float(remote.numel()) / max(1, part.n_pad - part.slab)

    def _check(self):
        if self._key is None:
            raise RuntimeError('SparseHalo.prepare(plan) must run before the first exchange')

    RING = 3       # gathered matrices alive at once under no_grad: the one in use, the prefetched one, one spare

    def _gather_buffer(self, slab, keep=False):
        """[N_pad, d] matrix whose unreferenced rows are zero.  Under no_grad (forward integration, adjoint solve) the
        result is consumed within the evaluation, so a small ring of persistent buffers is reused: only the owned and
        referenced rows are rewritten, the rest was zeroed once (no 4 N d byte memset per exchange).  With autograd
        recording, the gathered matrix is saved for backward and must not be overwritten: a fresh one each time."""
        shape = (self.part.n_pad,) + tuple(slab.shape[1:])
        if keep or torch.is_grad_enabled():
            return slab.new_zeros(shape)
        key = (shape, slab.dtype, slab.device)
        ring = self.__dict__.setdefault('_ring', {})
        bufs, nxt = ring.get(key, ([], 0))
        if len(bufs) < self.RING:
            bufs.append(slab.new_zeros(shape))
            buf = bufs[-1]
        else:
            buf = bufs[nxt % self.RING]
        ring[key] = (bufs, nxt + 1)
        return buf

    def _finish_gather(self, slab, recv, keep=False):
        part = self.part
        full = self._gather_buffer(slab, keep)
        full[part.rank * part.slab:(part.rank + 1) * part.slab] = slab
        if recv.shape[0]:
            full.index_copy_(0, self.recv_idx, recv)
        return full

    def _start_gather(self, slab, async_op):
        self._check()
        slab = slab.contiguous()
        send = slab.index_select(0, self.send_idx)
        recv = slab.new_empty((sum(self.recv_splits),) + tuple(slab.shape[1:]))
        work = self.dist.all_to_all_single(recv, send, self.recv_splits, self.send_splits, group=self.group,
                                           async_op=async_op)
        self.bytes_in += recv.numel() * recv.element_size()
        self.n_gathers += 1
        return slab, send, recv, work

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        slab, _send, recv, _ = self._start_gather(slab, False)
        return self._finish_gather(slab, recv, keep)

    def all_gather_rows_async(self, slab):
        slab_c, send, recv, work = self._start_gather(slab, True)
        self.n_prefetched = getattr(self, 'n_prefetched', 0) + 1
        return PendingExchange(slab, work, lambda: self._finish_gather(slab_c, recv), keep=(send,))

    def _finish_scatter(self, full, recv):
        part = self.part
        out = full[part.rank * part.slab:(part.rank + 1) * part.slab].clone()
        off = 0
        for cnt in self.send_splits:                                   # peer after peer, rank order: a fixed order of adds
            if cnt:
                out.index_add_(0, self.send_idx[off:off + cnt], recv[off:off + cnt])
            off += cnt
        return out

    def _start_scatter(self, full, async_op):
        self._check()
        full = full.contiguous()
        send = full.index_select(0, self.recv_idx)                    # my share for the columns the peers own
        recv = full.new_empty((sum(self.send_splits),) + tuple(full.shape[1:]))
        work = self.dist.all_to_all_single(recv, send, self.send_splits, self.recv_splits, group=self.group,
                                           async_op=async_op)
        self.bytes_in += recv.numel() * recv.element_size()
        self.n_scatters += 1
        return full, send, recv, work

    @_nvtx('halo.reduce_scatter')
    def reduce_scatter_rows(self, full):
        full, _send, recv, _ = self._start_scatter(full, False)
        return self._finish_scatter(full, recv)

    def reduce_scatter_rows_async(self, full):
        full, send, recv, work = self._start_scatter(full, True)
        return PendingExchange(None, work, lambda: self._finish_scatter(full, recv), keep=(send,))


class LocalRows:
    """the partition as ONE rank's kernels see it under a CompactHalo: a graph of n_local = S + R nodes whose first S
    ids are this rank's slab (rows [0, hi - lo) real, the rest padding) and whose last R ids are the remote rows its
    edges reference, in (owner, global id) order.  Same attribute names as RowPartition, so the compute entry points
    below take either."""

    def __init__(self, part, n_local):
        self.n, self.world, self.rank, self.slab = part.n, part.world, part.rank, part.slab
        self.lo, self.hi = 0, part.hi - part.lo
        self.n_pad = int(n_local)

    def __repr__(self):
        return f'LocalRows(rank={self.rank}, own={self.hi}, slab={self.slab}, n_local={self.n_pad})'


class CompactHalo(SparseHalo):
    """SparseHalo whose local problem is COMPACT: besides exchanging only the referenced rows, the rank's graph is
    relabelled (`localize`) to [own slab | referenced remote rows], so every per-rank buffer and every kernel that
    runs over "all columns" (column pass, [dq|dk] GEMMs, zero fill of the share) shrinks from N_pad rows to
    S + R rows - per-rank work is O(N/P + halo + E/P) instead of O(N + E/P).  The exchanges become copy-free on the
    receiving (gather) and sending (scatter) side: the all_to_all_v receives straight into the tail of the local
    state matrix and sends straight from the tail of the local share.

    Use: block.set_partition(part, CompactHalo(part)); ODEblock.reset_graph_data then relabels the rank's edges and
    hands (halo.view, halo) to the ODE function."""

    def __init__(self, part, group=None):
        super().__init__(part, group)
        self.view = None

    def localize(self, edge_index):
        """edges of this rank with GLOBAL ids [2, E_p] -> the same edges, same order, with LOCAL ids; builds the
        per-peer index lists on the way (collective: every rank of the group must call it)"""
        part = self.part
        ei = edge_index[0] if edge_index.dim() == 3 else edge_index
        rows, cols = ei[0].long(), ei[1].long()
        if rows.numel() and (int(rows.min()) < part.lo or int(rows.max()) >= part.hi):
            raise ValueError('CompactHalo.localize takes the edges whose row this rank owns (RowPartition.local_edges)')
        uniq = torch.unique(cols)                                     # sorted: grouped by owner, ascending
        owner = torch.div(uniq, part.slab, rounding_mode='floor')
        remote = uniq[owner != part.rank].contiguous()
        self._exchange_lists(remote, owner[owner != part.rank], ei.device)
        mine = (cols >= part.rank * part.slab) & (cols < (part.rank + 1) * part.slab)
        pos = torch.searchsorted(remote, cols)                        # only read where the column is remote
        local_cols = torch.where(mine, cols - part.rank * part.slab, part.slab + pos)
        self.view = LocalRows(part, part.slab + remote.numel())
        self._key = self.view
        return torch.stack([rows - part.lo, local_cols]).to(ei.dtype)

    def prepare(self, plan):
        if self.view is None:
            raise RuntimeError('CompactHalo: localize(edge_index) must run before the first exchange '
                               '(ODEblock.reset_graph_data does it)')
        if plan.n != self.view.n_pad:
            raise ValueError(f'CompactHalo: plan over {plan.n} nodes, local problem has {self.view.n_pad}')

    def _start_gather(self, slab, async_op):
        self._check()
        slab = slab.contiguous()
        s = self.part.slab
        local = slab.new_empty((self.view.n_pad,) + tuple(slab.shape[1:]))
        local[:s] = slab
        send = slab.index_select(0, self.send_idx)
        recv = local[s:]                                              # contiguous tail: received rows land in place
        work = self.dist.all_to_all_single(recv, send, self.recv_splits, self.send_splits, group=self.group,
                                           async_op=async_op)
        self.bytes_in += recv.numel() * recv.element_size()
        self.n_gathers += 1
        return local, send, recv, work

    def _finish_gather(self, local, recv, keep=False):
        return local                                                   # a fresh matrix every time: nothing is recycled

    def _start_scatter(self, full, async_op):
        self._check()
        full = full.contiguous()
        send = full[self.part.slab:]                                  # my share for the peers' rows, already in peer order
        recv = full.new_empty((sum(self.send_splits),) + tuple(full.shape[1:]))
        work = self.dist.all_to_all_single(recv, send, self.send_splits, self.recv_splits, group=self.group,
                                           async_op=async_op)
        self.bytes_in += recv.numel() * recv.element_size()
        self.n_scatters += 1
        return full, send, recv, work

    def _finish_scatter(self, full, recv):
        out = full[:self.part.slab].clone()
        off = 0
        for cnt in self.send_splits:                                   # peer after peer, rank order: a fixed order of adds
            if cnt:
                out.index_add_(0, self.send_idx[off:off + cnt], recv[off:off + cnt])
            off += cnt
        return out


class PushHalo(Halo):
    """Halo whose all-gather is a PUSH over NVLink peer memory instead of a NCCL collective (DESIGN.md "Multi-GPU").

    The gathered [N_pad, d] matrices live in symmetric memory (torch.distributed._symmetric_memory: every rank holds
    a pointer to every peer's copy).  A gather is ONE kernel on the producing rank - gb_stage_combine_multi with no
    stage terms = a P-way broadcast store of the rank's slab into its row range of every copy (posted writes: the
    kernel runs at local HBM speed while the copies drain over NVLink) - followed by a symmetric-memory barrier that
    orders every rank's reads after every rank's writes.  A ring of RING matrices per state width: a rank can only
    start writing matrix b again after it has passed a later barrier, which every peer reaches only after its reads
    of b were queued, so nobody overwrites what somebody still reads.

    The reduce-scatter of dL/dx is a PULL over the same mechanism: every rank forms its share for all columns in a
    symmetric buffer (`share_buffer`, handed to the VJP kernels as their output), a barrier, then ONE
    gb_stage_combine launch per rank sums its own slab rows of the P shares - own share first, then the peers' in
    rank order, read straight from their memory: deterministic, no NCCL on the data path.  Parameter gradients
    (a few KB) still go through all_reduce.

    Verified over NVLink on 2 B200s against the 1-GPU result (scripts/check_partition_nccl.py); opt-in,
    `bench.py --halo push`: within 3 % of the NCCL exchange at 2 - 8 GPUs."""

    RING = 3

    def __init__(self, part, group=None):
        super().__init__(part, group)
        self._rings = {}

    def _symmetric(self, rows, width, dtype, device):
        """(local tensor [rows, width], peer base pointers [world], barrier()) - one collective per allocation"""
        import warnings
        import torch.distributed._symmetric_memory as symm
        group = self.group if self.group is not None else self.dist.group.WORLD
        try:                                                   # older torch wants the group registered first
            with warnings.catch_warnings():
                warnings.simplefilter('ignore')
                if not symm.is_symm_mem_enabled_for_group(group.group_name):
                    symm.enable_symm_mem_for_group(group.group_name)
        except (AttributeError, RuntimeError):
            pass
        t = symm.empty(rows, width, dtype=dtype, device=device)
        handle = symm.rendezvous(t, group)
        return t, [int(p) for p in handle.buffer_ptrs], handle.barrier

    def _next(self, width, dtype, device):
        key = (int(width), dtype, str(device))
        ring = self._rings.get(key)
        if ring is None:
            ring = self._rings[key] = [[self._symmetric(self.part.n_pad, width, dtype, device)
                                        for _ in range(self.RING)], 0]
        entry = ring[0][ring[1] % self.RING]
        ring[1] += 1
        return entry

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        import ctypes as C
        slab = slab.contiguous()
        if slab.dim() != 2 or slab.dtype != torch.float32:
            raise TypeError('PushHalo exchanges fp32 [S, d] slabs')
        part = self.part
        full, ptrs, barrier = self._next(slab.shape[1], slab.dtype, slab.device)
        offset = 4 * part.rank * part.slab * slab.shape[1]                 # my rows inside every copy
        outs = (C.c_void_p * part.world)(*[p + offset for p in ptrs])
        lib = _lib.load()
        _lib.check(lib.gb_stage_combine_multi(slab.data_ptr(), None, None, 0, slab.numel(), outs, part.world,
                                              _lib.stream()), 'gb_stage_combine_multi')
        barrier()                                                          # all writes land before any rank reads
        self.bytes_in += (part.world - 1) * slab.numel() * slab.element_size()
        self.n_gathers += 1
        # a caller that keeps the gathered matrix (saved for backward) gets a private copy, else the ring buffer
        return full.clone() if (keep or torch.is_grad_enabled()) else full

    def all_gather_rows_async(self, slab):
        """the push and its barrier issued on a side stream (ordered after the work already queued on the current
        one), so that kernels launched before wait() overlap the barrier's wait for the slowest peer"""
        if not slab.is_cuda:
            full = self.all_gather_rows(slab)
            self.n_prefetched = getattr(self, 'n_prefetched', 0) + 1
            return PendingGather(slab, slab, full, None)
        side = self.__dict__.get('_side')
        if side is None:
            side = self._side = torch.cuda.Stream(device=slab.device)
        main = torch.cuda.current_stream(slab.device)
        side.wait_stream(main)
        slab_c = slab.contiguous()
        with torch.cuda.stream(side):
            full = self.all_gather_rows(slab_c)
        slab_c.record_stream(side)
        full.record_stream(side)
        self.n_prefetched = getattr(self, 'n_prefetched', 0) + 1
        return _StreamPending(slab, full, side)

    def share_buffer(self, rows, width, device):
        """next [rows, width] fp32 matrix of the ring the peers can read (nl_local_vjp & co. write dL/dx into it)"""
        key = ('share', int(rows), int(width), str(device))
        ring = self._rings.get(key)
        if ring is None:
            ring = self._rings[key] = [[self._symmetric(rows, width, torch.float32, device)
                                        for _ in range(self.RING)], 0]
        entry = ring[0][ring[1] % self.RING]
        ring[1] += 1
        self.__dict__.setdefault('_shares', {})[id(entry[0])] = entry
        return entry[0]

    @_nvtx('halo.reduce_scatter')
    def reduce_scatter_rows(self, full):
        import ctypes as C
        entry = self.__dict__.get('_shares', {}).get(id(full))
        if entry is None or entry[0] is not full:                          # not formed in a share buffer: NCCL
            return super().reduce_scatter_rows(full)
        part = self.part
        _t, ptrs, barrier = entry
        width = full.shape[1]
        if full.shape[0] != part.n_pad:
            raise ValueError('PushHalo.reduce_scatter_rows: share must cover all N_pad rows')
        barrier()                                                          # every rank's share is complete
        offset = 4 * part.rank * part.slab * width
        peers = [ptrs[p] + offset for p in range(part.world) if p != part.rank]
        out = full.new_empty((part.slab, width))
        ks = (C.c_void_p * max(1, len(peers)))(*peers)
        cs = (C.c_float * max(1, len(peers)))(*([1.0] * len(peers)))
        lib = _lib.load()
        _lib.check(lib.gb_stage_combine(ptrs[part.rank] + offset, ks, cs, len(peers), part.slab * width,
                                        out.data_ptr(), _lib.stream()), 'gb_stage_combine')
        self.bytes_in += (part.world - 1) * out.numel() * out.element_size()
        self.n_scatters += 1
        return out

    def reduce_scatter_rows_async(self, full):
        out = self.reduce_scatter_rows(full)
        return PendingScatter(full, None, lambda: out)


class CompactPushHalo(CompactHalo):
    """CompactHalo whose two exchanges are PUSHES over NVLink peer memory instead of NCCL all_to_all_v calls
    (293 GB/s into a rank at 8 GPUs on R-MAT 24: more than half of a GRAND-l step).

    gather   the local state matrices [S + R, d] live in symmetric memory.  ONE kernel per rank (gb_rows_push) picks the
             rows each peer asked for out of the rank's slab and stores them straight into the peer's tail, at the
             offset where that peer expects this rank's block (no packing buffer, no collective); a symmetric-memory
             barrier orders every rank's reads after every rank's writes.
    scatter  the tail of a rank's share (its contributions to rows the peers own, already in peer order) is pushed
             block by block into the owners' symmetric staging matrices (the same kernel without an index list);
             barrier; the owner adds the blocks to its own rows peer after peer in rank order (unique indices per
             peer: deterministic), exactly as CompactHalo does with what all_to_all_v delivered.
    Ring of RING buffers per kind and width, reused as in PushHalo (a rank writes a peer's buffer b again only after
    passing two more barriers, which that peer reaches after its reads of b were queued)."""

    RING = 3

    def __init__(self, part, group=None):
        super().__init__(part, group)
        self._sym = {}

    def _exchange_lists(self, remote, owner, dev):
        super()._exchange_lists(remote, owner, dev)
        part, dist = self.part, self.dist
        # need[q][r] = rows rank q needs from rank r, on every rank: where my block starts inside each peer's buffers
        mine = torch.tensor(self.recv_splits, dtype=torch.int64, device=dev)
        need = [torch.empty_like(mine) for _ in range(part.world)]
        dist.all_gather(need, mine, group=self.group)
        need = torch.stack(need).cpu()                                      # [q, r]
        me = part.rank
        self._tail_off = [int(need[q, :me].sum()) for q in range(part.world)]       # my block inside q's tail
        self._stage_off = [int(need[:me, q].sum()) for q in range(part.world)]      # my block inside q's staging rows
        sizes = torch.tensor([part.slab + int(need[me].sum()), int(need[:, me].sum())], dtype=torch.int64, device=dev)
        dist.all_reduce(sizes, op=dist.ReduceOp.MAX, group=self.group)
        self._rows_local, self._rows_stage = max(1, int(sizes[0])), max(1, int(sizes[1]))
        self._sym.clear()

    def _symmetric(self, kind, width, dtype, device):
        import warnings
        import torch.distributed._symmetric_memory as symm
        key = (kind, int(width), dtype, str(device))
        ring = self._sym.get(key)
        if ring is None:
            group = self.group if self.group is not None else self.dist.group.WORLD
            try:
                with warnings.catch_warnings():
                    warnings.simplefilter('ignore')
                    if not symm.is_symm_mem_enabled_for_group(group.group_name):
                        symm.enable_symm_mem_for_group(group.group_name)
            except (AttributeError, RuntimeError):
                pass
            rows = self._rows_local if kind == 'local' else self._rows_stage
            entries = []
            for _ in range(self.RING):
                t = symm.empty(rows, width, dtype=dtype, device=device)
                handle = symm.rendezvous(t, group)
                entries.append((t, [int(p) for p in handle.buffer_ptrs], handle.barrier))
            ring = self._sym[key] = [entries, 0]
        entry = ring[0][ring[1] % self.RING]
        ring[1] += 1
        return entry

    def _push(self, src, idx, counts, ptrs, row_off, width):
        """rows of `src` (picked by idx, or consecutive) -> peer q's matrix at row row_off[q], q in rank order"""
        import ctypes as C
        part = self.part
        peers = [q for q in range(part.world) if q != part.rank and counts[q] > 0]
        if not peers:
            return
        # the request list is ordered by peer; the kernel takes at most 8 segments per launch
        starts = [0]
        for q in range(part.world):
            starts.append(starts[-1] + counts[q])
        lib = _lib.load()
        for i in range(0, len(peers), 8):
            grp = peers[i:i + 8]
            base = starts[grp[0]]
            seg = [starts[q] - base for q in grp] + [starts[grp[-1] + 1] - base]
            # peers of one launch must be consecutive in the list: split where a skipped (empty / own) block intervenes
            ok = all(starts[grp[j] + 1] == starts[grp[j + 1]] for j in range(len(grp) - 1))
            if not ok:
                for q in grp:
                    self._push_one(lib, src, idx, starts[q], counts[q], ptrs[q] + 4 * row_off[q] * width, width)
                continue
            dst = (C.c_void_p * len(grp))(*[ptrs[q] + 4 * row_off[q] * width for q in grp])
            segs = (C.c_int64 * len(seg))(*seg)
            idx_ptr = None if idx is None else idx.data_ptr() + 8 * base
            src_ptr = src.data_ptr() if idx is not None else src.data_ptr() + 4 * base * width
            _lib.check(lib.gb_rows_push(src_ptr, width, idx_ptr, len(grp), segs, dst, width, width, _lib.stream()),
                       'gb_rows_push')

    def _push_one(self, lib, src, idx, start, cnt, dst_ptr, width):
        import ctypes as C
        dst = (C.c_void_p * 1)(dst_ptr)
        segs = (C.c_int64 * 2)(0, cnt)
        idx_ptr = None if idx is None else idx.data_ptr() + 8 * start
        src_ptr = src.data_ptr() if idx is not None else src.data_ptr() + 4 * start * width
        _lib.check(lib.gb_rows_push(src_ptr, width, idx_ptr, 1, segs, dst, width, width, _lib.stream()), 'gb_rows_push')

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        self._check()
        slab = slab.contiguous()
        if slab.dim() != 2 or slab.dtype != torch.float32 or slab.shape[1] % 4:
            raise TypeError('CompactPushHalo exchanges fp32 [S, d] slabs with d % 4 == 0')
        part, s, width = self.part, self.part.slab, slab.shape[1]
        buf, ptrs, barrier = self._symmetric('local', width, slab.dtype, slab.device)
        buf[:s] = slab
        self._push(slab, self.send_idx, self.send_splits, ptrs, [s + o for o in self._tail_off], width)
        barrier()                                                          # all writes landed before any rank reads
        self.bytes_in += sum(self.recv_splits) * width * slab.element_size()
        self.n_gathers += 1
        local = buf[:self.view.n_pad]
        return local.clone() if (keep or torch.is_grad_enabled()) else local

    def all_gather_rows_async(self, slab):
        side = self.__dict__.get('_side')
        if side is None:
            side = self._side = torch.cuda.Stream(device=slab.device)
        main = torch.cuda.current_stream(slab.device)
        side.wait_stream(main)
        slab_c = slab.contiguous()
        with torch.cuda.stream(side):
            local = self.all_gather_rows(slab_c)
        slab_c.record_stream(side)
        local.record_stream(side)
        self.n_prefetched = getattr(self, 'n_prefetched', 0) + 1
        return _StreamPending(slab, local, side)

    @_nvtx('halo.reduce_scatter')
    def reduce_scatter_rows(self, full):
        self._check()
        full = full.contiguous()
        part, s, width = self.part, self.part.slab, full.shape[1]
        if full.dtype != torch.float32 or width % 4:
            raise TypeError('CompactPushHalo exchanges fp32 rows with d % 4 == 0')
        stage, ptrs, barrier = self._symmetric('stage', width, full.dtype, full.device)
        tail = full[s:]                                                    # contributions to the peers' rows, peer order
        self._push(tail, None, self.recv_splits, ptrs, self._stage_off, width)
        barrier()
        self.bytes_in += sum(self.send_splits) * width * full.element_size()
        self.n_scatters += 1
        return self._finish_scatter(full, stage)

    def reduce_scatter_rows_async(self, full):
        out = self.reduce_scatter_rows(full)
        return PendingScatter(full, None, lambda: out)


class CopyHalo(PushHalo):
    """PushHalo whose all-gather moves the bytes with the COPY ENGINES instead of a kernel: every rank writes its slab
    into its own symmetric matrix and, with P - 1 peer-to-peer copies (cudaMemcpyAsync over NVLink, one stream per
    peer so the engines run side by side), into its row range of every peer's matrix; then a symmetric-memory barrier
    (GRANDB200_COPY_HALO=pull: barrier first, then every rank reads the peers' slabs instead).  No SM is involved apart from the barrier's few threads, so a gather prefetched on a side stream
    (GRANDB200_OVERLAP=1) does not take SMs or issue slots from the kernels it overlaps - what limited the overlapped
    NCCL / push exchanges at 8 GPUs (DESIGN.md "(e)").  Buffer reuse: a rank overwrites its slab in ring entry b three
    gathers later, after passing two more barriers, which every peer reaches only after its pulls from b were
    complete in stream order.  The reduce-scatter is PushHalo's pull kernel."""

    def _symmetric(self, rows, width, dtype, device):
        import warnings
        import torch.distributed._symmetric_memory as symm
        group = self.group if self.group is not None else self.dist.group.WORLD
        try:
            with warnings.catch_warnings():
                warnings.simplefilter('ignore')
                if not symm.is_symm_mem_enabled_for_group(group.group_name):
                    symm.enable_symm_mem_for_group(group.group_name)
        except (AttributeError, RuntimeError):
            pass
        t = symm.empty(rows, width, dtype=dtype, device=device)
        handle = symm.rendezvous(t, group)
        views = [t if p == self.part.rank else handle.get_buffer(p, (rows, width), dtype)
                 for p in range(self.part.world)]
        self.__dict__.setdefault('_views', {})[id(t)] = views
        return t, [int(p) for p in handle.buffer_ptrs], handle.barrier

    def _peer_streams(self, device):
        st = self.__dict__.get('_pstreams')
        if st is None:
            st = self._pstreams = [torch.cuda.Stream(device=device) for _ in range(max(1, self.part.world - 1))]
        return st

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        slab = slab.contiguous()
        if slab.dim() != 2 or slab.dtype != torch.float32:
            raise TypeError('CopyHalo exchanges fp32 [S, d] slabs')
        part = self.part
        full, _ptrs, barrier = self._next(slab.shape[1], slab.dtype, slab.device)
        views = self._views[id(full)]
        s = part.slab
        mine = slice(part.rank * s, (part.rank + 1) * s)
        cur = torch.cuda.current_stream(slab.device)
        streams = self._peer_streams(slab.device)
        pull = os.environ.get('GRANDB200_COPY_HALO', 'push') == 'pull'
        full[mine].copy_(slab)
        if pull:
            barrier()                                                      # every rank's own slab is in its matrix
        i = 0
        for p in range(part.world):
            if p == part.rank:
                continue
            st = streams[i % len(streams)]
            i += 1
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                if pull:                                                   # remote reads: measured 2.3x slower at P = 8
                    full[p * s:(p + 1) * s].copy_(views[p][p * s:(p + 1) * s], non_blocking=True)
                else:                                                      # posted writes into the peer's matrix
                    views[p][mine].copy_(slab, non_blocking=True)
        for st in streams[:i]:
            cur.wait_stream(st)
        if not pull:
            barrier()                                                      # all writes landed before any rank reads
        self.bytes_in += (part.world - 1) * slab.numel() * slab.element_size()
        self.n_gathers += 1
        return full.clone() if (keep or torch.is_grad_enabled()) else full


class _StreamPending:
    """work queued on a side stream; wait() makes the current stream wait for it (no host block)"""

    def __init__(self, src, result, stream):
        self.src, self._result, self._stream = src, result, stream

    def wait(self):
        if self._stream is not None:
            torch.cuda.current_stream(self._result.device).wait_stream(self._stream)
            self._stream = None
        return self._result


class PendingExchange:
    """an all_to_all_v in flight; wait() completes it and runs the local unpacking step"""

    def __init__(self, src, work, finish, keep=()):
        self.src, self._work, self._finish, self._keep = src, work, finish, keep

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
        out = self._finish()
        self._keep = ()
        return out


class PendingGather:
    """an all-gather in flight (Halo.all_gather_rows_async).  `src` is the tensor the caller asked to gather: the
    consumer checks identity before trusting the result."""

    def __init__(self, src, slab, full, work):
        self.src, self._slab, self._full, self._work = src, slab, full, work

    def wait(self):
        if self._work is not None:
            self._work.wait()              # NCCL: the current stream waits for the collective (no host block)
            self._work = None
        return self._full


class PendingScatter:
    """a reduce-scatter in flight (Halo.reduce_scatter_rows_async); keeps the send buffer alive until wait()"""

    def __init__(self, full, work, result):
        self._full, self._work, self._result = full, work, result

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
        out = self._result()
        self._full = None
        return out


class LoopbackHalo:
    """world == 1 stand-in (no process group needed): the collectives are identities"""

    def __init__(self, part):
        if part.world != 1:
            raise ValueError('LoopbackHalo is for a world of 1')
        self.part, self.bytes_in = part, 0

    def prepare(self, plan):
        pass

    @_nvtx('halo.all_gather')
    def all_gather_rows(self, slab, keep=False):
        return slab.contiguous()

    def all_gather_rows_async(self, slab):
        c = slab.contiguous()
        return PendingGather(slab, c, c, None)

    @_nvtx('halo.reduce_scatter')
    def reduce_scatter_rows(self, full):
        return full.contiguous()

    def reduce_scatter_rows_async(self, full):
        c = full.contiguous()
        return PendingScatter(c, None, lambda: c)

    def all_reduce_(self, t):
        return t


def _virt(t, lo, ld):
    """device address such that row `lo` of a GLOBAL-row-indexed matrix lands on row 0 of the slab tensor
    (the kernels only touch owned rows of row-indexed matrices)"""
    return None if t is None else t.data_ptr() - 4 * lo * ld


# ---- local compute of one rank (CUDA) ---------------------------------------------------------------
def nl_local_forward(part, plan, x_full, x_slab, wq, bq, wk, alpha_train, beta_train, x0_slab, heads, alpha_sigmoid):
    """rows [lo, hi) of one GRAND-nl evaluation given the gathered state: returns (f_slab [S,d], q_slab [S,A])"""
    from .ops import _nl_q, _nl_workspace
    lib = _lib.load()
    s, d = x_slab.shape
    a = wq.shape[0]
    q = _nl_q(x_slab, wq, bq)
    out = torch.zeros_like(x_slab) if part.hi - part.lo < s else torch.empty_like(x_slab)
    source = beta_train is not None and x0_slab is not None
    _lib.check(lib.gb_nl_fwd(plan.c, part.lo, part.hi, x_full.data_ptr(), d, _virt(q, part.lo, a), a, wk.data_ptr(), d,
                             heads, a // heads, d, _virt(x0_slab, part.lo, d) if source else None, d,
                             alpha_train.data_ptr(), int(alpha_sigmoid), beta_train.data_ptr() if source else None,
                             _virt(out, part.lo, d), d, None, _nl_workspace(x_slab.device).data_ptr(), _lib.stream()),
               'gb_nl_fwd')
    return out, q


def share_buffer(halo, like):
    """where a rank's share of dL/dx for all columns is formed: a buffer the halo can exchange in place (PushHalo:
    symmetric memory the peers pull from) or None (an ordinary allocation)"""
    make = getattr(halo, 'share_buffer', None)
    return None if make is None else make(like.shape[0], like.shape[1], like.device)


def nl_local_vjp(part, plan, x_full, q, wq, wk, alpha_train, beta_train, x0_slab, heads, alpha_sigmoid, g_slab,
                 want_f, want_w, dx_out=None):
    """this rank's share of the VJP.  Returns (f_slab|None, dx_share [N_pad,d] (to be reduce-scattered; already
    holds dq Wq for the owned rows and dk Wk for every column), grads: flat share of
    (dWq, dWk, dbq, dbk, dalpha, dbeta) [2A(d+1) + 2], or of (dalpha, dbeta) [2] when want_w is False) -
    sums over ranks give the single-GPU numbers."""
    from .ops import _nl_workspace
    lib = _lib.load()
    n_pad, d = x_full.shape
    s = part.slab
    a = wq.shape[0]
    dev = x_full.device
    f32 = dict(dtype=torch.float32, device=dev)
    ds = torch.empty(plan.e, heads, **f32)
    dabar = torch.empty(plan.e, **f32)
    abar = torch.empty(plan.e, **f32)
    # [dq | dk] for ALL rows in one buffer: dq is non-zero on the owned rows only, dk is this rank's share for
    # every column - so ONE tensor-core GEMM each gives dx += [dq|dk] [Wq;Wk] and d[Wq;Wk] = [dq|dk]^T x
    dqk = torch.empty(n_pad, 2 * a, **f32)
    dqk[:part.lo, :a].zero_()
    dqk[part.hi:, :a].zero_()
    dx = torch.empty(n_pad, d, **f32) if dx_out is None else dx_out
    out = None
    if want_f:
        out = torch.zeros(s, d, **f32) if part.hi - part.lo < s else torch.empty(s, d, **f32)
    partials = torch.empty(int(lib.gb_att_bwd_fused_partials(plan.c)), **f32)
    grads = torch.empty(2 * a * d + 2 * a + 2 if want_w else 2, **f32)
    dparams = grads[-2:]
    source = beta_train is not None and x0_slab is not None
    _lib.check(lib.gb_nl_adjoint(plan.c, part.lo, part.hi, x_full.data_ptr(), d, _virt(q, part.lo, a), a,
                                 wk.data_ptr(), d, heads, a // heads, d,
                                 _virt(x0_slab, part.lo, d) if source else None, d, alpha_train.data_ptr(),
                                 int(alpha_sigmoid), beta_train.data_ptr() if source else None,
                                 _virt(g_slab, part.lo, d), d, _virt(out, part.lo, d), d, ds.data_ptr(),
                                 dabar.data_ptr(), abar.data_ptr(), dqk.data_ptr(), 2 * a, dqk.data_ptr() + 4 * a,
                                 2 * a, dx.data_ptr(), d, partials.data_ptr(), dparams.data_ptr(),
                                 _nl_workspace(dev).data_ptr(), _lib.stream()), 'gb_nl_adjoint')
    w_cat = torch.cat([wq, wk], dim=0)
    _lib.check(lib.gb_linear_bwd_x(dqk.data_ptr(), 2 * a, n_pad, 2 * a, w_cat.data_ptr(), d, dx.data_ptr(), d, 1,
                                   _lib.stream()), 'gb_linear_bwd_x')
    if want_w:
        ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n_pad, d, 2 * a)), dtype=torch.uint8, device=dev)
        _lib.check(lib.gb_linear_bwd_w(dqk.data_ptr(), 2 * a, x_full.data_ptr(), d, n_pad, d, 2 * a, grads.data_ptr(),
                                       grads.data_ptr() + 4 * 2 * a * d, ws.data_ptr(), ws.numel(), _lib.stream()),
                   'gb_linear_bwd_w')
    return out, dx, grads


class _NlVjpHalf:
    """buffers of one evaluation + VJP between its row pass and its column pass (nl_local_vjp_rows / _cols)"""
    __slots__ = ('part', 'plan', 'x_full', 'q', 'wq', 'wk', 'alpha_train', 'beta_train', 'x0_slab', 'heads', 'sig',
                 'g_slab', 'f', 'ds', 'dabar', 'abar', 'dqk', 'dx', 'partials', 'grads', 'want_w')


def nl_local_vjp_rows(part, plan, x_full, q, wq, wk, alpha_train, beta_train, x0_slab, heads, alpha_sigmoid, g_slab,
                      want_w, dx_out=None):
    """first half of nl_local_vjp: the row pass only (gb_nl_adjoint_phase, GB_NL_PHASE_ROWS).  On return (i.e. once
    the launch is queued) `.f` holds f for the owned rows - all the solver needs to form the NEXT stage's state, whose
    all-gather can then run while nl_local_vjp_cols() works."""
    from .ops import _nl_workspace
    lib = _lib.load()
    n_pad, d = x_full.shape
    s = part.slab
    a = wq.shape[0]
    dev = x_full.device
    f32 = dict(dtype=torch.float32, device=dev)
    h = _NlVjpHalf()
    h.part, h.plan, h.x_full, h.q, h.wq, h.wk = part, plan, x_full, q, wq, wk
    h.alpha_train, h.beta_train, h.x0_slab, h.heads, h.sig, h.g_slab, h.want_w = \
        alpha_train, beta_train, x0_slab, heads, int(alpha_sigmoid), g_slab, want_w
    h.ds = torch.empty(plan.e, heads, **f32)
    h.dabar = torch.empty(plan.e, **f32)
    h.abar = torch.empty(plan.e, **f32)
    h.dqk = torch.empty(n_pad, 2 * a, **f32)
    h.dqk[:part.lo, :a].zero_()
    h.dqk[part.hi:, :a].zero_()
    h.dx = torch.empty(n_pad, d, **f32) if dx_out is None else dx_out
    h.f = torch.zeros(s, d, **f32) if part.hi - part.lo < s else torch.empty(s, d, **f32)
    h.partials = torch.empty(int(lib.gb_att_bwd_fused_partials(plan.c)), **f32)
    h.grads = torch.empty(2 * a * d + 2 * a + 2 if want_w else 2, **f32)
    _nl_adjoint_phase(lib, h, 1, _nl_workspace(dev))
    return h


def _nl_adjoint_phase(lib, h, phase, workspace):
    part, plan, x_full = h.part, h.plan, h.x_full
    d = x_full.shape[1]
    a = h.wq.shape[0]
    source = h.beta_train is not None and h.x0_slab is not None
    dparams = h.grads[-2:]
    _lib.check(lib.gb_nl_adjoint_phase(plan.c, part.lo, part.hi, x_full.data_ptr(), d, _virt(h.q, part.lo, a), a,
                                       h.wk.data_ptr(), d, h.heads, a // h.heads, d,
                                       _virt(h.x0_slab, part.lo, d) if source else None, d, h.alpha_train.data_ptr(),
                                       h.sig, h.beta_train.data_ptr() if source else None,
                                       _virt(h.g_slab, part.lo, d), d, _virt(h.f, part.lo, d), d, h.ds.data_ptr(),
                                       h.dabar.data_ptr(), h.abar.data_ptr(), h.dqk.data_ptr(), 2 * a,
                                       h.dqk.data_ptr() + 4 * a, 2 * a, h.dx.data_ptr(), d, h.partials.data_ptr(),
                                       dparams.data_ptr(), workspace.data_ptr(), phase, _lib.stream()),
               'gb_nl_adjoint_phase')


def nl_local_vjp_cols(h):
    """second half, first part: column pass + dx += [dq|dk] [Wq;Wk].  Returns dx_share [N_pad, d] (what
    nl_local_vjp returns as its second value) - ready for the reduce-scatter, which can then run while
    nl_local_vjp_weights() forms the weight gradients."""
    from .ops import _nl_workspace
    lib = _lib.load()
    x_full = h.x_full
    n_pad, d = x_full.shape
    a = h.wq.shape[0]
    _nl_adjoint_phase(lib, h, 2, _nl_workspace(x_full.device))
    w_cat = torch.cat([h.wq, h.wk], dim=0)
    _lib.check(lib.gb_linear_bwd_x(h.dqk.data_ptr(), 2 * a, n_pad, 2 * a, w_cat.data_ptr(), d, h.dx.data_ptr(), d, 1,
                                   _lib.stream()), 'gb_linear_bwd_x')
    return h.dx


def nl_local_vjp_weights(h):
    """second half, second part: d[Wq;Wk] = [dq|dk]^T x, d[bq;bk] = column sums -> the flat parameter-gradient share
    (third value of nl_local_vjp; its last two entries, dalpha / dbeta, were written by the column pass)"""
    lib = _lib.load()
    x_full = h.x_full
    n_pad, d = x_full.shape
    a = h.wq.shape[0]
    if h.want_w:
        ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(n_pad, d, 2 * a)), dtype=torch.uint8,
                         device=x_full.device)
        _lib.check(lib.gb_linear_bwd_w(h.dqk.data_ptr(), 2 * a, x_full.data_ptr(), d, n_pad, d, 2 * a, h.grads.data_ptr(),
                                       h.grads.data_ptr() + 4 * 2 * a * d, ws.data_ptr(), ws.numel(), _lib.stream()),
                   'gb_linear_bwd_w')
    return h.grads


def diffusion_local_forward(part, plan, vals, x_full, x0_slab, alpha_train, beta_train, alpha_sigmoid):
    """rows [lo, hi) of alpha (A x - x) + beta x0 (GRAND-l) given the gathered state -> [S, d]"""
    lib = _lib.load()
    n_pad, d = x_full.shape
    s = part.slab
    out = torch.zeros(s, d, dtype=torch.float32, device=x_full.device) if part.hi - part.lo < s else \
        torch.empty(s, d, dtype=torch.float32, device=x_full.device)
    source = beta_train is not None and x0_slab is not None
    _lib.check(lib.gb_spmm_fwd_part(plan.c, part.lo, part.hi, vals.data_ptr(), x_full.data_ptr(), d,
                                    _virt(x0_slab, part.lo, d) if source else None, d, alpha_train.data_ptr(),
                                    int(alpha_sigmoid), beta_train.data_ptr() if source else None, d,
                                    _virt(out, part.lo, d), d, None, _lib.stream()), 'gb_spmm_fwd_part')
    return out


def diffusion_local_vjp(part, plan, vals, x_full, x0_slab, alpha_train, alpha_sigmoid, g_slab, dx_out=None):
    """this rank's share of dx for all columns [N_pad, d] and of (dalpha_train, dbeta_train) [2]"""
    lib = _lib.load()
    n_pad, d = x_full.shape
    dev = x_full.device
    dx = torch.empty(n_pad, d, dtype=torch.float32, device=dev) if dx_out is None else dx_out
    partials = torch.empty(int(lib.gb_spmm_bwd_partials(plan.c, d)), dtype=torch.float32, device=dev)
    dparams = torch.empty(2, dtype=torch.float32, device=dev)
    own_hi = part.hi if part.hi > 0 else -1                  # (0, 0) means "own everything" to the library
    _lib.check(lib.gb_spmm_bwd_part(plan.c, part.lo, own_hi, vals.data_ptr(), _virt(g_slab, part.lo, d), d,
                                    x_full.data_ptr(), d, _virt(x0_slab, part.lo, d), d, alpha_train.data_ptr(),
                                    int(alpha_sigmoid), d, dx.data_ptr(), d, 0, partials.data_ptr(),
                                    dparams.data_ptr(), _lib.stream()), 'gb_spmm_bwd_part')
    return dx, dparams


# ---- autograd over the exchange + local compute -------------------------------------------------------
from torch.autograd.function import once_differentiable  # noqa: E402


class PartNlFn(torch.autograd.Function):
    """partitioned ODEFuncTransformerAtt.forward: slab in, slab out"""

    @staticmethod
    def forward(ctx, x_slab, wq, bq, wk, bk, alpha_train, beta_train, x0_slab, part, halo, plan, heads, alpha_sigmoid):
        x_slab = x_slab.contiguous()
        halo.prepare(plan)
        # Function.forward runs with grad mode off: tell the halo itself whether this state is saved for a backward
        x_full = halo.all_gather_rows(x_slab, keep=any(ctx.needs_input_grad))
        out, q = nl_local_forward(part, plan, x_full, x_slab, wq.contiguous(), bq.contiguous(), wk.contiguous(),
                                  alpha_train, beta_train, x0_slab, heads, alpha_sigmoid)
        ctx.save_for_backward(x_full, q, wq, wk, alpha_train, beta_train, x0_slab)
        ctx.part, ctx.halo, ctx.plan, ctx.heads, ctx.alpha_sigmoid = part, halo, plan, heads, alpha_sigmoid
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g_slab):
        x_full, q, wq, wk, alpha_train, beta_train, x0_slab = ctx.saved_tensors
        a, d = wq.shape
        _, dx_share, grads = nl_local_vjp(ctx.part, ctx.plan, x_full, q, wq.contiguous(), wk.contiguous(), alpha_train,
                                          beta_train, x0_slab, ctx.heads, ctx.alpha_sigmoid, g_slab.contiguous(),
                                          False, True, dx_out=share_buffer(ctx.halo, x_full))
        dx = ctx.halo.reduce_scatter_rows(dx_share)
        ctx.halo.all_reduce_(grads)
        dwq, dwk, dbq, dbk, dalpha, dbeta = unpack_nl_grads(grads, a, d)
        return (dx, dwq, dbq, dwk, dbk, dalpha.reshape(alpha_train.shape),
                dbeta.reshape(()) if beta_train is not None else None) + (None,) * 6


def unpack_nl_grads(grads, a, d):
    """flat (dWq, dWk, dbq, dbk, dalpha, dbeta) as written by nl_local_vjp"""
    w = 2 * a * d
    return (grads[:a * d].view(a, d), grads[a * d:w].view(a, d), grads[w:w + a], grads[w + a:w + 2 * a],
            grads[w + 2 * a], grads[w + 2 * a + 1])


class PartDiffusionFn(torch.autograd.Function):
    """partitioned LaplacianODEFunc.forward with fixed edge values (constant block)"""

    @staticmethod
    def forward(ctx, x_slab, vals, alpha_train, beta_train, x0_slab, part, halo, plan, alpha_sigmoid):
        halo.prepare(plan)
        x_full = halo.all_gather_rows(x_slab.contiguous(), keep=any(ctx.needs_input_grad))
        out = diffusion_local_forward(part, plan, vals, x_full, x0_slab, alpha_train, beta_train, alpha_sigmoid)
        ctx.save_for_backward(x_full, vals, alpha_train, beta_train, x0_slab)
        ctx.part, ctx.halo, ctx.plan, ctx.alpha_sigmoid = part, halo, plan, alpha_sigmoid
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, g_slab):
        x_full, vals, alpha_train, beta_train, x0_slab = ctx.saved_tensors
        if ctx.needs_input_grad[1]:
            raise NotImplementedError('gradients w.r.t. edge values are not available on the partitioned path')
        dx_share, dparams = diffusion_local_vjp(ctx.part, ctx.plan, vals, x_full,
                                                x0_slab if beta_train is not None else None, alpha_train,
                                                ctx.alpha_sigmoid, g_slab.contiguous(),
                                                dx_out=share_buffer(ctx.halo, x_full))
        dx = ctx.halo.reduce_scatter_rows(dx_share)
        ctx.halo.all_reduce_(dparams)
        return (dx, None, dparams[0].reshape(alpha_train.shape),
                dparams[1].reshape(()) if beta_train is not None else None) + (None,) * 5


def nl_eval_vjp_part(part, halo, plan, x_slab, g_slab, wq, bq, wk, alpha_train, beta_train, x0_slab, heads,
                     alpha_sigmoid):
    """partitioned f and VJP in one fused pass (adjoint solver): (f_slab, dx_slab, flat parameter gradients)"""
    x_slab = x_slab.contiguous()
    halo.prepare(plan)
    x_full = halo.all_gather_rows(x_slab)
    from .ops import _nl_q
    q = _nl_q(x_slab, wq, bq)
    f, dx_share, grads = nl_local_vjp(part, plan, x_full, q, wq, wk, alpha_train, beta_train, x0_slab, heads,
                                      alpha_sigmoid, g_slab.contiguous(), True, True, dx_out=share_buffer(halo, x_full))
    dx = halo.reduce_scatter_rows(dx_share)
    return f, dx, grads
