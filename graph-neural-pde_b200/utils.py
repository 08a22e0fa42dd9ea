"""Host-side mirror of the hot-path helpers of the reference's `src/utils.py`.

Same names and argument meaning as the reference (`MaxNFEException`, `get_rw_adj`,
`gcn_norm_fill_val`, `add_remaining_self_loops`, `Meter`), un-broken (the fork's
versions raise IndexError / scramble edges) and following the intent pinned by the reference's own
tests (test/test_utils.py:62-77).  Accepts the fork's batched layout ([1,2,E], [1,E]) and
upstream's ([2,E], [E]); returns the layout it was given.  Degree sums run on the GPU in the
reference's summation order (ordered segment sums, csrc/plan.cu), so weights are bit-identical
to oracle/graph.py.
"""
import torch

from . import _lib
from .plan import degree


class MaxNFEException(Exception):
    """src/utils.py:13"""
    pass


def _unbatch(edge_index, edge_weight):
    batched = edge_index.dim() == 3
    if batched:
        if edge_index.shape[0] != 1:
            raise NotImplementedError('batched graphs (B > 1): normalise one graph at a time')
        edge_index = edge_index[0]
        if edge_weight is not None:
            edge_weight = edge_weight.reshape(-1)
    return batched, edge_index, edge_weight


def _rebatch(batched, edge_index, edge_weight):
    if batched:
        return edge_index[None], (None if edge_weight is None else edge_weight[None])
    return edge_index, edge_weight


def maybe_num_nodes(edge_index, num_nodes=None):
    """src/utils.py:95-98"""
    return num_nodes if num_nodes is not None else int(edge_index.max()) + 1


def softmax(src, index, num_nodes=None):
    """utils.py:116-127 (PyG 1.7 `softmax`): per head, out = exp(src - max_g) / (sum_g exp(src - max_g) + 1e-16) over the
    groups g = index[e].  src [B,E,H] / index [B,E] (the fork's layout, B == 1) or [E,H] / [E].  The reference does
    scatter_max / exp / scatter_add / divide with float atomics; here the group ids are sorted once (stable) and one
    kernel makes the three passes per group (csrc/gat.cu, gb_segment_softmax_fwd).  The ODE functions do not call
    this - their softmax is fused with the score kernels; it is the stand-alone entry for callers with their own
    scores."""
    from .ops import SegmentSoftmaxFn
    batched = src.dim() == 3
    if batched and src.shape[0] != 1:
        raise NotImplementedError('batched scores (B > 1) are normalised one graph at a time; pass B == 1')
    s2 = src[0] if batched else src
    idx = (index[0] if index.dim() == 2 else index).long()
    n = maybe_num_nodes(idx, num_nodes)
    order = torch.argsort(idx, stable=True)
    ptr = torch.zeros(n + 1, dtype=torch.int32, device=idx.device)
    counts = torch.bincount(idx, minlength=n)
    ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    long_groups = torch.nonzero(counts > SegmentSoftmaxFn.LONG).reshape(-1).to(torch.int32)
    out = SegmentSoftmaxFn.apply(s2, ptr, order.to(torch.int32).contiguous(), idx, long_groups)
    return out[None] if batched else out


def add_remaining_self_loops(edge_index, edge_attr, fill_value, num_nodes):
    """PyG 1.7 semantics (what upstream called; the fork's src/utils.py:16-42 is broken): existing
    self-loops are moved to the end, every node without one gets (i,i) with weight `fill_value`."""
    batched, ei, ew = _unbatch(edge_index, edge_attr)
    row, col = ei[0], ei[1]
    mask = row != col
    loops = torch.arange(num_nodes, dtype=ei.dtype, device=ei.device)
    new_index = torch.cat([ei[:, mask], torch.stack([loops, loops])], dim=1)
    new_w = None
    if ew is not None:
        loop_w = torch.full((num_nodes,), float(fill_value), dtype=ew.dtype, device=ew.device)
        inv = ~mask
        if bool(inv.any()):
            # several self-loops on one node: the LAST one wins (CPU index-assignment order), made
            # deterministic on the GPU with an integer max-reduction over edge ids
            ids = torch.nonzero(inv, as_tuple=False).reshape(-1)
            last = torch.full((num_nodes,), -1, dtype=torch.int64, device=ei.device)
            last.scatter_reduce_(0, row[ids], ids, reduce='amax', include_self=True)
            has = last >= 0
            loop_w[has] = ew[last[has]]
        new_w = torch.cat([ew[mask], loop_w])
    return _rebatch(batched, new_index, new_w)


def get_rw_adj(edge_index, edge_weight=None, norm_dim=1, fill_value=0., num_nodes=None, dtype=None):
    """src/utils.py:215-233: w_e / deg[idx_e], deg = sum of w over edges sharing idx (col for norm_dim 1)."""
    lib = _lib.load()
    batched, ei, ew = _unbatch(edge_index, edge_weight)
    _lib.require_cuda(ei)
    num_nodes = maybe_num_nodes(ei, num_nodes)
    if ew is None:
        ew = torch.ones(ei.shape[1], dtype=torch.float32, device=ei.device)
    ew = ew.to(torch.float32)
    if not fill_value == 0:
        ei, ew = add_remaining_self_loops(ei, ew, fill_value, num_nodes)
    ei = ei.to(torch.int64).contiguous()
    ew = ew.contiguous()
    idx = (ei[0] if norm_dim == 0 else ei[1]).contiguous()
    deg = degree(idx, ew, num_nodes)
    out = torch.empty_like(ew)
    _lib.check(lib.gb_rw_normalize(idx.data_ptr(), ew.data_ptr(), deg.data_ptr(), ew.numel(), out.data_ptr(),
                                   _lib.stream()), 'gb_rw_normalize')
    return _rebatch(batched, ei, out)


def gcn_norm_fill_val(edge_index, edge_weight=None, fill_value=0., num_nodes=None, dtype=None):
    """src/utils.py:177-194: deg^-1/2[row] * w * deg^-1/2[col], deg over col, inf -> 0."""
    lib = _lib.load()
    batched, ei, ew = _unbatch(edge_index, edge_weight)
    _lib.require_cuda(ei)
    num_nodes = maybe_num_nodes(ei, num_nodes)
    if ew is None:
        ew = torch.ones(ei.shape[1], dtype=torch.float32, device=ei.device)
    ew = ew.to(torch.float32)
    if not int(fill_value) == 0:
        ei, ew = add_remaining_self_loops(ei, ew, fill_value, num_nodes)
    ei = ei.to(torch.int64).contiguous()
    ew = ew.contiguous()
    row, col = ei[0].contiguous(), ei[1].contiguous()
    deg = degree(col, ew, num_nodes)
    out = torch.empty_like(ew)
    _lib.check(lib.gb_gcn_normalize(row.data_ptr(), col.data_ptr(), ew.data_ptr(), deg.data_ptr(), ew.numel(),
                                    out.data_ptr(), _lib.stream()), 'gb_gcn_normalize')
    return _rebatch(batched, ei, out)


class Meter(object):
    """src/utils.py:281-302: forward / backward NFE counters"""

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = None
        self.sum = 0
        self.cnt = 0

    def update(self, val):
        self.val = val
        self.sum += val
        self.cnt += 1

    def get_average(self):
        if self.cnt == 0:
            return 0
        return self.sum / self.cnt

    def get_value(self):
        return self.val
