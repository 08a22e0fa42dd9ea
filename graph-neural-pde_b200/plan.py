"""GraphPlan: the immutable device-side form of one graph (CSR + CSC + long-segment lists).

Built once per `ODEblock.reset_graph_data` (reference: src/base_classes.py:70-90) instead of the
reference's per-evaluation COO -> dense conversion (src/function_laplacian_diffusion.py:39-58), and
cached on the ODE function while `edge_index` is the same tensor (identity + version).
All arrays are int32 device tensors allocated by torch; the CUDA library only borrows them.
"""
import ctypes as C

import torch

from . import _lib

LONG_THRESHOLD = 256       # segments longer than this get one CTA each instead of a sub-warp group


def _as_2xE(edge_index):
    """accept the fork's batched [1,2,E] layout and upstream's [2,E]"""
    if edge_index.dim() == 3:
        if edge_index.shape[0] != 1:
            raise NotImplementedError('batched graphs (B > 1) are evaluated one graph at a time; pass B == 1')
        edge_index = edge_index[0]
    if edge_index.dim() != 2 or edge_index.shape[0] != 2:
        raise ValueError(f'edge_index must be [2,E] or [1,2,E], got {tuple(edge_index.shape)}')
    return edge_index


_ARRAYS = ('rowptr', 'colptr', 'colidx', 'eperm', 'rowidx', 'tperm', 'long_rows', 'long_cols')


def fingerprint(edge_index, num_nodes):
    """content key of (edge list IN ITS ORDER, node count) for the on-disk plan cache: E, n and three wrapping int64
    sums, one of them position weighted (the plan's `eperm` depends on the caller's edge order).  Four reductions
    over the edge list on its device, one host read."""
    ei = _as_2xE(edge_index).to(torch.int64)
    e = int(ei.shape[1])
    if e == 0:
        return (0, int(num_nodes), 0, 0, 0)
    pos = torch.arange(1, e + 1, dtype=torch.int64, device=ei.device)
    mixed = ei[0] * -7046029254386353131 + ei[1]                      # 0x9E3779B97F4A7C15 as int64; products wrap
    sums = torch.stack([ei[0].sum(), ei[1].sum(), (mixed * pos).sum()]).tolist()
    return (e, int(num_nodes)) + tuple(int(v) for v in sums)


class GraphPlan:
    def __init__(self, edge_index, num_nodes, long_threshold=LONG_THRESHOLD, validate=True):
        lib = _lib.load()
        ei = _as_2xE(edge_index)
        _lib.require_cuda(ei)
        ei = ei.to(torch.int64).contiguous()
        self.device = ei.device
        self.n = int(num_nodes)
        self.e = int(ei.shape[1])
        if validate and self.e > 0:
            lo, hi = int(ei.min()), int(ei.max())
            if lo < 0 or hi >= self.n:
                raise ValueError(f'edge_index entries must lie in [0, {self.n}); got [{lo}, {hi}]')
        i32 = dict(dtype=torch.int32, device=self.device)
        self.rowptr = torch.empty(self.n + 1, **i32)
        self.colptr = torch.empty(self.n + 1, **i32)
        self.colidx = torch.empty(self.e, **i32)
        self.eperm = torch.empty(self.e, **i32)
        self.rowidx = torch.empty(self.e, **i32)
        self.tperm = torch.empty(self.e, **i32)
        with torch.cuda.device(self.device):
            ws_bytes = int(lib.gb_plan_workspace_bytes(self.n, self.e))
            ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device=self.device)
            _lib.check(lib.gb_plan_build(ei[0].data_ptr(), ei[1].data_ptr(), self.e, self.n,
                                         self.rowptr.data_ptr(), self.colidx.data_ptr(), self.eperm.data_ptr(),
                                         self.colptr.data_ptr(), self.rowidx.data_ptr(), self.tperm.data_ptr(),
                                         ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_plan_build')
            self.long_threshold = int(long_threshold)
            self.long_rows = self._long_segments(lib, self.rowptr, ws)
            self.long_cols = self._long_segments(lib, self.colptr, ws)
        del ws
        self._finish()

    def _finish(self):
        self.identity_order = bool(self.e == 0 or torch.equal(
            self.eperm, torch.arange(self.e, dtype=torch.int32, device=self.device)))
        self._c = _lib.GbGraph(
            n=self.n, e=self.e, rowptr=self.rowptr.data_ptr(), colidx=_lib.ptr(self.colidx) or 0,
            eperm=_lib.ptr(self.eperm) or 0, colptr=self.colptr.data_ptr(), rowidx=_lib.ptr(self.rowidx) or 0,
            tperm=_lib.ptr(self.tperm) or 0,
            long_rows=self.long_rows.data_ptr() if self.long_rows.numel() else None,
            long_cols=self.long_cols.data_ptr() if self.long_cols.numel() else None,
            n_long_rows=self.long_rows.numel(), n_long_cols=self.long_cols.numel(),
            long_threshold=self.long_threshold, reserved=0)

    def _long_segments(self, lib, ptr_t, ws):
        if self.long_threshold <= 0 or self.e <= self.long_threshold:
            return torch.empty(0, dtype=torch.int32, device=self.device)
        cap = self.e // (self.long_threshold + 1) + 1
        out = torch.empty(cap, dtype=torch.int32, device=self.device)
        count = C.c_int32(0)
        _lib.check(lib.gb_plan_long_segments(ptr_t.data_ptr(), self.n, self.long_threshold, out.data_ptr(), cap,
                                             C.byref(count), ws.data_ptr(), ws.numel(), _lib.stream()),
                   'gb_plan_long_segments')
        return out[:count.value].clone()

    @property
    def c(self):
        """ctypes pointer to the gb_graph struct"""
        return C.byref(self._c)

    # ---- on-disk form (the plan of a 10^8..10^9-edge graph outlives the process) -----------------
    def save(self, path, key=None):
        """the eight int32 arrays + sizes (+ the fingerprint of the edge list they were built from) as one torch file"""
        blob = {name: getattr(self, name).cpu() for name in _ARRAYS}
        blob.update(n=self.n, e=self.e, long_threshold=self.long_threshold, key=key, format=1)
        tmp = f'{path}.tmp{id(self)}'
        torch.save(blob, tmp)
        import os
        os.replace(tmp, path)                                          # readers never see a half-written file

    @classmethod
    def load(cls, path, device, key=None):
        """plan from a file written by save(); no kernel runs.  `key`: refuse a file built from another edge list."""
        blob = torch.load(path, map_location='cpu')
        if blob.get('format') != 1:
            raise ValueError(f'{path}: not a GraphPlan file of this version')
        if key is not None and tuple(blob.get('key') or ()) != tuple(key):
            raise ValueError(f'{path}: plan was built from a different edge list')
        self = cls.__new__(cls)
        self.device = torch.device(device)
        self.n, self.e, self.long_threshold = int(blob['n']), int(blob['e']), int(blob['long_threshold'])
        for name in _ARRAYS:
            t = blob[name]
            if t.dtype != torch.int32:
                raise ValueError(f'{path}: {name} is {t.dtype}, expected int32')
            setattr(self, name, t.to(self.device).contiguous())
        if self.rowptr.numel() != self.n + 1 or self.colptr.numel() != self.n + 1 or \
                any(getattr(self, k).numel() != self.e for k in ('colidx', 'eperm', 'rowidx', 'tperm')):
            raise ValueError(f'{path}: array sizes do not match n = {self.n}, e = {self.e}')
        self._finish()
        return self

    # ---- per-edge reordering between the caller's COO order and CSR-slot order -----------------
    def to_csr_order(self, per_edge):
        """[E] caller order -> [E] CSR-slot order"""
        per_edge = per_edge.reshape(-1).to(torch.float32).contiguous()
        if per_edge.numel() != self.e:
            raise ValueError(f'expected {self.e} per-edge values, got {per_edge.numel()}')
        if self.identity_order:
            return per_edge
        out = torch.empty_like(per_edge)
        _lib.check(_lib.load().gb_gather_f32(per_edge.data_ptr(), self.eperm.data_ptr(), self.e, out.data_ptr(),
                                             _lib.stream()), 'gb_gather_f32')
        return out

    def head_mean_to_csr(self, att_coo):
        """[E,H] caller order -> mean over heads, [E] CSR-slot order (function_laplacian_diffusion.py:46)"""
        att_coo = att_coo.reshape(self.e, -1).to(torch.float32).contiguous()
        out = torch.empty(self.e, dtype=torch.float32, device=self.device)
        if self.e:
            _lib.check(_lib.load().gb_head_mean_gather(att_coo.data_ptr(), att_coo.shape[1], self.eperm.data_ptr(),
                                                       self.e, out.data_ptr(), _lib.stream()), 'gb_head_mean_gather')
        return out

    def to_coo_order(self, rows_csr):
        """[E,W] CSR-slot order -> caller order"""
        if self.identity_order:
            return rows_csr
        rows_csr = rows_csr.contiguous()
        width = rows_csr.shape[1] if rows_csr.dim() == 2 else 1
        out = torch.empty_like(rows_csr)
        _lib.check(_lib.load().gb_scatter_rows_f32(rows_csr.data_ptr(), width, self.eperm.data_ptr(), self.e,
                                                   out.data_ptr(), _lib.stream()), 'gb_scatter_rows_f32')
        return out


def degree(idx, weight, num_nodes):
    """ordered fp32 scatter_add (src/utils.py:191,230): deg[i] = sum_{e: idx_e = i} w_e in edge order"""
    lib = _lib.load()
    _lib.require_cuda(idx, weight)
    idx = idx.to(torch.int64).contiguous()
    e = idx.numel()
    w = None if weight is None else weight.to(torch.float32).contiguous()
    deg = torch.empty(num_nodes, dtype=torch.float32, device=idx.device)
    with torch.cuda.device(idx.device):
        ws = torch.empty(int(lib.gb_degree_workspace_bytes(num_nodes, e)), dtype=torch.uint8, device=idx.device)
        _lib.check(lib.gb_degree(idx.data_ptr(), _lib.ptr(w), e, num_nodes, deg.data_ptr(), ws.data_ptr(),
                                 ws.numel(), _lib.stream()), 'gb_degree')
    return deg


# ---- caches ------------------------------------------------------------------------------------
# ODEblock.reset_graph_data re-assigns edge_index / edge_weight on every block forward
# (src/base_classes.py:86-89) and tests assign them directly, so plans are looked up by tensor
# identity + in-place version, never assumed immutable.
import weakref

_PLAN_CACHE = {}
_PLAN_CACHE_MAX = 8


def get_plan(edge_index, num_nodes):
    key = id(edge_index)
    hit = _PLAN_CACHE.get(key)
    if hit is not None:
        ref, version, n, plan = hit
        if ref() is edge_index and version == edge_index._version and n == num_nodes:
            return plan
    plan = _build_or_load(edge_index, num_nodes)
    if len(_PLAN_CACHE) >= _PLAN_CACHE_MAX:
        for k in [k for k, v in _PLAN_CACHE.items() if v[0]() is None] or list(_PLAN_CACHE)[:1]:
            _PLAN_CACHE.pop(k, None)
    _PLAN_CACHE[key] = (weakref.ref(edge_index), edge_index._version, num_nodes, plan)
    return plan


def _build_or_load(edge_index, num_nodes):
    """GRANDB200_PLAN_CACHE=<dir>: plans are kept on disk under the fingerprint of their edge list and re-read by
    later processes instead of being rebuilt (sort + two permutations over E edges)"""
    import os
    root = os.environ.get('GRANDB200_PLAN_CACHE')
    if not root:
        return GraphPlan(edge_index, num_nodes)
    key = fingerprint(edge_index, num_nodes)
    path = os.path.join(root, 'plan_' + '_'.join(f'{v & 0xffffffffffffffff:x}' for v in key) + '.pt')
    if os.path.exists(path):
        try:
            return GraphPlan.load(path, _as_2xE(edge_index).device, key=key)
        except Exception:                                              # unreadable, damaged or foreign file:
            pass                                                       # rebuild and overwrite it
    plan = GraphPlan(edge_index, num_nodes)
    os.makedirs(root, exist_ok=True)
    plan.save(path, key=key)
    return plan


def cached_per_edge(plan, tensor, tag, make):
    """memoise `make(tensor)` (a CSR-order re-layout of a per-edge tensor without grad) on the plan"""
    cache = plan.__dict__.setdefault('_edge_cache', {})
    key = (tag, id(tensor))
    hit = cache.get(key)
    if hit is not None and hit[0]() is tensor and hit[1] == tensor._version:
        return hit[2]
    out = make(tensor)
    if len(cache) >= 8:
        cache.clear()
    cache[key] = (weakref.ref(tensor), tensor._version, out)
    return out
