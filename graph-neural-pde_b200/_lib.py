"""ctypes binding of libgrandb200.so (the C ABI declared in include/grandb200.h).

Thin by design: every entry point takes raw device pointers (`tensor.data_ptr()`), sizes and the
current CUDA stream; torch only owns the memory.  There is NO fallback: if the shared library is
missing or a call fails, a RuntimeError carrying the C error string is raised.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GRANDB200_LIB: load a tuning variant (csrc/Makefile `variant`, scripts/sweep_variants.py) instead of the default build
LIB_PATH = os.environ.get('GRANDB200_LIB') or os.path.join(_HERE, 'libgrandb200.so')
HEADER_PATH = os.path.join(_HERE, '..', 'include', 'grandb200.h')

GB_ATT = {'scaled_dot': 0, 'cosine_sim': 1, 'pearson': 2, 'exp_kernel': 3}


class GbGraph(C.Structure):
    """mirror of `gb_graph` (include/grandb200.h)"""
    _fields_ = [('n', C.c_int32), ('e', C.c_int32),
                ('rowptr', C.c_void_p), ('colidx', C.c_void_p), ('eperm', C.c_void_p),
                ('colptr', C.c_void_p), ('rowidx', C.c_void_p), ('tperm', C.c_void_p),
                ('long_rows', C.c_void_p), ('long_cols', C.c_void_p),
                ('n_long_rows', C.c_int32), ('n_long_cols', C.c_int32),
                ('long_threshold', C.c_int32), ('reserved', C.c_int32)]


class GbStage(C.Structure):
    """mirror of `gb_stage`"""
    _fields_ = [('y', C.c_void_p), ('k', C.c_void_p * 6), ('c', C.c_float * 7),
                ('n_prev', C.c_int32), ('write_f', C.c_int32), ('ynext', C.c_void_p),
                ('ld', C.c_int32), ('reserved', C.c_int32)]


_P, _I32, _I64, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
_G = C.POINTER(GbGraph)

# name -> (restype, argtypes); keep in the order of include/grandb200.h
SIGNATURES = {
    'gb_abi_version': (C.c_int, []),
    'gb_last_error': (C.c_char_p, []),
    'gb_device_info': (C.c_int, [C.POINTER(_I32), C.POINTER(_I32), C.POINTER(_I32)]),
    'gb_launch_count': (_I64, []),
    'gb_prof_enable': (C.c_int, [C.c_int]),
    'gb_prof_reset': (C.c_int, []),
    'gb_prof_collect': (C.c_int, []),
    'gb_prof_get': (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(_I64)]),
    'gb_prof_num_kernels': (C.c_int, []),
    'gb_prof_overlapped': (C.c_int64, [C.c_int]),
    'gb_prof_name': (C.c_char_p, [C.c_int]),
    'gb_plan_workspace_bytes': (_SZ, [_I32, _I64]),
    'gb_plan_build': (C.c_int, [_P, _P, _I64, _I32, _P, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    'gb_plan_long_segments': (C.c_int, [_P, _I32, _I32, _P, _I32, C.POINTER(_I32), _P, _SZ, _P]),
    'gb_degree_workspace_bytes': (_SZ, [_I32, _I64]),
    'gb_degree': (C.c_int, [_P, _P, _I64, _I32, _P, _P, _SZ, _P]),
    'gb_rw_normalize': (C.c_int, [_P, _P, _P, _I64, _P, _P]),
    'gb_gcn_normalize': (C.c_int, [_P, _P, _P, _P, _I64, _P, _P]),
    'gb_gather_f32': (C.c_int, [_P, _P, _I64, _P, _P]),
    'gb_head_mean_gather': (C.c_int, [_P, _I32, _P, _I64, _P, _P]),
    'gb_scatter_rows_f32': (C.c_int, [_P, _I32, _P, _I64, _P, _P]),
    'gb_spmm_fwd': (C.c_int, [_G, _P, _P, _I32, _P, _I32, _P, _I32, _P, _I32, _P, _I32, C.POINTER(GbStage), _P]),
    'gb_spmm_fwd_part': (C.c_int, [_G, _I32, _I32, _P, _P, _I32, _P, _I32, _P, _I32, _P, _I32, _P, _I32,
                                   C.POINTER(GbStage), _P]),
    'gb_spmm_bwd_partials': (_SZ, [_G, _I32]),
    'gb_spmm_bwd': (C.c_int, [_G, _P, _P, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _P, _I32, _I32, _P, _P, _P]),
    'gb_spmm_bwd_static': (C.c_int, [_G, _P, _P, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _P, _I32, _I32, _P, _P]),
    'gb_spmm_bwd_part': (C.c_int, [_G, _I32, _I32, _P, _P, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _P, _I32, _I32, _P,
                                   _P, _P]),
    'gb_linear_tc_eligible': (C.c_int, [_P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32]),
    'gb_linear_fwd': (C.c_int, [_P, _I32, _I32, _I32, _P, _P, _I32, _P, _I32, _P]),
    'gb_linear_bwd_x': (C.c_int, [_P, _I32, _I32, _I32, _P, _I32, _P, _I32, _I32, _P]),
    'gb_linear_bwd_w_workspace_bytes': (_SZ, [_I32, _I32, _I32]),
    'gb_linear_bwd_w': (C.c_int, [_P, _I32, _P, _I32, _I32, _I32, _I32, _P, _P, _P, _SZ, _P]),
    'gb_edge_softmax_fwd': (C.c_int, [_G, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _P, _P, _P, _P, _P]),
    'gb_sddmm': (C.c_int, [_G, _P, _I32, _P, _I32, _I32, _P, _I32, _P, _P]),
    'gb_edge_softmax_bwd': (C.c_int, [_G, _I32, _P, _P, _P, _I32, _P, _P, _P]),
    'gb_att_fused_supported': (C.c_int, [_P, _I32, _P, _I32, _I32, _I32, _P, _I32, _I32]),
    'gb_att_fwd_fused': (C.c_int, [_G, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _P, _P, _P, _P, _I32,
                                   _P, _I32, _P, _I32, _P, _I32, _P, _I32, _P, _I32, C.POINTER(GbStage), _P]),
    'gb_att_bwd_fused_partials': (_SZ, [_G]),
    'gb_att_bwd_fused': (C.c_int, [_G, _P, _I32, _P, _I32, _I32, _I32, _P, _P, _P, _I32, _P, _I32, _P, _I32, _P,
                                   _I32, _I32, _P, _P, _P, _I32, _P, _I32, _P, _I32, _P, _P, _P]),
    'gb_stage_combine': (C.c_int, [_P, C.POINTER(C.c_void_p), C.POINTER(C.c_float), _I32, _I64, _P, _P]),
    'gb_stage_combine_multi': (C.c_int, [_P, C.POINTER(C.c_void_p), C.POINTER(C.c_float), _I32, _I64,
                                         C.POINTER(C.c_void_p), _I32, _P]),
    'gb_nl_supported': (C.c_int, [_I32, _I32, _I32]),
    'gb_nl_workspace_bytes': (_SZ, []),
    'gb_nl_fwd': (C.c_int, [_G, _I32, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _I32, _P, _I32, _P,
                            _P, _I32, C.POINTER(GbStage), _P, _P]),
    'gb_nl_adjoint': (C.c_int, [_G, _I32, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _I32, _P, _I32,
                                _P, _P, _I32, _P, _I32, _P, _P, _P, _P, _I32, _P, _I32, _P, _I32, _P, _P, _P, _P]),
    'gb_nl_adjoint_phase': (C.c_int, [_G, _I32, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _I32, _P,
                                      _I32, _P, _P, _I32, _P, _I32, _P, _P, _P, _P, _I32, _P, _I32, _P, _I32, _P, _P, _P,
                                      _I32, _P]),
    'gb_nlk_supported': (C.c_int, [_I32, _I32, _I32]),
    'gb_nlk_fwd': (C.c_int, [_G, _I32, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _I32, _P, _I32, _P,
                             _P, _I32, C.POINTER(GbStage), _P, _P]),
    'gb_nlk_adjoint_phase': (C.c_int, [_G, _I32, _I32, _P, _I32, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _I32, _P,
                                       _I32, _P, _P, _I32, _P, _I32, _P, _P, _P, _P, _I32, _P, _I32, _P, _I32, _P, _P,
                                       _P, _I32, _P]),
    'gb_edge_score_bwd': (C.c_int, [_G, _P, _P, _I32, _P, _I32, _I32, _I32, _I32, _P, _P, _P, _I32, _P, _I32, _P, _P, _P]),
    'gb_gat_softmax_fwd': (C.c_int, [_G, _I32, _P, _P, _I32, _I32, C.c_float, _P, _P, _P, _P]),
    'gb_segment_softmax_fwd': (C.c_int, [_P, _P, _I32, _I32, _P, _P, _P, _P, _I32, _I32, _P]),
    'gb_rows_push': (C.c_int, [_P, _I32, _P, _I32, C.POINTER(C.c_int64), C.POINTER(C.c_void_p), _I32, _I32, _P]),
    'gb_decode_eval_max_classes': (C.c_int, []),
    'gb_decode_eval': (C.c_int, [_P, _I32, _I32, _I32, _P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _I32, _P, _P]),
}

_lib = None


def load():
    """dlopen the library (once) and attach the signatures. Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
            '(or `make -C graph-neural-pde_b200/csrc`). There is no CPU or torch fallback.')
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    return load().gb_last_error().decode('utf-8', 'replace')


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f'{what} failed with gb_status {rc}: {last_error()}')


def ptr(t):
    """device pointer of a tensor (None -> NULL)"""
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def stream():
    """the current CUDA stream as an integer handle.  torch.cuda.current_stream() builds a Stream object (~17 us, more
    than the launch it precedes on the launch-bound configurations); the raw getter is the same query without it"""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError('grand_b200 operators run on CUDA tensors only (no CPU fallback); got a '
                               f'{t.device} tensor')


def launch_count():
    return int(load().gb_launch_count())


class Profiler:
    """per-kernel CUDA-event timing kept inside the library (bench.py roofline numbers)"""

    def __enter__(self):
        lib = load()
        lib.gb_prof_reset()
        lib.gb_prof_enable(1)
        return self

    def __exit__(self, *exc):
        load().gb_prof_enable(0)

    @staticmethod
    def results():
        lib = load()
        check(lib.gb_prof_collect(), 'gb_prof_collect')
        out = {}
        for i in range(lib.gb_prof_num_kernels()):
            ms, cnt = C.c_double(0), C.c_int64(0)
            lib.gb_prof_get(i, C.byref(ms), C.byref(cnt))
            if cnt.value:
                out[lib.gb_prof_name(i).decode()] = {'ms': ms.value, 'launches': cnt.value,
                                                     'overlapped': int(lib.gb_prof_overlapped(i))}
        return out


class nvtx_range:
    """NVTX range around a phase of the solve (forward integration, adjoint solve, halo exchanges) when
    GRANDB200_NVTX=1; a no-op otherwise (no nsys on the build box: the ranges are for timelines taken elsewhere)."""
    _on = os.environ.get('GRANDB200_NVTX') == '1'

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if self._on:
            import torch
            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        if self._on:
            import torch
            torch.cuda.nvtx.range_pop()

