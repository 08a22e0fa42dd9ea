"""GPU: gb_spmm_bwd_static - the allocation-free variant of gb_spmm_bwd (reduction scratch in a file-scope device array,
include/grandb200.h; no counterpart in the reference) - called through
the C ABI and compared with gb_spmm_bwd on the same inputs, two calls in a row (the second call finds the first
call's contents in the shared array).  Collected last on purpose."""
import ctypes as C
import re
import os

import numpy as np
import pytest
import torch

from oracle import graph as og
from helpers import dev, device

pytestmark = pytest.mark.gpu
G_RTOL = 1e-4                       # the gradient tolerance of the other GPU tests
HEADER = os.path.join(os.path.dirname(__file__), '..', 'include', 'grandb200.h')


def _max_floats():
    m = re.search(r'#define\s+GB_SPMM_BWD_STATIC_MAX_FLOATS\s+\(1\s*<<\s*(\d+)\)', open(HEADER).read())
    return 1 << int(m.group(1))


def _inputs(n, e, d, seed):
    import grand_b200
    from grand_b200 import plan as P
    ei, w = og.reset_graph_data(og.shaped_graph(n, e, seed), None, n, 'rw', 1.0)
    plan = P.get_plan(dev(ei), n)
    rng = np.random.default_rng(seed)
    x, x0, gout = (dev(rng.standard_normal((n, d)).astype(np.float32)) for _ in range(3))
    vals = plan.to_csr_order(dev(np.asarray(w, dtype=np.float32)))
    alpha = torch.tensor(0.3, device=device())
    return grand_b200, plan, vals, x, x0, gout, alpha


def _reference(lib, _lib, plan, vals, x, x0, gout, alpha, d):
    n = x.shape[0]
    dx = torch.empty(n, d, device=device())
    dparams = torch.zeros(2, device=device())
    partials = torch.empty(int(lib.gb_spmm_bwd_partials(plan.c, d)), dtype=torch.float32, device=device())
    _lib.check(lib.gb_spmm_bwd(plan.c, vals.data_ptr(), gout.data_ptr(), d, x.data_ptr(), d, x0.data_ptr(), d,
                               alpha.data_ptr(), 1, d, dx.data_ptr(), d, 0, partials.data_ptr(), dparams.data_ptr(),
                               _lib.stream()), 'gb_spmm_bwd')
    return dx, dparams


def _static(lib, _lib, plan, vals, x, x0, gout, alpha, d):
    n = x.shape[0]
    dx = torch.empty(n, d, device=device())                 # outputs are the caller's; the library allocates nothing
    dparams = torch.zeros(2, device=device())
    _lib.check(lib.gb_spmm_bwd_static(plan.c, vals.data_ptr(), gout.data_ptr(), d, x.data_ptr(), d, x0.data_ptr(), d,
                                      alpha.data_ptr(), 1, d, dx.data_ptr(), d, 0, dparams.data_ptr(), _lib.stream()),
               'gb_spmm_bwd_static')
    return dx, dparams


def _close(a, b, what):
    a, b = a.double().cpu(), b.double().cpu()
    err = float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
    assert err <= G_RTOL, f'{what}: max rel err {err:.3e}'


def test_static_scratch_variant_matches_gb_spmm_bwd_on_two_calls_in_a_row():
    # two different problems back to back: the second call must not see the first call's leftovers
    for call, (n, e, d, seed) in enumerate([(3000, 24000, 64, 7), (1700, 9000, 96, 8)]):
        gb, plan, vals, x, x0, gout, alpha = _inputs(n, e, d, seed)
        lib, _lib = gb._lib.load(), gb._lib
        ref_dx, ref_dp = _reference(lib, _lib, plan, vals, x, x0, gout, alpha, d)
        dx, dp = _static(lib, _lib, plan, vals, x, x0, gout, alpha, d)
        torch.cuda.synchronize()
        _close(dx, ref_dx, f'call {call}: dx')
        _close(dp, ref_dp, f'call {call}: dparams')
        assert torch.equal(dx, ref_dx) and torch.equal(dp, ref_dp), 'same kernels, same order: bit-identical'


def test_static_scratch_variant_repeats_bit_identically_and_allocates_nothing():
    gb, plan, vals, x, x0, gout, alpha = _inputs(2500, 30000, 128, 3)
    lib, _lib = gb._lib.load(), gb._lib
    first = _static(lib, _lib, plan, vals, x, x0, gout, alpha, 128)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    n, d = x.shape
    dx = torch.empty(n, d, device=device())
    dparams = torch.zeros(2, device=device())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    for _ in range(2):
        _lib.check(lib.gb_spmm_bwd_static(plan.c, vals.data_ptr(), gout.data_ptr(), d, x.data_ptr(), d, x0.data_ptr(),
                                          d, alpha.data_ptr(), 1, d, dx.data_ptr(), d, 0, dparams.data_ptr(),
                                          _lib.stream()), 'gb_spmm_bwd_static')
        torch.cuda.synchronize()
        assert torch.equal(dx, first[0]) and torch.equal(dparams, first[1])
    free2, _ = torch.cuda.mem_get_info()
    assert free2 == free1, f'device memory changed across the calls: {free1 - free2} bytes (torch side: {free0 - free1})'


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    import grand_b200
    from grand_b200 import plan as P
    lib, _lib = grand_b200._lib.load(), grand_b200._lib
    limit = _max_floats()
    # a graph of self loops only (cheap to build); the slot count grows with n / 32 per 128-column slice, so a wide
    # enough d exceeds the built-in array (the entry point refuses before it looks at any other argument)
    n = 1 << 16
    loops = torch.arange(n, device=device())
    plan = P.get_plan(torch.stack([loops, loops]), n)
    d = 128
    while int(lib.gb_spmm_bwd_partials(plan.c, d)) <= limit:
        d *= 2
        assert d <= (1 << 24), 'no width exceeded the limit'
    assert int(lib.gb_spmm_bwd_partials(plan.c, 128)) <= limit   # ... while the usual widths fit
    t = torch.zeros(8, device=device())
    rc = lib.gb_spmm_bwd_static(plan.c, t.data_ptr(), t.data_ptr(), d, t.data_ptr(), d, t.data_ptr(), d, t.data_ptr(),
                                1, d, t.data_ptr(), d, 0, t.data_ptr(), _lib.stream())
    assert rc == 4, 'GB_ERR_WORKSPACE'                        # refused before anything is launched
    msg = lib.gb_last_error().decode()
    assert 'gb_spmm_bwd_static' in msg and 'floats' in msg and str(limit) in msg
    with pytest.raises(RuntimeError, match='gb_spmm_bwd_static'):
        _lib.check(rc, 'gb_spmm_bwd_static')
