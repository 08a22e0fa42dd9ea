"""GPU: the tcgen05 (3xTF32, TMEM accumulators) projection GEMMs of csrc/linear_tc.cu against an fp64 matmul,
called through the C ABI exactly as the ODE function calls them (gb_linear_fwd / gb_linear_bwd_x)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 2e-6          # of the result scale: fp32-class accuracy (plain TF32 would be ~1e-3)


def _close(got, ref, what):
    scale = float(ref.abs().max())
    err = float((got.double() - ref).abs().max())
    assert err <= RTOL * scale, f'{what}: max err {err:.3e} at scale {scale:.3e}'


@pytest.mark.parametrize('rows,k,nout', [(1000, 128, 64), (130, 64, 128), (389, 96, 64), (5, 32, 16), (128, 128, 128),
                                         (40000, 128, 64), (2708, 64, 64)])
def test_linear_fwd_tensor_core(rows, k, nout):
    from grand_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device='cuda').manual_seed(rows + k)
    x = torch.randn(rows, k, device='cuda', generator=g)
    w = torch.randn(nout, k, device='cuda', generator=g) / k ** 0.5
    b = torch.randn(nout, device='cuda', generator=g)
    out = torch.full((rows, nout), float('nan'), device='cuda')
    assert lib.gb_linear_tc_eligible(x.data_ptr(), k, w.data_ptr(), k, out.data_ptr(), nout, nout, k, 0) == 1
    _lib.check(lib.gb_linear_fwd(x.data_ptr(), k, rows, k, w.data_ptr(), b.data_ptr(), nout, out.data_ptr(), nout,
                                 _lib.stream()), 'gb_linear_fwd')
    torch.cuda.synchronize()
    _close(out, x.double() @ w.double().t() + b.double(), 'x W^T + b')
    out2 = torch.empty_like(out)
    _lib.check(lib.gb_linear_fwd(x.data_ptr(), k, rows, k, w.data_ptr(), b.data_ptr(), nout, out2.data_ptr(), nout,
                                 _lib.stream()), 'gb_linear_fwd')
    assert torch.equal(out, out2)                               # deterministic


@pytest.mark.parametrize('rows,nout,d', [(1000, 128, 128), (777, 128, 96), (300, 64, 64), (33, 32, 128)])
@pytest.mark.parametrize('accumulate', [0, 1])
def test_linear_bwd_x_tensor_core(rows, nout, d, accumulate):
    from grand_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device='cuda').manual_seed(rows + d)
    dy = torch.randn(rows, nout, device='cuda', generator=g)
    w = torch.randn(nout, d, device='cuda', generator=g) / d ** 0.5
    dx0 = torch.randn(rows, d, device='cuda', generator=g)
    dx = dx0.clone()
    assert lib.gb_linear_tc_eligible(dy.data_ptr(), nout, w.data_ptr(), d, dx.data_ptr(), d, d, nout, 1) == 1
    _lib.check(lib.gb_linear_bwd_x(dy.data_ptr(), nout, rows, nout, w.data_ptr(), d, dx.data_ptr(), d, accumulate,
                                   _lib.stream()), 'gb_linear_bwd_x')
    torch.cuda.synchronize()
    ref = dy.double() @ w.double() + (dx0.double() if accumulate else 0)
    _close(dx, ref, 'dy W')


def test_strided_operands_and_fallback_shapes():
    """q|k packed buffers (leading dimension 2A) go through the tensor-core path; odd widths fall back"""
    from grand_b200 import _lib
    lib = _lib.load()
    rows, k, a = 500, 128, 64
    x = torch.randn(rows, k, device='cuda')
    w = torch.randn(a, k, device='cuda') / k ** 0.5
    b = torch.randn(a, device='cuda')
    buf = torch.zeros(rows, 2 * a, device='cuda')
    _lib.check(lib.gb_linear_fwd(x.data_ptr(), k, rows, k, w.data_ptr(), b.data_ptr(), a, buf.data_ptr() + 4 * a, 2 * a,
                                 _lib.stream()), 'gb_linear_fwd')
    _close(buf[:, a:], x.double() @ w.double().t() + b.double(), 'strided out')
    assert float(buf[:, :a].abs().max()) == 0.0
    assert lib.gb_linear_tc_eligible(x.data_ptr(), 6, w.data_ptr(), 6, buf.data_ptr(), 6, 6, 6, 0) == 0


@pytest.mark.parametrize('rows,d', [(1000, 128), (777, 96), (31, 64), (300001, 128), (4097, 32), (1, 128)])
def test_linear_bwd_w_tensor_core(rows, d):
    """dW = dY^T X and db = column sums on the tensor cores (MN-major operands, per-SM partials added in a
    fixed order); the packed Q|K projection has nout = 2 * attention_dim = 128"""
    from grand_b200 import _lib
    lib = _lib.load()
    nout = 128
    g = torch.Generator(device='cuda').manual_seed(rows + d)
    dy = torch.randn(rows, nout, device='cuda', generator=g)
    x = torch.randn(rows, d, device='cuda', generator=g)
    dw = torch.full((nout, d), float('nan'), device='cuda')
    db = torch.full((nout,), float('nan'), device='cuda')
    ws = torch.empty(int(lib.gb_linear_bwd_w_workspace_bytes(rows, d, nout)), dtype=torch.uint8, device='cuda')
    for _ in range(2):
        _lib.check(lib.gb_linear_bwd_w(dy.data_ptr(), nout, x.data_ptr(), d, rows, d, nout, dw.data_ptr(), db.data_ptr(),
                                       ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
    torch.cuda.synchronize()
    ref_w = dy.double().t() @ x.double()
    ref_b = dy.double().sum(0)
    scale = float((dy.double().abs().t() @ x.double().abs()).max())          # sum of |terms|: the fp32 accumulation scale
    assert float((dw.double() - ref_w).abs().max()) <= 2e-6 * scale
    assert float((db.double() - ref_b).abs().max()) <= 2e-6 * float(dy.double().abs().sum(0).max())
    dw2, db2 = torch.empty_like(dw), torch.empty_like(db)
    _lib.check(lib.gb_linear_bwd_w(dy.data_ptr(), nout, x.data_ptr(), d, rows, d, nout, dw2.data_ptr(), db2.data_ptr(),
                                   ws.data_ptr(), ws.numel(), _lib.stream()), 'gb_linear_bwd_w')
    assert torch.equal(dw, dw2) and torch.equal(db, db2)
