import sys; sys.path.insert(0, '.')
import torch, torch.nn.functional as F
import grand_b200
from grand_b200 import ops
def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(it): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / it
for n, d, c in ((2708, 64, 7), (169343, 96, 40), (2449029, 128, 47)):
    z = torch.randn(n, d, device='cuda'); w = torch.randn(c, d, device='cuda'); b = torch.randn(c, device='cuda')
    y = torch.randint(0, c, (n,), device='cuda'); r = torch.rand(n, device='cuda')
    masks = [r < 0.1, (r >= 0.1) & (r < 0.2), r >= 0.2]
    t_ours = timeit(lambda: ops.decode_eval(z, w, b, y, masks)[0])
    def ref():
        lg = F.linear(F.relu(z), w, b)
        return [(lg[m].argmax(-1) == y[m]).sum() / m.sum() for m in masks]
    t_ref = timeit(ref)
    def ref_sync():
        lg = F.linear(F.relu(z), w, b)
        return [(lg[m].argmax(-1) == y[m]).sum().item() / m.sum().item() for m in masks]
    t_ref_sync = timeit(ref_sync)
    t_ours_sync = timeit(lambda: ops.decode_eval(z, w, b, y, masks)[0].tolist())
    bytes_ = n * (4 * d + 8 + 3)
    print(f'decode n={n} d={d} C={c}: ours {t_ours*1e3:.1f} us ({bytes_/t_ours/1e6:.0f} GB/s), with read-back {t_ours_sync*1e3:.1f} us; torch composition {t_ref*1e3:.1f} us, with its 6 .item() {t_ref_sync*1e3:.1f} us')
for n, f, d in ((2708, 1433, 64), (19717, 500, 128), (2449029, 100, 128), (169343, 128, 96)):
    x = torch.randn(n, f, device='cuda'); lin = ops.Linear(f, d).cuda(); ref = torch.nn.Linear(f, d).cuda()
    t1 = timeit(lambda: lin(x)); t2 = timeit(lambda: ref(x))
    by = 4 * n * (f + d)
    print(f'encoder [{n},{f}]x[{f},{d}]: ours {t1*1e3:.1f} us ({by/t1/1e6:.0f} GB/s), cuBLAS nn.Linear (TF32 off) {t2*1e3:.1f} us')
