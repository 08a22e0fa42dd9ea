"""CPU: the traffic model bench.py reports against reproduces the survey's figures, and the
bench's reporting helpers behave (no GPU, no compute)."""
import importlib.util
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location('bench_module', os.path.join(ROOT, 'bench.py'))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)

C1 = dict(n=2708, e=13264, d=64, a=64, h=4)
C4 = dict(n=2449029, e=64308169, d=128, a=64, h=4)
C5 = dict(n=1 << 24, e=285212672, d=64, a=64, h=4)


def test_laplacian_bytes_match_the_survey():
    b = bench.algorithmic_bytes(**C1)
    assert b['spmm_fwd'] == 8 * 13264 + 12 * 2708 * 64 + 4 * 2708 == 2196688        # C1 B_L = 2.20 MB
    assert abs(bench.algorithmic_bytes(**C5)['spmm_fwd'] / 1e9 - 15.23) < 0.01        # C5 B_L = 15.23 GB
    assert b['spmm_fwd'] + b['spmm_bwd'] > 16 * C1['e'] + 24 * C1['n'] * C1['d']      # adjoint eval = f + VJP


def test_single_gather_path_moves_no_more_than_the_three_kernel_plan():
    n, e, d, a = C4['n'], C4['e'], C4['d'], C4['a']
    plan = 16 * n * d + 16 * n * a + 12 * e + 4 * n                                   # B_NL
    assert abs(plan / 1e9 - 8.305) < 0.001
    b = bench.algorithmic_bytes(**C4)
    ours = (4 * n * d + 4 * n * a) + b['nl_fwd']                                      # Q projection + one graph kernel
    assert ours < plan


def test_gather_model_adds_one_row_per_extra_edge():
    g = bench.gather_bytes(**C4)
    assert g['nl_fwd'] == 4 * (C4['e'] - C4['n']) * C4['d']
    assert g['fused_att_bwd_col'] == 4 * (C4['e'] - C4['n']) * (C4['a'] + C4['d'])


def test_side_stream_kernels_are_left_out_of_the_roofline_choice():
    k = {'nl_adj_rows': {'ms': 10.0, 'launches': 2, 'overlapped': 0},
         'spmm_bwd_long': {'ms': 30.0, 'launches': 2, 'overlapped': 2}}
    assert list(bench.main_kernels(k)) == ['nl_adj_rows']


def test_measured_traffic_table_is_readable():
    assert bench.measured_traffic('c4', 'fused_att_bwd_col') > 6e10
    assert bench.measured_traffic('c4', 'no_such_kernel') is None
    json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))


def test_per_evaluation_bytes_match_the_survey():
    """C2 B_NL = 61.95 MB; C3 B_L = 206.5 MB, B_NL = 450.2 MB; C4 B_NL = 8.305 GB, B_L = 4.286 GB;
    C5 B_L = 15.23 GB, B_NL = 37.85 GB; adjoint GRAND-l = 16E' + 24Nd + 8N (C5 30.5 GB)"""
    c2 = bench.eval_bytes(n=19717, e=108365, d=128, a=64, h=4)
    c3 = bench.eval_bytes(n=169343, e=1335586, d=96, a=64, h=4)
    c4 = bench.eval_bytes(**C4)
    c5 = bench.eval_bytes(**C5)
    assert abs(c2['transformer']['forward'] / 1e6 - 61.95) < 0.01
    assert abs(c3['laplacian']['forward'] / 1e6 - 206.5) < 0.1 and abs(c3['transformer']['forward'] / 1e6 - 450.2) < 0.1
    assert abs(c4['transformer']['forward'] / 1e9 - 8.305) < 0.001 and abs(c4['laplacian']['forward'] / 1e9 - 4.286) < 0.001
    assert abs(c5['laplacian']['forward'] / 1e9 - 15.23) < 0.01 and abs(c5['transformer']['forward'] / 1e9 - 37.85) < 0.01
    assert abs(c5['laplacian']['adjoint'] / 1e9 - 30.5) < 0.1
    assert c4['transformer']['adjoint'] == 2 * c4['transformer']['forward'] + 8 * C4['e'] * C4['h']
