#!/usr/bin/env python
"""Regenerate profiles/traffic.json from the committed `ncu --set full` summaries (profiles/*ncu_full*.csv, written by
the summariser used in round 2 from `ncu -i <rep> --page raw --csv`): per workload and kernel id of bench.py, the
measured DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, mean over the captured launches) and where
they come from.  bench.py reads this file for `roofline.traffic`.

    python scripts/make_traffic.py          # rewrites profiles/traffic.json
"""
import csv
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# (file glob, workload, {regex on the kernel name: bench.py kernel id}); later files override earlier ones
SOURCES = [
    ('r01zl_ncu_full_col_nl_c4.csv', 'c4', {r'att_col_bwd_kernel': 'fused_att_bwd_col'}),
    ('r02h_ncu_full_nl2_c4.csv', 'c4', {r'nl2_rows_kernel<.*?, 0, ': 'nl_fwd@single-gather', r'nl2_rows_kernel<.*?, 1, ': 'nl_adj_rows@single-gather'}),
    ('r02j_ncu_full_nlk_c4.csv', 'c4', {r'nlk_rows_kernel<\d+, 0, ': 'nl_fwd', r'nlk_rows_kernel<\d+, 1, ': 'nl_adj_rows'}),
    ('r02r_ncu_full_spmm2_c4.csv', 'c4', {r'spmm2_kernel<\d+, 0, ': 'spmm_fwd', r'spmm2_kernel<\d+, 1, ': 'spmm_bwd'}),
    # end of round 2: the three kernels of a C4 step in their final form (forward launches include the fused stage store)
    ('r02zz_ncu_full_c4.csv', 'c4', {r'nlk_rows_kernel<\d+, 0, ': 'nl_fwd', r'nlk_rows_kernel<\d+, 1, ': 'nl_adj_rows',
                                     r'att_col_bwd_kernel': 'fused_att_bwd_col'}),
]


def to_bytes(value, unit):
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}[unit]
    return float(value.replace(',', '')) * scale


def main():
    out = {}
    for name, workload, kernels in SOURCES:
        path = os.path.join(ROOT, 'profiles', name)
        if not os.path.exists(path):
            continue
        rows = list(csv.reader(open(path)))
        if rows[0][0] == 'metric':                         # round-1 summaries: one line per metric, one column per launch
            names = [r[0] for r in rows[1:]]
            hdr, units = names, [r[1] for r in rows[1:]]
            data = [[r[2 + j] for r in rows[1:]] for j in range(len(rows[0]) - 2)]
        else:
            hdr, units, data = rows[0], rows[1], rows[2:]
        col = {h: i for i, h in enumerate(hdr)}
        if 'dram__bytes_read.sum' not in col:
            continue
        for pattern, kid in kernels.items():
            hits = [r for r in data if re.search(pattern, r[col['Kernel Name']])]
            if not hits:
                continue
            tot = [to_bytes(r[col['dram__bytes_read.sum']], units[col['dram__bytes_read.sum']]) +
                   to_bytes(r[col['dram__bytes_write.sum']], units[col['dram__bytes_write.sum']]) for r in hits]
            ms = [float(r[col['gpu__time_duration.sum']].replace(',', '')) *
                  {'ms': 1.0, 'us': 1e-3, 's': 1e3, 'ns': 1e-6}.get(units[col['gpu__time_duration.sum']], 1.0) for r in hits]
            out.setdefault(workload, {})[kid] = round(sum(tot) / len(tot))
            out.setdefault('_sources', {})[f'{workload}/{kid}'] = {
                'file': f'profiles/{name}', 'launches': len(hits), 'ms_under_ncu': round(sum(ms) / len(ms), 3)}
    with open(os.path.join(ROOT, 'profiles', 'traffic.json'), 'w') as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print(json.dumps(out, indent=1, sort_keys=True))


if __name__ == '__main__':
    main()
