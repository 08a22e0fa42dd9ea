#!/usr/bin/env python
"""Tuning-variant sweep of the single-gather GRAND-nl kernels (csrc/nl_kernels.cuh, knobs GB_NL_*).

    python scripts/sweep_variants.py build            # here (no GPU): compile every variant, print registers / spills
    python scripts/sweep_variants.py run [--workload c4] [--steps 3]   # on the GPU box, ONE gpurun call:
        for every built variant: (with --check) a parity spot check (tests/test_gpu_transformer.py), then
        bench.py with GRANDB200_LIB pointing at it; lines land in gpurun_out/sweep_<workload>_<tag>.json and a
        table in gpurun_out/sweep_<workload>.md

Variants are whole libraries (graph-neural-pde_b200/variants/libgrandb200_<tag>.so, git-ignored like every
.so; they travel to the GPU box with the snapshot).  The default library is variant `default`.
"""
import argparse
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'graph-neural-pde_b200', 'csrc')
VDIR = os.path.join(ROOT, 'graph-neural-pde_b200', 'variants')

# tag -> -D flags.  U = gathers in flight per lane group, CTAS = resident CTAs per SM compiled for,
# WARPS = warps per CTA, CHUNK_ROWS = rows per scheduler grab.  Defaults: fwd U4 / 3 CTAs, adj U8 / 2 CTAs, 8 warps, 4 rows.
VARIANTS = {
    'default': '',
    # adjoint row kernel: more warps per scheduler (DESIGN.md "what comes next" item 1)
    'adj_w10': '-DGB_NL_WARPS_ADJ=10',                               # 20 warps/SM
    'adj_u4': '-DGB_NL_UADJ=4',                                      # fewer gathers in flight, same occupancy
    'adj_u4_w10': '-DGB_NL_UADJ=4 -DGB_NL_WARPS_ADJ=10',             # 20 warps/SM at 96 registers
    'adj_u4_w12': '-DGB_NL_UADJ=4 -DGB_NL_WARPS_ADJ=12',             # 24 warps/SM at 80 registers
    'adj_u4_w6_c3': '-DGB_NL_UADJ=4 -DGB_NL_WARPS_ADJ=6 -DGB_NL_CTAS_ADJ=3',   # 18 warps/SM, 3 smaller CTAs
    'adj_u2_w12': '-DGB_NL_UADJ=2 -DGB_NL_WARPS_ADJ=12',
    'adj_u2_w16': '-DGB_NL_UADJ=2 -DGB_NL_WARPS_ADJ=16',             # 32 warps/SM at 64 registers
    # forward row kernel (default: U4, 3 CTAs of 8 warps = 24 warps/SM at 80 registers)
    'fwd_u8_c2': '-DGB_NL_UFWD=8 -DGB_NL_CTAS_FWD=2',
    'fwd_u2_c4': '-DGB_NL_UFWD=2 -DGB_NL_CTAS_FWD=4',                # 32 warps/SM at 64 registers
    'fwd_w10': '-DGB_NL_WARPS_FWD=10',                               # 30 warps/SM
    'fwd_u2_w10': '-DGB_NL_UFWD=2 -DGB_NL_WARPS_FWD=10',
    # scheduler granularity (small-degree graphs: c1-c3)
    'chunk8': '-DGB_NL_CHUNK_ROWS=8',
    'chunk16': '-DGB_NL_CHUNK_ROWS=16',
    'chunk1': '-DGB_NL_CHUNK_ROWS=1',
}

HEADLINE = ('Li32ELi4ELb0E', 'Li32ELi4ELb1E')          # nl_rows_kernel<32, 4, fwd> / <32, 4, adj> (d in (64, 128], 4 heads)


def lib_path(tag):
    if tag == 'default':
        return os.path.join(ROOT, 'graph-neural-pde_b200', 'libgrandb200.so')
    return os.path.join(VDIR, f'libgrandb200_{tag}.so')


def build(tags):
    rows = []
    subprocess.run(['make', '-C', CSRC, '-j', '8'], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for tag in tags:
        defs = VARIANTS[tag]
        out = subprocess.run(['make', '-C', CSRC, '-j', '8', '-B', 'variant', f'TAG={tag}', f'DEFS={defs} -Xptxas -v'],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if out.returncode != 0:
            rows.append((tag, defs, 'BUILD FAILED', '', ''))
            sys.stderr.write(out.stdout[-2000:])
            continue
        info = {}
        cur = None
        for line in out.stdout.splitlines():
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                cur = m.group(1)
                continue
            if cur and 'spill stores' in line:
                info.setdefault(cur, {})['spill'] = tuple(int(v) for v in re.findall(r'(\d+) bytes', line)[1:3])
            m = re.search(r'Used (\d+) registers', line)
            if cur and m:
                info.setdefault(cur, {})['regs'] = int(m.group(1))
        cells = []
        for key in HEADLINE:
            hit = [v for k, v in info.items() if 'nl_rows_kernel' in k and key in k]
            cells.append(f"{hit[0].get('regs')} regs, spill {hit[0].get('spill', (0, 0))[0]}/{hit[0].get('spill', (0, 0))[1]} B"
                         if hit else 'n/a')
        rows.append((tag, defs or '(defaults)', 'ok', cells[0], cells[1]))
    print('| variant | flags | build | nl_rows_kernel<32,4,fwd> | nl_rows_kernel<32,4,adj> |')
    print('|---|---|---|---|---|')
    for r in rows:
        print('| ' + ' | '.join(r) + ' |')


def run(tags, workload, steps, check=False):
    out_dir = os.path.join(ROOT, 'gpurun_out')
    os.makedirs(out_dir, exist_ok=True)
    table = ['| variant | parity spot check | ms/step | evals/s | nl_fwd ms | nl_adj_rows ms |', '|---|---|---|---|---|---|']
    for tag in tags:
        path = lib_path(tag)
        if not os.path.exists(path):
            table.append(f'| {tag} | library not built | | | | |')
            continue
        env = dict(os.environ, GRANDB200_LIB=path)
        ok = 'not run'
        if check:                                      # ~1 GPU-minute per variant: use it on the candidates worth keeping
            chk = subprocess.run([sys.executable, '-m', 'pytest', os.path.join(ROOT, 'tests', 'test_gpu_transformer.py'), '-q',
                                  '-x', '-m', 'gpu', '-k', 'fused_path or hub_segments or scaled_dot_forward'], env=env,
                                 cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            ok = 'pass' if chk.returncode == 0 else 'FAIL'
            if chk.returncode != 0:
                with open(os.path.join(out_dir, f'sweep_{workload}_{tag}_pytest.log'), 'w') as fh:
                    fh.write(chk.stdout)
        cmd = [sys.executable, os.path.join(ROOT, 'bench.py'), '--workload', workload, '--steps', str(steps), '--warmup', '3',
               '--no-cpu-baseline', '--no-gpu-reference']
        if workload in ('c1', 'c2', 'c3'):
            cmd.append('--cuda-graph')
        res = subprocess.run(cmd, env=env, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        line = res.stdout.strip().splitlines()[-1] if res.stdout.strip() else ''
        with open(os.path.join(out_dir, f'sweep_{workload}_{tag}.json'), 'w') as fh:
            fh.write(line + '\n')
        try:
            j = json.loads(line)
            k = j['roofline']['kernels']
            table.append(f"| {tag} | {ok} | {j['ms_per_step']} | {j['value']} | {k.get('nl_fwd', {}).get('ms_per_launch')} | "
                         f"{k.get('nl_adj_rows', {}).get('ms_per_launch')} |")
        except (ValueError, KeyError):
            table.append(f'| {tag} | {ok} | bench failed (rc {res.returncode}) | | | |')
            with open(os.path.join(out_dir, f'sweep_{workload}_{tag}_stderr.log'), 'w') as fh:
                fh.write(res.stderr[-4000:])
    text = '\n'.join(table) + '\n'
    with open(os.path.join(out_dir, f'sweep_{workload}.md'), 'w') as fh:
        fh.write(text)
    print(text)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('mode', choices=['build', 'run'])
    ap.add_argument('--only', default='', help='comma-separated variant tags (default: all)')
    ap.add_argument('--workload', default='c4')
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--check', action='store_true', help='run the parity spot check before each bench (about 1 GPU-minute each)')
    args = ap.parse_args()
    tags = [t for t in args.only.split(',') if t] or list(VARIANTS)
    for t in tags:
        if t not in VARIANTS:
            sys.exit(f'unknown variant {t}; known: {list(VARIANTS)}')
    if args.mode == 'build':
        build([t for t in tags if t != 'default'])
    else:
        run(tags, args.workload, args.steps, args.check)


if __name__ == '__main__':
    main()
