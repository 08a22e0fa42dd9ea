#!/usr/bin/env python
"""BASELINE.md section 3 "B-CPU-literal": time the REFERENCE'S OWN modules (imported unmodified from /root/reference/src
behind the import-only stubs of oracle/ref_stubs, B = 1) on the shapes where their dense [N,N] form fits:
C1 (Cora-shaped) LaplacianODEFunc and ODEFuncTransformerAtt, C2 (PubMed-shaped) LaplacianODEFunc.
Runs only where /root/reference exists (the build container - NOT the GPU box); writes profiles/r02_cpu_literal.json.

    python scripts/time_reference_literal.py
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import graph as og                      # noqa: E402
from oracle.make_golden import load_reference, BASE_OPT   # noqa: E402


def timed(fn, reps):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts))


def main():
    ref_utils, ref_lap, ref_att = load_reference()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dev = torch.device('cpu')
    out = {'where': 'build container (no GPU box has /root/reference)', 'cores': cores, 'torch': torch.__version__,
           'note': 'the fork evaluates A x through a dense [N,N] matrix per call; evals/s of ONE forward evaluation under '
                   'no_grad and of one forward + autograd backward', 'cases': {}}
    shapes = {'c1': (2708, 10556, 64), 'c2': (19717, 88648, 128)}
    for name, (n, e, d) in shapes.items():
        ei, w = og.reset_graph_data(og.shaped_graph(n, e, 1234), None, n, 'rw', 1.0)
        ei_t = torch.tensor(ei)[None]
        g = torch.Generator().manual_seed(3)
        x, x0, gout = (torch.randn(1, n, d, generator=g) for _ in range(3))
        opt = dict(BASE_OPT, block='constant', function='laplacian', hidden_dim=d)
        lap = ref_lap.LaplacianODEFunc(d, d, opt, dev)
        lap.edge_index, lap.edge_weight, lap.x0 = ei_t, torch.tensor(w)[None], x0
        with torch.no_grad():
            lap.alpha_train.fill_(0.3)
            lap.beta_train.fill_(0.1)

        def fwd(f=lap):
            with torch.no_grad():
                f.nfe = 0
                f(torch.tensor(0.0), x)

        def fwd_bwd(f=lap):
            f.nfe = 0
            xs = x.clone().requires_grad_(True)
            torch.autograd.grad(f(torch.tensor(0.0), xs), xs, gout)
        reps = 5 if name == 'c1' else 3
        t_f, t_fb = timed(fwd, reps), timed(fwd_bwd, reps)
        out['cases'][f'{name}/LaplacianODEFunc'] = {'N': n, "E'": int(ei.shape[1]), 'd': d, 'ms_per_eval': round(t_f * 1e3, 2),
                                                    'evals_per_s': round(1 / t_f, 2), 'ms_fwd_plus_bwd': round(t_fb * 1e3, 2)}
        print(name, 'laplacian', out['cases'][f'{name}/LaplacianODEFunc'], flush=True)
        if name != 'c1':
            continue                                  # the fork's E x E score matmul needs E'^2 floats beyond C1
        opt = dict(BASE_OPT, block='attention', function='transformer', hidden_dim=d, heads=4, attention_dim=64,
                   attention_type='cosine_sim')       # a per-edge score type (the fork's scaled_dot is the E x E bug)
        att = ref_att.ODEFuncTransformerAtt(d, d, opt, dev)
        att.edge_index, att.x0, att.y = ei_t, x0, None

        def fwd_a():
            with torch.no_grad():
                att.nfe = 0
                att(torch.tensor(0.0), x)

        def fwd_bwd_a():
            att.nfe = 0
            xs = x.clone().requires_grad_(True)
            torch.autograd.grad(att(torch.tensor(0.0), xs), xs, gout)
        t_f, t_fb = timed(fwd_a, 3), timed(fwd_bwd_a, 3)
        out['cases'][f'{name}/ODEFuncTransformerAtt(cosine_sim)'] = {
            'N': n, "E'": int(ei.shape[1]), 'd': d, 'ms_per_eval': round(t_f * 1e3, 2), 'evals_per_s': round(1 / t_f, 2),
            'ms_fwd_plus_bwd': round(t_fb * 1e3, 2)}
        print(name, 'transformer', out['cases'][f'{name}/ODEFuncTransformerAtt(cosine_sim)'], flush=True)
    with open(os.path.join(ROOT, 'profiles', 'r02_cpu_literal.json'), 'w') as fh:
        json.dump(out, fh, indent=1)


if __name__ == '__main__':
    main()
