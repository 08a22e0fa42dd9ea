#!/usr/bin/env python
"""bench.py - headline benchmark of the B200 graph-diffusion operator (see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1..c5] [--impl ours|reference]

Metric (BASELINE.json): ODE-function evaluations per second, forward + adjoint backward.
A "step" is one training pass of the reference's call pattern over one synthetic graph:
`AttODEblock(ODEFuncTransformerAtt)` (GRAND-nl, 4 heads) integrated with rk4 over [0, T] under
`odeint_adjoint` (adjoint_method rk4), forward then backward of sum(z * g).  Evaluations are counted
exactly as the reference counts them (`odefunc.nfe`, +1 per forward call; one adjoint-dynamics call
= 1 evaluation + 1 VJP).  value = evals / device time with the state already in HBM; e2e = the same
pass with the state copied from pinned host memory and the result read back, inside the timed region.
One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

T_END = 2.0            # integration time -> 2 rk4 steps forward, 2 backward (--t-end / the workload's preset)
STEP_SIZE = 1.0
# BASELINE.json's configs as the reference names them; flags override single fields
PRESETS = {
    'c1': dict(function='laplacian', block='constant', method='rk4', t_end=10.0),     # GRAND-l, Cora shape, T = 10
    'c2': dict(function='transformer', block='attention', method='dopri5', t_end=2.0),  # GRAND-nl, PubMed shape
    'c3': dict(function='laplacian', block='attention', method='rk4', t_end=2.0),     # BLEND primary, ogbn-arxiv shape
    'c4': dict(function='transformer', block='attention', method='rk4', t_end=2.0),   # GRAND-nl + adjoint (headline)
    'c5': dict(function='transformer', block='attention', method='rk4', t_end=2.0),   # R-MAT scale 24 sweep
}
BLOCK = 'attention'
METHOD = 'rk4'
TOL_SCALE = 1.0


def read_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return float(p['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


def make_opt(cfg, adjoint=True, function='transformer'):
    return {'hidden_dim': cfg['d'], 'block': BLOCK, 'function': function, 'max_nfe': 10 ** 9,
            'no_alpha_sigmoid': False, 'add_source': True, 'augment': False, 'adjoint': adjoint, 'tol_scale': TOL_SCALE,
            'tol_scale_adjoint': TOL_SCALE, 'data_norm': 'rw', 'self_loop_weight': 1.0, 'method': METHOD,
            'step_size': STEP_SIZE, 'max_iters': 100, 'adjoint_method': METHOD, 'adjoint_step_size': STEP_SIZE,
            'leaky_relu_slope': 0.2, 'heads': cfg['heads'], 'attention_dim': cfg['adim'], 'attention_norm_idx': 0,
            'reweight_attention': False, 'beltrami': False, 'attention_type': 'scaled_dot', 'square_plus': False,
            'mix_features': False, 'multi_modal': False, 'kinetic_energy': None, 'jacobian_norm2': None,
            'total_deriv': None, 'directional_penalty': None}


# ---- algorithmic bytes per launch of each kernel (fp32 = 4 B, int32 index = 4 B; every matrix read
# ---- once per kernel that needs it, perfect reuse of gathered rows) - DESIGN.md "Kernels" ---------
def algorithmic_bytes(n, e, d, a, h, kgather=False):
    return {
        'spmm_fwd': 8 * e + 12 * n * d + 4 * n,
        'spmm_bwd': 12 * e + 16 * n * d + 4 * n,
        'edge_softmax_fwd': 8 * n * a + 4 * n + 8 * e + 8 * e * h,
        'sddmm': 8 * n * d + 8 * e + 4 * n,
        'edge_softmax_bwd': 8 * e * h + 4 * e + 4 * n,
        'edge_score_bwd_q': 4 * e * h + 4 * e + 8 * n * a + 4 * n,
        'edge_score_bwd_k': 4 * e * h + 8 * e + 8 * n * a + 4 * n,
        'linear_fwd': 4 * n * d + 8 * n * a,
        'linear_bwd_x': 4 * n * d + 8 * n * a,
        'linear_bwd_w': 4 * n * d + 8 * n * a,
        # fused GRAND-nl kernels (csrc/fused_att.cu).  Forward: the no-store figure (forward integration);
        # the adjoint's recompute launches also write att [E,H] + abar [E] (+ 4EH + 4E).
        'fused_att_spmm_fwd': 8 * n * a + 4 * n + 4 * e + 12 * n * d,
        'fused_att_bwd_row': 8 * n * d + 8 * n * a + 4 * e + 8 * e * h + 4 * n,
        'fused_att_bwd_col': 16 * n * d + 8 * n * a + 12 * e + 4 * e * h + 4 * n,
        # single-gather kernels (csrc/nl_kernels.cuh, nl2_kernels.cuh); k-gather kernels (csrc/nlk_kernels.cuh): + 4NA (k)
        'nl_fwd': 4 * e + 4 * n + 4 * n * a + 12 * n * d + (4 * n * a if kgather else 0),
        'nl_adj_rows': 4 * e + 4 * n + 8 * n * a + 16 * n * d + 4 * e * h + 4 * e + (4 * n * a if kgather else 0),
        'edge_softmax_fwd_long': 0, 'stage_axpy': 0,
    }


def gather_bytes(n, e, d, a, h, kgather=False):
    """no-reuse gather model: every edge fetches its endpoint's rows from HBM.  Added to the
    algorithmic bytes minus the one compulsory read of those matrices."""
    extra = 4 * (e - n) * (a + d)
    row = 4 * (e - n) * (d + a) if kgather else 4 * (e - n) * d
    return {'nl_fwd': row, 'nl_adj_rows': row, 'fused_att_spmm_fwd': extra, 'fused_att_bwd_row': extra, 'fused_att_bwd_col': extra,
            'spmm_fwd': 4 * (e - n) * d, 'spmm_bwd': 4 * (e - n) * d}


def measured_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` on `workload`, from the committed
    `ncu --set full` summaries (profiles/traffic.json, generated by scripts/make_traffic.py); None if not captured"""
    path = os.path.join(ROOT, 'profiles', 'traffic.json')
    try:
        with open(path) as fh:
            return json.load(fh).get(workload, {}).get(kernel)
    except (OSError, ValueError):
        return None


class ClockSampler:
    """samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)"""
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': max(mx), 'reasons': sorted(reasons), 'samples': len(sm)}


# ---- our arm ------------------------------------------------------------------------------------
FUNCTION = 'transformer'      # --function: transformer = GRAND-nl (headline), laplacian = GRAND-l (constant block)


def build_block(cfg, device):
    import grand_b200
    torch.manual_seed(1234)                     # every parameter not set below (V, Wout, ...) is seeded as well
    opt = make_opt(cfg, function=FUNCTION)
    block = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, device,
                                      t=torch.tensor([0., T_END])).to(device)
    block.train()
    g = torch.Generator().manual_seed(7)
    with torch.no_grad():                       # random-init weights of the named architecture (N(0, 1/d))
        layers = [getattr(block.odefunc, 'multihead_att_layer', None), getattr(block, 'multihead_att_layer', None)]
        if True:
            for layer in [l for l in layers if l is not None]:
                for lin in (layer.Q, layer.K):
                    lin.weight.copy_((torch.randn(lin.weight.shape, generator=g) / cfg['d'] ** 0.5).to(device))
                    lin.bias.copy_((0.1 * torch.randn(lin.bias.shape, generator=g)).to(device))   # seeded too: the
                    # default init draws from the global generator and made every process a different model
        block.odefunc.alpha_train.fill_(0.3)
        block.odefunc.beta_train.fill_(0.1)
    return block


def model_name():
    if FUNCTION == 'transformer':
        return 'GRAND-nl'
    return 'GRAND-l' if BLOCK == 'constant' else 'BLEND-style GRAND-l (block-level attention, laplacian function)'


def block_name():
    blk = {'constant': 'ConstantODEblock', 'attention': 'AttODEblock'}[BLOCK]
    return blk + (' + ODEFuncTransformerAtt' if FUNCTION == 'transformer' else ' + LaplacianODEFunc')


def solver_name():
    tol = '' if METHOD in ('rk4', 'euler', 'midpoint') else f', atol {1e-7 * TOL_SCALE:g} rtol {1e-9 * TOL_SCALE:g}'
    step = f' step {STEP_SIZE}' if METHOD in ('rk4', 'euler', 'midpoint') else ''
    return f'{METHOD}{step} T={T_END}{tol}, odeint_adjoint ({METHOD})'


def main_kernels(kernels):
    """kernels timed on the launching stream; the hub-row / hub-column kernels that run on the library's side
    stream next to the main kernel are left out of shares and of the roofline choice (their event interval
    includes the wait for free SMs)"""
    return {k: v for k, v in kernels.items() if not v.get('overlapped')}


def train_pass(block, data, x_dev, g_dev):
    """one step: forward integrate + adjoint backward; returns (evals, loss tensor)"""
    func = block.odefunc
    func.nfe = 0
    x = x_dev.detach().requires_grad_(True)
    block.set_x0(x)
    z = block(x, data)
    loss = (z * g_dev).sum()
    params = [p for p in block.parameters() if p.requires_grad]
    for p in params:
        p.grad = None
    loss.backward()
    return func.nfe, loss


def input_checksums(ei, x_dev, g_dev):
    """order-independent fingerprints of the synthetic inputs (so that two JSON lines can be compared at all)"""
    e = ei.reshape(2, -1)
    return {'graph': int(((e[0] * 1000003 + e[1]) % 2147483647).sum()), 'edges': int(e.shape[1]),
            'x': round(float(x_dev.double().sum()), 6), 'g': round(float(g_dev.double().sum()), 6)}


def loss_checksum(block, data, x_dev, g_dev):
    """sum(z(T) * g) of one extra no-grad solve, accumulated in fp64 (the fp32 loss of the timed passes carries
    ~1e-7 * sum|z g| of summation noise, which hides real differences between runs at different N)"""
    with torch.no_grad():
        block.set_x0(x_dev)
        z = block(x_dev, data)
        tot = torch.zeros((), dtype=torch.float64, device=x_dev.device)
        for lo in range(0, z.shape[0], 1 << 18):
            tot += (z[lo:lo + (1 << 18)].double() * g_dev[lo:lo + (1 << 18)].double()).sum()
    return tot


def run_ours(args):
    import grand_b200
    from grand_b200 import synth, _lib
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    name = args.workload
    cfg = dict(synth.CONFIGS[name])
    if world > 1:
        import torch.distributed as dist
        import datetime
        pg_opts = None
        if os.environ.get('GRANDB200_NCCL_PRIORITY', '1') == '1':
            # NCCL's kernels on high-priority streams: a collective issued next to our kernels (overlapped exchange)
            # gets SMs as soon as thread blocks retire instead of after the queued grid has drained
            pg_opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dist.init_process_group('nccl', device_id=device, timeout=datetime.timedelta(seconds=600), pg_options=pg_opts)
        try:
            return run_partitioned(args, cfg, device, rank, world)
        finally:
            dist.destroy_process_group()
    ei, n = synth.make_graph(name, device)
    data = grand_b200.GraphData()
    data.new_graph(ei, n)
    d = cfg['d']
    gen = torch.Generator(device=device).manual_seed(99)
    x_dev = torch.randn(n, d, generator=gen, device=device)
    g_dev = torch.randn(n, d, generator=gen, device=device)
    x_host = x_dev.cpu().pin_memory()
    block = build_block(cfg, device)

    graphed, profiled_ms_per_step = None, None
    if not args.cuda_graph:
        for _ in range(max(args.warmup, 3)):
            evals, loss = train_pass(block, data, x_dev, g_dev)
    else:
        # launch-bound shapes: the whole pass (forward + adjoint, ~130 launches) replayed as ONE CUDA graph.
        # The warm-up passes run inside GraphedPass on its capture stream (an eager pass on the legacy stream
        # first would leave autograd leaf nodes tied to that stream, which a capture must not touch).
        params = [p for p in block.parameters() if p.requires_grad]
        block.odefunc.nfe = 0

        def whole_pass(x_in, g_in):
            x = x_in.detach().requires_grad_(True)
            block.set_x0(x)
            z = block(x, data)
            loss = (z * g_in).sum()
            return (loss,) + torch.autograd.grad(loss, [x] + params, allow_unused=True)
        graphed = grand_b200.GraphedPass(whole_pass, [x_dev, g_dev], warmup=max(args.warmup, 3))
        evals = evals_per_pass = block.odefunc.nfe // (max(args.warmup, 3) + 1)
    torch.cuda.synchronize()
    e_prime = int(block.odefunc.edge_index.shape[-1])

    # ---- device-resident timing (value) + per-kernel event timing (roofline) ---------------------
    launches0 = _lib.launch_count()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graphed is None:
        with ClockSampler(local) as clocks, _lib.Profiler() as prof:
            torch.cuda.synchronize()
            start.record()
            total_evals = 0
            for _ in range(args.steps):
                evals, loss = train_pass(block, data, x_dev, g_dev)
                total_evals += evals
            stop.record()
            torch.cuda.synchronize()
            kernels = prof.results()
        launches = _lib.launch_count() - launches0
        if start.elapsed_time(stop) / args.steps < 100.0:
            # launch-bound step: the two CUDA events the profiler records around EVERY kernel launch are a measurable part
            # of it (c2: 11 %).  The per-kernel table comes from the profiled pass above, the step time from a second
            # pass of the same K steps without it.
            profiled_ms_per_step = start.elapsed_time(stop) / args.steps
            with ClockSampler(local) as clocks:
                torch.cuda.synchronize()
                start.record()
                for _ in range(args.steps):
                    train_pass(block, data, x_dev, g_dev)
                stop.record()
                torch.cuda.synchronize()
    else:
        with ClockSampler(local) as clocks:
            torch.cuda.synchronize()
            start.record()
            for _ in range(args.steps):
                out = graphed(x_dev, g_dev)
            stop.record()
            torch.cuda.synchronize()
        with _lib.Profiler() as prof:                       # per-kernel times: one eager pass, outside the timing
            train_pass(block, data, x_dev, g_dev)
            torch.cuda.synchronize()
            kernels = prof.results()
        launches = (_lib.launch_count() - launches0) * args.steps      # the graph replays exactly these launches
        total_evals, evals, loss = evals_per_pass * args.steps, evals_per_pass, out[0]
    ms = start.elapsed_time(stop)

    # ---- end-to-end timing: pinned host state -> device, pass, loss + dL/dx back to host --------
    # Every step copies ITS input from pinned host memory and reads its loss and dL/dx back, all inside the
    # timed region.  The copies run on a copy stream like a data loader's: the input of step i+1 is uploaded
    # while step i computes (two device buffers), dL/dx of step i is downloaded while step i+1 computes; the
    # first upload and the last download are exposed.
    gx_host = torch.empty(n, d).pin_memory()
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main_s, copy_s = torch.cuda.current_stream(), torch.cuda.Stream()
    x_bufs = [torch.empty_like(x_dev) for _ in range(2)] if graphed is None else None
    uploaded = [torch.cuda.Event(), torch.cuda.Event()]
    torch.cuda.synchronize()
    s2.record()
    evals_e2e = 0

    def upload(i):
        copy_s.wait_stream(main_s) if i == 0 else None       # step 0: the copy starts inside the timed region
        with torch.cuda.stream(copy_s):
            x_bufs[i % 2].copy_(x_host, non_blocking=True)
            uploaded[i % 2].record(copy_s)

    if graphed is None:
        upload(0)
    for i in range(args.steps):
        if graphed is not None:
            graphed.static_in[0].copy_(x_host, non_blocking=True)
            out = graphed(graphed.static_in[0], g_dev)
            gx_host.copy_(out[1], non_blocking=True)
            loss_host = float(out[0])
            evals_e2e += evals_per_pass
            continue
        main_s.wait_event(uploaded[i % 2])
        if i + 1 < args.steps:
            upload(i + 1)            # the other buffer: its last reader (step i-1) ended at that step's loss read
        func = block.odefunc
        func.nfe = 0
        xg = x_bufs[i % 2].detach().requires_grad_(True)
        block.set_x0(xg)
        z = block(xg, data)
        loss = (z * g_dev).sum()
        gx, = torch.autograd.grad(loss, xg)
        done = torch.cuda.Event()
        done.record(main_s)
        with torch.cuda.stream(copy_s):
            copy_s.wait_event(done)
            gx_host.copy_(gx, non_blocking=True)
            gx.record_stream(copy_s)
        loss_host = float(loss.detach())                    # device -> host read of the step's result
        evals_e2e += func.nfe
    main_s.wait_stream(copy_s)                              # the last download is inside the timed region
    e2.record()
    torch.cuda.synchronize()
    ms_e2e = s2.elapsed_time(e2)

    # ---- roofline of the dominant kernel ---------------------------------------------------------
    peak, peak_src = read_peaks()
    from grand_b200 import ops as _ops
    kgather = FUNCTION == 'transformer' and _ops.nlk_eligible(cfg['heads'], cfg['adim'], d)
    alg = algorithmic_bytes(n, e_prime, d, cfg['adim'], cfg['heads'], kgather)
    gat = gather_bytes(n, e_prime, d, cfg['adim'], cfg['heads'], kgather)
    # the dominant kernel AMONG THOSE WITH A BYTE MODEL (a kernel without one is listed but never chosen)
    modelled = {k: v for k, v in main_kernels(kernels).items() if alg.get(k)}
    top = max((modelled or main_kernels(kernels)).items(), key=lambda kv: kv[1]['ms'])
    kname, kstat = top
    per_launch_ms = kstat['ms'] / kstat['launches']
    achieved = alg.get(kname, 0) / (per_launch_ms * 1e-3) / 1e9 if per_launch_ms > 0 else 0.0
    ksum = sum(v['ms'] for v in main_kernels(kernels).values())
    roofline = {'bound': 'hbm', 'kernel': kname, 'achieved': round(achieved, 1), 'peak': peak, 'unit': 'GB/s',
                'frac': round(achieved / peak, 4), 'peak_source': peak_src,
                'traffic': measured_traffic(name, kname + ('@single-gather' if kname in ('nl_fwd', 'nl_adj_rows') and not kgather else '')),
                'algorithmic_bytes_per_launch': alg.get(kname, 0), 'avg_launch_ms': round(per_launch_ms, 4),
                'share_of_kernel_time': round(kstat['ms'] / ksum, 4) if ksum else None,
                'kernels': {k: {'ms_per_launch': round(v['ms'] / v['launches'], 4), 'launches': v['launches'],
                                'share': None if v.get('overlapped') else round(v['ms'] / ksum, 4),
                                'overlapped_with_main_kernel': bool(v.get('overlapped')),
                                'alg_GBps': round(alg[k] / (v['ms'] / v['launches'] * 1e-3) / 1e9, 1) if alg.get(k) else None,
                                'gather_model_GBps': round((alg[k] + gat[k]) / (v['ms'] / v['launches'] * 1e-3) / 1e9, 1)
                                if k in alg and k in gat else None}
                            for k, v in sorted(kernels.items(), key=lambda kv: -kv[1]['ms'])}}

    line = {
        'metric': 'ODE-func evals/sec (fwd + adjoint bwd)', 'value': round(total_evals / (ms * 1e-3), 3),
        'unit': 'evals/s', 'n_gpus': 1, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': round(ms / args.steps, 3), 'higher_is_better': True, 'scaling': 'strong',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': f"{name}: {model_name()} ({cfg['desc']}), N={n}, E'={e_prime}, d={d}, heads={cfg['heads']}, "
                               f"attention_dim={cfg['adim']}, {block_name()}, {solver_name()}",
                   'row_kernels': 'k-gather (nlk)' if kgather else 'single-gather (nl2)' if FUNCTION == 'transformer' else 'spmm',
                   'evals_per_step': evals, 'cuda_graph': bool(args.cuda_graph),
                   **({'ms_per_step_with_kernel_events': round(profiled_ms_per_step, 3)} if profiled_ms_per_step else {}), 'l2': 'inputs larger than L2 (no flush)' if name in ('c4', 'c5')
                   else 'working set partly L2-resident; no flush (state reused across evaluations as in the solver)'},
        'e2e': {'value': round(evals_e2e / (ms_e2e * 1e-3), 3), 'unit': 'evals/s',
                'h2d_bytes_per_step': n * d * 4, 'd2h_bytes_per_step': n * d * 4 + 4,
                'ms_per_step': round(ms_e2e / args.steps, 3),
                'copies': 'every step uploads its input from pinned host memory and downloads loss + dL/dx; '
                          'copy stream, overlapped with the neighbouring steps (first upload / last download exposed)'
                          if graphed is None else 'in line on the compute stream (CUDA-graph replay)'},
        'gpu_launches': int(launches), 'loss_checksum': round(float(loss_checksum(block, data, x_dev, g_dev)), 6),
        'input_checksums': input_checksums(ei, x_dev, g_dev), 'clocks': clocks.summary(), 'roofline': roofline,
    }
    if args.gpu_reference:
        del block, x_dev, g_dev, graphed
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        line['gpu_reference'] = {'unavailable': 'out of memory in the torch-ops comparator at every sample size'}
        for div in [dv for dv in (1, 2, 4, 8, 16, 32) if dv <= GPU_REF_SCALE[name]]:   # full size first; its [E,d] temporaries may not fit
            try:
                line['gpu_reference'] = gpu_reference(name, cfg, device, div)
                line['gpu_reference']['speedup_of_value'] = round(line['value'] / line['gpu_reference']['value'], 1)
                break
            except torch.cuda.OutOfMemoryError:                              # the comparator's temporaries, not ours
                gc.collect()
                torch.cuda.empty_cache()
    if args.cpu_baseline:
        line['cpu_baseline'] = cpu_baseline(name, cfg, steps=2, warmup=1)
    emit(line)


def eval_bytes(n, e, d, a, h):
    """SURVEY.md 8d algorithmic bytes of ONE evaluation (not one kernel launch): GRAND-l / GRAND-nl, forward
    evaluation and adjoint evaluation (f recomputed + VJP)"""
    b_l = 8 * e + 12 * n * d + 4 * n
    b_nl = 16 * n * d + 16 * n * a + 12 * e + 4 * n
    return {'laplacian': {'forward': b_l, 'adjoint': 16 * e + 24 * n * d + 8 * n},
            'transformer': {'forward': b_nl, 'adjoint': 2 * b_nl + 8 * e * h}}


def run_microbench(args):
    """--microbench: the isolated per-evaluation measurement of SURVEY.md 8d(i).  One forward evaluation
    (`func(t, x)` under no_grad) and one adjoint evaluation (f + VJP: `func.eval_vjp`, or forward + autograd.grad
    where the function has no fused form), each timed alone with CUDA events: `iters` (>= 100) calls after 10
    warm-ups, once back to back (L2 hot: the state stays wherever the previous call left it) and once with L2
    flushed before every call (a buffer of 2 x L2 size is overwritten; every call timed with its own event pair)."""
    import grand_b200
    from grand_b200 import synth, _lib
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    name = args.workload
    cfg = dict(synth.CONFIGS[name])
    ei, n = synth.make_graph(name, device)
    data = grand_b200.GraphData()
    data.new_graph(ei, n)
    d = cfg['d']
    gen = torch.Generator(device=device).manual_seed(99)
    x = torch.randn(n, d, generator=gen, device=device)
    g = torch.randn(n, d, generator=gen, device=device)
    block = build_block(cfg, device)
    block.reset_graph_data(data, x.dtype)
    block.set_x0(x)
    func = block.odefunc
    if FUNCTION == 'laplacian' and block.opt['block'] == 'attention':
        func.attention_weights = block.get_attention_weights(x)
    e_prime = int(func.edge_index.shape[-1])
    params = [p for p in func.parameters() if p.requires_grad]
    t0 = torch.zeros((), device=device)
    iters = max(100, args.steps)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)      # 2 x the 126 MB L2

    def forward_eval():
        with torch.no_grad():
            return func(t0, x)

    def adjoint_eval():
        fused = getattr(func, 'eval_vjp', None)
        res = fused(t0, x, g, params) if fused is not None else None
        if res is not None:
            return res
        with torch.enable_grad():
            xg = x.detach().requires_grad_(True)
            f = func(t0, xg)
            return f, torch.autograd.grad(f, [xg] + params, g, allow_unused=True)

    def timed(fn):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        l0 = _lib.launch_count()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record()
        torch.cuda.synchronize()
        hot_ms = s.elapsed_time(e) / iters
        launches = (_lib.launch_count() - l0) / iters
        pairs = []
        for i in range(iters):
            flush.fill_(i & 0xff)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            pairs.append((a, b))
        torch.cuda.synchronize()
        cold = sorted(a.elapsed_time(b) for a, b in pairs)
        return {'hot_ms': round(hot_ms, 4), 'flushed_ms': round(sum(cold) / iters, 4),
                'flushed_ms_median': round(cold[iters // 2], 4), 'kernels_per_eval': round(launches, 1)}

    func.nfe = 0
    peak, peak_src = read_peaks()
    with ClockSampler(local) as clocks:
        out = {'forward': timed(forward_eval), 'adjoint': timed(adjoint_eval)}
    eb = eval_bytes(n, e_prime, d, cfg['adim'], cfg['heads'])[FUNCTION]
    for kind, r in out.items():
        r['algorithmic_bytes'] = eb[kind]
        for key in ('hot', 'flushed'):
            gbs = eb[kind] / (r[key + '_ms'] * 1e-3) / 1e9
            r[f'alg_GBps_{key}'] = round(gbs, 1)
            r[f'frac_of_peak_{key}'] = round(gbs / peak, 4)
            r[f'evals_per_s_{key}'] = round(1e3 / r[key + '_ms'], 2)
    pair_ms = out['forward']['flushed_ms'] + out['adjoint']['flushed_ms']
    emit({'metric': 'isolated ODE-func evaluations/sec (1 forward + 1 adjoint evaluation, L2 flushed)',
          'value': round(2e3 / pair_ms, 3), 'unit': 'evals/s', 'n_gpus': 1, 'steps': iters, 'warmup': 10,
          'ms_per_step': round(pair_ms, 4), 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None,
          'dtype': 'f32', 'data': 'synthetic',
          'config': {'workload': f"{name}: {model_name()} ({cfg['desc']}), N={n}, E'={e_prime}, d={d}, "
                                 f"heads={cfg['heads']}, attention_dim={cfg['adim']}, isolated evaluation microbench",
                     'l2': 'both: back to back (hot) and a 256 MB buffer overwritten before every call (flushed)'},
          'microbench': out, 'peak_GBps': peak, 'peak_source': peak_src, 'clocks': clocks.summary()})


def run_partitioned(args, cfg, device, rank, world):
    """N > 1: the same training pass with the graph split by rows over the ranks (grand_b200/partition.py):
    one all-gather of the state per evaluation, one reduce-scatter of dL/dx per VJP, all over NCCL.  Total
    work is fixed as N grows (strong scaling).  Timed on the device, max over ranks."""
    import torch.distributed as dist
    import grand_b200
    from grand_b200 import synth, _lib
    name = args.workload
    local = device.index
    n = cfg['n'] if name != 'c5' else 1 << 24
    # rank 0 draws the graph, everybody receives it (identical global graph on every rank)
    if rank == 0:
        ei, n = synth.make_graph(name, device)
        meta = torch.tensor([ei.shape[1], n], device=device)
    else:
        meta = torch.zeros(2, dtype=torch.int64, device=device)
    dist.broadcast(meta, 0)
    if rank != 0:
        ei = torch.empty(2, int(meta[0]), dtype=torch.int64, device=device)
    n = int(meta[1])
    dist.broadcast(ei, 0)
    data = grand_b200.GraphData()
    data.new_graph(ei, n)
    d = cfg['d']
    part = grand_b200.RowPartition(n, world, rank)
    halo_kind = args.halo
    halo = {'dense': grand_b200.Halo, 'compact': grand_b200.CompactHalo, 'push': grand_b200.PushHalo, 'copy': grand_b200.CopyHalo,
            'compact_push': grand_b200.CompactPushHalo}.get(
        halo_kind, grand_b200.SparseHalo)(part)
    # the SAME global state and cotangent as the 1-GPU run (seed 99), sliced to this rank's slab: the loss checksum of
    # the JSON line is comparable across N
    gen = torch.Generator(device=device).manual_seed(99)
    x_dev = part.to_slab(torch.randn(n, d, generator=gen, device=device)).contiguous()
    g_dev = part.to_slab(torch.randn(n, d, generator=gen, device=device)).contiguous()
    x_host = x_dev.cpu().pin_memory()
    block = build_block(cfg, device)
    block.set_partition(part, halo)
    referenced = None
    if halo_kind not in ('dense', 'push', 'copy'):
        train_pass(block, data, x_dev, g_dev)           # builds the per-peer index lists (SparseHalo.prepare)
        frac = torch.tensor([halo.referenced_fraction], device=device)
        dist.all_reduce(frac, op=dist.ReduceOp.MAX)
        referenced = round(float(frac), 4)
        if halo_kind == 'auto' and referenced > 0.7:   # dense halo (e.g. c4): whole slabs are cheaper than index lists
            halo_kind, halo = 'dense', grand_b200.Halo(part)
            block.set_partition(part, halo)
        elif halo_kind == 'auto':
            halo_kind = 'sparse'

    for _ in range(max(args.warmup, 3)):
        evals, loss = train_pass(block, data, x_dev, g_dev)
    torch.cuda.synchronize()
    gen2 = torch.Generator(device=device).manual_seed(99)
    in_sums = input_checksums(ei, torch.randn(n, d, generator=gen2, device=device), torch.randn(n, d, generator=gen2, device=device))
    del ei
    e_local = int(block.odefunc.edge_index.shape[-1])
    e_total = torch.tensor([e_local], device=device)
    dist.all_reduce(e_total)

    def timed(fn):
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        start.record()
        out = fn()
        stop.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([start.elapsed_time(stop)], device=device)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)       # max over ranks
        return float(ms), out

    launches0 = _lib.launch_count()
    halo.bytes_in = halo.n_gathers = halo.n_scatters = halo.n_prefetched = 0

    last_loss = [None]

    def loop():
        tot = 0
        for _ in range(args.steps):
            ev, last_loss[0] = train_pass(block, data, x_dev, g_dev)
            tot += ev
        return tot

    with ClockSampler(local) as clocks, _lib.Profiler() as prof:
        ms, total_evals = timed(loop)
        kernels = prof.results()
    launches = _lib.launch_count() - launches0
    halo_bytes = halo.bytes_in
    n_gathers, n_scatters = halo.n_gathers // args.steps, halo.n_scatters // args.steps

    gx_host = torch.empty(part.slab, d).pin_memory()
    copy_s = torch.cuda.Stream()
    x_bufs = [torch.empty_like(x_dev) for _ in range(2)]

    def loop_e2e():
        tot = 0
        # as in the 1-GPU arm: uploads / downloads on a copy stream, overlapped with the neighbouring steps
        main_s = torch.cuda.current_stream()
        copy_s.wait_stream(main_s)
        uploaded = [torch.cuda.Event(), torch.cuda.Event()]

        def upload(i):
            with torch.cuda.stream(copy_s):
                x_bufs[i % 2].copy_(x_host, non_blocking=True)
                uploaded[i % 2].record(copy_s)
        upload(0)
        for i in range(args.steps):
            main_s.wait_event(uploaded[i % 2])
            if i + 1 < args.steps:
                upload(i + 1)
            func = block.odefunc
            func.nfe = 0
            xg = x_bufs[i % 2].detach().requires_grad_(True)
            block.set_x0(xg)
            z = block(xg, data)
            loss = (z * g_dev).sum()
            gx, = torch.autograd.grad(loss, xg)
            done = torch.cuda.Event()
            done.record(main_s)
            with torch.cuda.stream(copy_s):
                copy_s.wait_event(done)
                gx_host.copy_(gx, non_blocking=True)
                gx.record_stream(copy_s)
            float(loss.detach())
            tot += func.nfe
        main_s.wait_stream(copy_s)
        return tot

    ms_e2e, evals_e2e = timed(loop_e2e)
    # collective: EVERY rank takes part (before the rank-0-only reporting below)
    exchange = exchange_times(halo, part, d, device, n_gathers, n_scatters, ms / args.steps)
    loss_sum = loss_checksum(block, data, x_dev, g_dev).reshape(1)
    dist.all_reduce(loss_sum)                           # = the 1-GPU checksum: same graph, weights, state and cotangent
    if rank != 0:
        return
    peak, peak_src = read_peaks()
    rows = part.hi - part.lo
    alg = algorithmic_bytes(rows, e_local, d, cfg['adim'], cfg['heads'])
    gat = gather_bytes(rows, e_local, d, cfg['adim'], cfg['heads'])
    modelled = {k: v for k, v in main_kernels(kernels).items() if alg.get(k)}
    kname, kstat = max((modelled or main_kernels(kernels)).items(), key=lambda kv: kv[1]['ms'])
    per_launch_ms = kstat['ms'] / kstat['launches']
    achieved = alg.get(kname, 0) / (per_launch_ms * 1e-3) / 1e9 if per_launch_ms > 0 else 0.0
    ksum = sum(v['ms'] for v in main_kernels(kernels).values())
    roofline = {'bound': 'hbm', 'kernel': kname, 'achieved': round(achieved, 1), 'peak': peak, 'unit': 'GB/s',
                'frac': round(achieved / peak, 4), 'traffic': None, 'peak_source': peak_src,
                'algorithmic_bytes_per_launch': alg.get(kname, 0), 'avg_launch_ms': round(per_launch_ms, 4),
                'note': "rank 0's kernels on its row slab (algorithmic bytes of the slab)",
                'kernels': {k: {'ms_per_launch': round(v['ms'] / v['launches'], 4), 'launches': v['launches'],
                                'share': None if v.get('overlapped') else round(v['ms'] / ksum, 4),
                                'overlapped_with_main_kernel': bool(v.get('overlapped')),
                                'gather_model_GBps': round((alg[k] + gat[k]) / (v['ms'] / v['launches'] * 1e-3) / 1e9, 1)
                                if k in alg and k in gat else None}
                            for k, v in sorted(kernels.items(), key=lambda kv: -kv[1]['ms'])}}
    line = {
        'metric': 'ODE-func evals/sec (fwd + adjoint bwd)', 'value': round(total_evals / (ms * 1e-3), 3),
        'unit': 'evals/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': round(ms / args.steps, 3), 'higher_is_better': True, 'scaling': 'strong',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': f"{name}: {model_name()} ({cfg['desc']}), N={n}, E'={int(e_total)}, d={d}, heads={cfg['heads']}, "
                               f"attention_dim={cfg['adim']}, {block_name()}, {solver_name()}",
                   'parallelism': f'1-D row partition over {world} GPUs ({part.slab} rows each), all-gather of the '
                                  f'state per evaluation + reduce-scatter of dL/dx per VJP (NCCL)',
                   'evals_per_step': evals, 'l2': 'inputs larger than L2 (no flush)',
                   'overlapped_exchange': os.environ.get('GRANDB200_OVERLAP', '1') == '1',
                   'halo': halo_kind + {'sparse': ' (all_to_all_v of the referenced rows only)',
                                        'compact': ' (all_to_all_v of the referenced rows only; local graph relabelled to '
                                                   '[own slab | referenced rows])',
                                        'push': ' (all-gather = one broadcast-store kernel into symmetric memory + barrier; '
                                                'reduce-scatter = barrier + one kernel pulling the peers\' shares)',
                                        'compact_push': ' (compact halo; both exchanges are one gb_rows_push kernel per rank '
                                                        'storing the requested rows straight into the peers\' symmetric matrices + a barrier)',
                                        'copy': ' (all-gather = P-1 copy-engine copies into the peers\' symmetric matrices + '
                                                'barrier, no SMs; reduce-scatter as push)'}.get(
                                            halo_kind, ' (all-gather / reduce-scatter of whole slabs)'),
                   'rows_of_the_local_problem_rank0': getattr(halo, 'view', part).n_pad,
                   'remote_rows_referenced_max_over_ranks': referenced},
        'e2e': {'value': round(evals_e2e / (ms_e2e * 1e-3), 3), 'unit': 'evals/s',
                'h2d_bytes_per_step': world * part.slab * d * 4, 'd2h_bytes_per_step': world * (part.slab * d * 4 + 4),
                'ms_per_step': round(ms_e2e / args.steps, 3)},
        'gpu_launches': int(launches), 'loss_checksum': round(float(loss_sum), 6), 'input_checksums': in_sums, 'clocks': clocks.summary(), 'roofline': roofline,
        'nvlink': {'bytes_in_per_rank_per_step': halo_bytes // args.steps,
                   'GBps_in_per_rank_if_serial': round(halo_bytes / (ms * 1e-3) / 1e9, 1), 'peak_GBps': 770.0,
                   'prefetched_gathers_per_step': getattr(halo, 'n_prefetched', 0) // args.steps,
                   'note': 'received bytes of the all-gathers / reduce-scatters divided by the WHOLE step time; with '
                           'GRANDB200_OVERLAP=1 the all-gather of the next adjoint stage runs under the column pass of '
                           'the current one (prefetched_gathers_per_step of them), otherwise exchanges are serial'},
    }
    line['nvlink'].update(exchange)
    emit(line)


@torch.no_grad()
def exchange_times(halo, part, d, device, n_gathers, n_scatters, step_ms):
    """the two collectives of the partitioned path timed alone (CUDA events, max over ranks), outside the timed
    region: how much of a step is exchange, and the NVLink rate the exchanges themselves reach"""
    import torch.distributed as dist
    slab = torch.zeros(part.slab, d, device=device)
    full = torch.zeros(getattr(halo, 'view', part).n_pad, d, device=device)
    out = {}
    for key, fn in (('all_gather_ms', lambda: halo.all_gather_rows(slab)),
                    ('reduce_scatter_ms', lambda: halo.reduce_scatter_rows(full))):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        a.record()
        for _ in range(10):
            fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / 10], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[key] = round(float(t), 3)
    recv_g = recv_s = (part.world - 1) * part.slab * d * 4
    if hasattr(halo, 'recv_splits'):                     # SparseHalo: only the referenced rows travel
        recv_g, recv_s = sum(halo.recv_splits) * d * 4, sum(halo.send_splits) * d * 4
    out['all_gather_GBps_in'] = round(recv_g / (out['all_gather_ms'] * 1e-3) / 1e9, 1)
    out['reduce_scatter_GBps_in'] = round(recv_s / (out['reduce_scatter_ms'] * 1e-3) / 1e9, 1)
    out['bytes_in_per_all_gather'], out['bytes_in_per_reduce_scatter'] = recv_g, recv_s
    out['exchanges_per_step'] = {'all_gather': n_gathers, 'reduce_scatter': n_scatters}
    ex = n_gathers * out['all_gather_ms'] + n_scatters * out['reduce_scatter_ms']
    out['exchange_ms_per_step'] = round(ex, 2)
    out['exchange_share_of_step'] = round(ex / step_ms, 3)
    return out


# ---- GPU comparator: the upstream-semantics path with stock torch CUDA ops -----------------------------
GPU_REF_SCALE = {'c1': 1, 'c2': 1, 'c3': 1, 'c4': 8, 'c5': 32}


def gpu_reference(name, cfg, device, div_override=None):
    """SURVEY.md 8d comparator (3): what the reference's GPU path does when its sparse dependencies are present -
    index_select gathers, [E,d] temporaries, scatter_add / scatter_max (float atomics), autograd backward -
    written with plain torch CUDA ops (torch_sparse / torch_scatter / PyG are not installable here; they
    lower to exactly these kernels).  Timed on a down-scaled sample of the same shape (its [E,d] temporaries do
    not fit otherwise), 1 forward evaluation + 1 adjoint evaluation, extrapolated linearly.  None of our kernels."""
    import math
    from grand_b200 import synth
    div = div_override or GPU_REF_SCALE[name]
    ei, n = synth.make_graph(name, device, seed=77, scale_div=div)
    loops = torch.arange(n, device=device)
    ei = torch.cat([ei[:, ei[0] != ei[1]], torch.stack([loops, loops])], dim=1)
    row, col = ei[0], ei[1]
    d, a, h = cfg['d'], cfg['adim'], cfg['heads']
    dk = a // h
    g = torch.Generator(device=device).manual_seed(3)
    mk = lambda *shape: torch.randn(*shape, generator=g, device=device)
    x, x0, gout = mk(n, d), mk(n, d), mk(n, d)
    wq, wk = (mk(a, d) / d ** 0.5).requires_grad_(True), (mk(a, d) / d ** 0.5).requires_grad_(True)
    bq, bk = mk(a).requires_grad_(True), mk(a).requires_grad_(True)
    alpha, beta = torch.tensor(0.3, device=device, requires_grad=True), torch.tensor(0.1, device=device, requires_grad=True)

    def f_eval(xx):
        if FUNCTION == 'transformer':
            q, k = xx @ wq.t() + bq, xx @ wk.t() + bk
            prods = (q.view(n, h, dk)[row] * k.view(n, h, dk)[col]).sum(dim=2) / math.sqrt(dk)
            idx = row[:, None].expand(-1, h)
            mx = torch.zeros(n, h, device=device).scatter_reduce(0, idx, prods.detach(), reduce='amax', include_self=False)
            ex = (prods - mx[row]).exp()
            den = torch.zeros(n, h, device=device).index_add_(0, row, ex)
            vals = (ex / (den[row] + 1e-16)).mean(dim=1)
        else:
            vals = torch.full((row.numel(),), 0.05, device=device)
        ax = torch.zeros_like(xx).index_add_(0, row, vals[:, None] * xx[col])
        return torch.sigmoid(alpha) * (ax - xx) + beta * x0

    params = (wq, bq, wk, bk, alpha, beta) if FUNCTION == 'transformer' else (alpha, beta)

    def one_pass():
        with torch.no_grad():
            f_eval(x)
        xs = x.detach().requires_grad_(True)
        torch.autograd.grad(f_eval(xs), (xs,) + params, gout)

    one_pass()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    start.record()
    for _ in range(reps):
        one_pass()
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / reps
    del x, x0, gout, ei
    torch.cuda.empty_cache()
    return {'value': round(2 / (ms * 1e-3) / div, 3), 'unit': 'evals/s', 'extrapolated': div > 1, 'kind': 'torch CUDA ops (index_select, [E,d] '
            'temporaries, scatter_add / scatter_reduce atomics, autograd) = what torch_sparse.spmm + PyG softmax lower to',
            'sample': (f'EXTRAPOLATED from a 1/{div}-scale sample: ' if div > 1 else 'full size: ') +
                      f'{name} shape scaled 1/{div} (N={n}, E\'={row.numel()}), 1 forward + 1 adjoint evaluation, '
                      f'{ms:.2f} ms per pair, evals/s divided by {div}'}


# ---- CPU baseline / reference arm (the only place bench.py executes oracle/) ----------------------
def cpu_pass(state):
    """one forward evaluation + one adjoint-dynamics evaluation (f recomputed + VJP) of the oracle port"""
    from oracle import funcs
    x, ei, wq, bq, wk, bk, x0, g, heads = state
    if FUNCTION == 'laplacian':
        with torch.no_grad():
            funcs.laplacian_f(x, ei, state.w, state.alpha, state.beta, x0)
        xs = x.detach().requires_grad_(True)
        f = funcs.laplacian_f(xs, ei, state.w, state.alpha, state.beta, x0)
        torch.autograd.grad(f, (xs, state.alpha, state.beta), g)
        return 2
    with torch.no_grad():
        funcs.transformer_f(x, ei, wq, bq, wk, bk, heads, state.alpha, state.beta, x0)
    xs = x.detach().requires_grad_(True)
    f = funcs.transformer_f(xs, ei, wq, bq, wk, bk, heads, state.alpha, state.beta, x0)
    torch.autograd.grad(f, (xs, wq, bq, wk, bk, state.alpha, state.beta), g)
    return 2


class CpuState(tuple):
    pass


def make_cpu_state(name, cfg, scale_div):
    from oracle import graph as og
    import numpy as np
    if name == 'c5':
        import math
        ei = og.rmat_graph(24 - int(round(math.log2(scale_div))), 16, 1234)
        n = 1 << (24 - int(round(math.log2(scale_div))))
    else:
        n = cfg['n'] // scale_div
        ei = og.shaped_graph(n, (cfg['e'] // scale_div) // 2 * 2, 1234)
    ei, w_rw = og.reset_graph_data(ei, None, n, 'rw', 1.0)
    d, a = cfg['d'], cfg['adim']
    g = torch.Generator().manual_seed(5)
    mk = lambda *s: torch.randn(*s, generator=g)
    st = CpuState((mk(n, d), torch.tensor(ei), (mk(a, d) / d ** 0.5).requires_grad_(True), mk(a).requires_grad_(True),
                   (mk(a, d) / d ** 0.5).requires_grad_(True), mk(a).requires_grad_(True), mk(n, d), mk(n, d),
                   cfg['heads']))
    st.alpha = torch.tensor(0.3, requires_grad=True)
    st.beta = torch.tensor(0.1, requires_grad=True)
    st.n, st.e = n, ei.shape[1]
    st.w = torch.tensor(w_rw)
    return st


CPU_SCALE = {'c1': 1, 'c2': 1, 'c3': 2, 'c4': 16, 'c5': 64}      # keeps the CPU leg at ~10-30 s


def cpu_baseline(name, cfg, steps=1, warmup=1):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    div = CPU_SCALE[name]
    st = make_cpu_state(name, cfg, div)
    for _ in range(warmup):
        cpu_pass(st)
    t0 = time.perf_counter()
    evals = 0
    for _ in range(steps):
        evals += cpu_pass(st)
    dt = time.perf_counter() - t0
    return {'value': round(evals / dt / div, 4), 'unit': 'evals/s', 'cores': cores, 'kind': 'port',
            'extrapolated': div > 1,
            'sample': ('EXTRAPOLATED from a 1/%d-scale sample: ' % div if div > 1 else '') + f'oracle port (torch CPU fp32, upstream-semantics sparse restatement, {model_name()}) on the {name} shape scaled '
                      f'1/{div} (N={st.n}, E\'={st.e}); {steps} x (1 forward eval + 1 adjoint eval); evals/s divided '
                      f'by {div} (linear-cost extrapolation to the full shape)',
            'seconds': round(dt, 2)}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from grand_b200 import synth
    cfg = dict(synth.CONFIGS[args.workload])
    base = cpu_baseline(args.workload, cfg, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 1)))
    line = {'impl': 'reference', 'metric': 'ODE-func evals/sec (fwd + adjoint bwd)', 'value': base['value'],
            'unit': 'evals/s', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': round(base['seconds'] / max(1, args.steps) * 1e3, 3), 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': f"{args.workload}: {model_name()} ({cfg['desc']}), N={cfg['n']}, E'={cfg['e'] + cfg['n']}, "
                                   f"d={cfg['d']}, heads={cfg['heads']}, attention_dim={cfg['adim']}, {block_name()}, "
                                   f"{solver_name()}",
                       'arm': 'reference path on the host cores (oracle port of the upstream-semantics sparse path; the '
                              'reference itself needs dense [N,N] memory and uninstallable sparse packages), bounded sample, '
                              'see cpu_baseline.sample'},
            'cpu_baseline': base,
            'e2e': {'value': base['value'], 'unit': 'evals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    emit(line)


def emit(line):
    """the ONE JSON line goes to the real stdout; everything else (NCCL banners, library chatter written to
    fd 1) was redirected to stderr by main()"""
    os.write(_REAL_STDOUT, (json.dumps(line) + '\n').encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='c4', choices=['c1', 'c2', 'c3', 'c4', 'c5'])
    ap.add_argument('--cuda-graph', dest='cuda_graph', action='store_true', default=None,
                    help='replay the whole pass as one CUDA graph (1 GPU).  Default: on for the launch-bound fixed-step '
                         'presets c1 and c3, off otherwise (an adaptive solve decides on the host after every step)')
    ap.add_argument('--eager', dest='cuda_graph', action='store_false', help='no CUDA-graph replay')
    ap.add_argument('--no-cpu-baseline', dest='cpu_baseline', action='store_false')
    ap.add_argument('--no-gpu-reference', dest='gpu_reference', action='store_false',
                    help='skip the torch-CUDA-ops comparator (N = 1 only)')
    ap.add_argument('--function', default=None, choices=['transformer', 'laplacian'],
                    help='transformer = GRAND-nl (BASELINE.json headline); laplacian = GRAND-l sweep of config 5')
    ap.add_argument('--halo', default=os.environ.get('GRANDB200_HALO'), choices=['dense', 'sparse', 'compact', 'compact_push', 'push', 'copy', 'auto'],
                    help='N > 1: exchange whole slabs (dense), only the rows the peers reference (sparse: R-MAT slabs '
                         'reference ~25 %% of the others at 8 ranks), sparse exchange + the rank\'s graph relabelled to '
                         '[own slab | referenced rows] so all per-rank work is O(N/P + halo) (compact), or dense / sparse '
                         'decided from the measured fraction (auto); push = whole slabs stored straight into the peers\' symmetric memory '
                         'by one kernel instead of a NCCL all-gather')
    ap.add_argument('--microbench', action='store_true',
                    help='isolated per-evaluation timing (SURVEY.md 8d): forward and adjoint evaluation alone, '
                         'L2 hot and L2 flushed, >= 100 iterations; 1 GPU')
    ap.add_argument('--watchdog', type=int, default=1500,
                    help='seconds after which a still-running bench process kills itself (a rank stuck in a '
                         'collective must not hold the GPUs)')
    ap.add_argument('--block', default=None, choices=['constant', 'attention'], help='default: the workload preset')
    ap.add_argument('--method', default=None, choices=['rk4', 'euler', 'midpoint', 'dopri5', 'adaptive_heun'],
                    help='solver of the forward and of the adjoint solve; default: the workload preset')
    ap.add_argument('--t-end', type=float, default=None, help='integration time; default: the workload preset')
    ap.add_argument('--tol-scale', type=float, default=None,
                    help='tol_scale / tol_scale_adjoint of the adaptive solvers (atol 1e-7, rtol 1e-9 times this); '
                         'default 1e3 for fp32 states')
    args = ap.parse_args()
    global FUNCTION, BLOCK, METHOD, T_END, TOL_SCALE
    preset = PRESETS[args.workload]
    FUNCTION = args.function or preset['function']
    BLOCK = args.block or (preset['block'] if FUNCTION == preset['function'] else
                           ('attention' if FUNCTION == 'transformer' else 'constant'))
    METHOD = args.method or preset['method']
    T_END = args.t_end if args.t_end is not None else preset['t_end']
    TOL_SCALE = args.tol_scale if args.tol_scale is not None else (1.0 if METHOD in ('rk4', 'euler', 'midpoint') else 1e3)
    if args.halo is None:
        # R-MAT slabs reference ~21 % of the remote rows at 8 ranks: the compact halo (exchange + per-rank problem over
        # [own slab | referenced rows]) is 2.3x - 2.8x the dense exchange there (profiles/r02z_c5_*_n8_compact.json);
        # the c4 shape references 97 %: whole slabs
        args.halo = 'compact' if args.workload == 'c5' else 'dense'
    if args.cuda_graph is None:
        args.cuda_graph = (args.workload in ('c1', 'c3') and METHOD in ('rk4', 'euler', 'midpoint') and args.gpus == 1
                           and int(os.environ.get('WORLD_SIZE', '1')) == 1 and not args.microbench
                           and args.impl == 'ours')
    if int(os.environ.get('WORLD_SIZE', '1')) > 1 and args.function is None and args.workload in ('c1', 'c2', 'c3'):
        pass                                           # the small shapes stay on one GPU; a partitioned run is by request
    if args.watchdog > 0:
        import threading

        def _expired():
            sys.stderr.write(f'bench.py: watchdog expired after {args.watchdog} s, exiting\n')
            sys.stderr.flush()
            os._exit(3)
        t = threading.Timer(args.watchdog, _expired)
        t.daemon = True
        t.start()
    if args.impl == 'reference':
        run_reference(args)
    elif args.microbench:
        run_microbench(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
