"""Seeded synthetic graphs of the BASELINE.json shapes, generated on the GPU with torch ops (bench and
test plumbing).  Same distributions as oracle/graph.py (`shaped_graph`, `rmat_graph`);
not bit-identical to the numpy versions (different RNG), which is stated wherever it matters.

CONFIGS maps the names used by bench.py / DESIGN.md to (N, E, d, heads, attention_dim)."""
import torch

CONFIGS = {
    # name: nodes, directed edges (before self loops), state width d, heads, attention_dim
    'c1': dict(n=2708, e=10556, d=64, heads=4, adim=64, desc='Cora-shaped'),
    'c2': dict(n=19717, e=88648, d=128, heads=4, adim=64, desc='PubMed-shaped'),
    'c3': dict(n=169343, e=1166243 - 1, d=96, heads=4, adim=64, desc='ogbn-arxiv-shaped (BLEND: 64 feat + 32 pos)'),
    'c4': dict(n=2449029, e=61859140, d=128, heads=4, adim=64, desc='ogbn-products-shaped'),
    'c5': dict(n=1 << 24, e=(1 << 24) * 16, d=64, heads=4, adim=64, desc='RMAT scale 24, edge factor 16'),
}


def shaped_graph(n, e, seed, device):
    """exactly n nodes and e (even) directed edges: e/2 distinct unordered pairs u != v, u uniform,
    v ~ (rank+1)^-0.5 under a fixed random permutation, symmetrised.  int64 [2, e] on `device`."""
    assert e % 2 == 0
    g = torch.Generator(device=device).manual_seed(seed)
    half = e // 2
    perm = torch.randperm(n, generator=g, device=device)
    keys = torch.empty(0, dtype=torch.int64, device=device)
    while keys.numel() < half:
        need = int((half - keys.numel()) * 1.25) + 1024
        u = torch.randint(0, n, (need,), generator=g, device=device)
        r = (n * torch.rand(need, generator=g, device=device, dtype=torch.float64) ** 2).long().clamp_(max=n - 1)
        v = perm[r]
        keep = u != v
        lo, hi = torch.minimum(u, v)[keep], torch.maximum(u, v)[keep]
        keys = torch.unique(torch.cat([keys, lo * n + hi]))
    if keys.numel() > half:
        keys = keys[torch.randperm(keys.numel(), generator=g, device=device)[:half]]
    lo, hi = keys // n, keys % n
    return torch.stack([torch.cat([lo, hi]), torch.cat([hi, lo])])


def rmat_graph(scale, edge_factor, seed, device, a=0.57, b=0.19, c=0.19, chunk=1 << 26):
    """Graph500 R-MAT (a,b,c,d) = (0.57,0.19,0.19,0.05), vertex ids permuted, duplicates and self
    loops kept.  int64 [2, E]."""
    g = torch.Generator(device=device).manual_seed(seed)
    n = 1 << scale
    e = n * edge_factor
    perm = torch.randperm(n, generator=g, device=device)
    ab, c_norm, a_norm = a + b, c / (1.0 - (a + b)), a / (a + b)
    rows, cols = [], []
    for s in range(0, e, chunk):
        m = min(chunk, e - s)
        row = torch.zeros(m, dtype=torch.int64, device=device)
        col = torch.zeros(m, dtype=torch.int64, device=device)
        for _ in range(scale):
            i_bit = torch.rand(m, generator=g, device=device) > ab
            thr = torch.where(i_bit, torch.full((), c_norm, device=device), torch.full((), a_norm, device=device))
            j_bit = torch.rand(m, generator=g, device=device) > thr
            row = (row << 1) | i_bit
            col = (col << 1) | j_bit
        rows.append(perm[row])
        cols.append(perm[col])
    return torch.stack([torch.cat(rows), torch.cat(cols)])


def make_graph(name, device, seed=1234, scale_div=1):
    """edge_index of config `name`; scale_div > 1 shrinks N and E by that factor (CPU baseline samples)"""
    cfg = CONFIGS[name]
    if name == 'c5':
        import math
        scale = 24 - int(round(math.log2(scale_div)))
        return rmat_graph(scale, 16, seed, device), 1 << scale
    n = cfg['n'] // scale_div
    e = (cfg['e'] // scale_div) // 2 * 2
    return shaped_graph(n, e, seed, device), n
