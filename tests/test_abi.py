"""CPU-only checks of the drop-in boundary: the shared library loads and exports exactly the
symbols include/grandb200.h declares, and the ctypes table covers all of them (no compute calls)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
HEADER = os.path.join(ROOT, 'include', 'grandb200.h')
LIB = os.path.join(ROOT, 'graph-neural-pde_b200', 'libgrandb200.so')


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(gb_[a-z0-9_]+)\s*\(', text)))


def test_header_declares_symbols():
    syms = declared_symbols()
    assert 'gb_spmm_fwd' in syms and 'gb_edge_softmax_fwd' in syms and len(syms) >= 30


@pytest.mark.skipif(not os.path.exists(LIB), reason='libgrandb200.so not built (run __graft_entry__.build())')
def test_library_exports_every_declared_symbol():
    out = subprocess.run(['nm', '-D', '--defined-only', LIB], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r' T (gb_[a-z0-9_]+)', out))
    missing = [s for s in declared_symbols() if s not in exported]
    assert not missing, f'declared in the header but not exported: {missing}'


@pytest.mark.skipif(not os.path.exists(LIB), reason='libgrandb200.so not built')
def test_ctypes_table_matches_header_and_loads():
    import grand_b200
    lib = grand_b200._lib.load()
    assert lib.gb_abi_version() == 2
    assert sorted(grand_b200._lib.SIGNATURES) == declared_symbols()
    assert lib.gb_prof_num_kernels() > 0 and lib.gb_prof_name(0) == b'spmm_fwd'


def test_cpu_tensors_are_rejected_loudly():
    """no CPU fallback: the operators refuse CPU tensors instead of computing something else"""
    import torch
    import grand_b200
    if not os.path.exists(LIB):
        pytest.skip('library not built')
    opt = {'hidden_dim': 4, 'block': 'constant', 'max_nfe': 10, 'no_alpha_sigmoid': False, 'add_source': False}
    f = grand_b200.LaplacianODEFunc(4, 4, opt, torch.device('cpu'))
    f.edge_index = torch.tensor([[0, 1], [1, 0]])
    f.edge_weight = torch.ones(2)
    with pytest.raises(RuntimeError, match='CUDA'):
        f(0.0, torch.zeros(2, 4))


def test_heads_must_divide_attention_dim():
    """test/test_gnn.py:61-65, function_transformer_attention.py:86-87"""
    import torch
    import grand_b200
    opt = {'leaky_relu_slope': 0.2, 'heads': 3, 'attention_dim': 32, 'attention_type': 'scaled_dot', 'beltrami': False}
    with pytest.raises(AssertionError):
        grand_b200.SpGraphTransAttentionLayer(6, 6, opt, torch.device('cpu'))


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: no module of the package may import it, and bench.py may do so only
    inside its CPU-baseline / reference-arm functions"""
    import ast
    import glob
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def oracle_imports(path):
        tree = ast.parse(open(path).read())
        hits = []
        for node in ast.walk(tree):
            if isinstance(node, ast.Import) and any(a.name.split('.')[0] == 'oracle' for a in node.names):
                hits.append(node.lineno)
            if isinstance(node, ast.ImportFrom) and (node.module or '').split('.')[0] == 'oracle':
                hits.append(node.lineno)
        return tree, hits

    for path in glob.glob(os.path.join(root, 'graph-neural-pde_b200', '*.py')) + [os.path.join(root, 'grand_b200.py')]:
        assert oracle_imports(path)[1] == [], f'{path} imports the oracle'
    tree, hits = oracle_imports(os.path.join(root, 'bench.py'))
    allowed = set()
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in ('cpu_pass', 'make_cpu_state', 'cpu_baseline', 'run_reference'):
            allowed.update(range(node.lineno, node.end_lineno + 1))
    assert all(h in allowed for h in hits), f'bench.py imports the oracle outside the CPU baseline: lines {hits}'
