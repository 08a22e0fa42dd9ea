"""Mirror of the reference's src/model_configurations.py:17-44: the string -> class registry behind
the `--block` and `--function` flags (src/graph_datasets/run_GNN.py:318-319).  Blocks: constant, attention, mixed, hard_attention; functions: laplacian, transformer.  The others
(rewire_attention, GAT) raise, naming the key."""
from .function_transformer_attention import ODEFuncTransformerAtt
from .function_laplacian_diffusion import LaplacianODEFunc
from .block_transformer_attention import AttODEblock
from .block_constant import ConstantODEblock
from .block_mixed import MixedODEblock
from .block_transformer_hard_attention import HardAttODEblock


class BlockNotDefined(Exception):
    pass


class FunctionNotDefined(Exception):
    pass


def set_block(opt):
    ode_str = opt['block']
    if ode_str == 'attention':
        block = AttODEblock
    elif ode_str == 'constant':
        block = ConstantODEblock
    elif ode_str == 'mixed':
        block = MixedODEblock
    elif ode_str == 'hard_attention':
        block = HardAttODEblock
    elif ode_str == 'rewire_attention':
        raise NotImplementedError("block 'rewire_attention' (block_transformer_rewiring.py: kNN / GDC rewiring inside "
                                  'the block) is outside the accelerated path (SURVEY.md 2.1)')
    else:
        raise BlockNotDefined
    return block


def set_function(opt):
    ode_str = opt['function']
    if ode_str == 'laplacian':
        f = LaplacianODEFunc
    elif ode_str == 'transformer':
        f = ODEFuncTransformerAtt
    elif ode_str == 'GAT':
        raise NotImplementedError("function 'GAT' (function_GAT_attention.py) is outside the accelerated path")
    else:
        raise FunctionNotDefined
    return f
