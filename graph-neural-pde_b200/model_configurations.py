"""The string -> class registry behind the `--block` and `--function` flags: same two entry points and the same two
exception types as the reference's src/model_configurations.py:17-44 (callers: src/GNN.py:11-13,
src/graph_datasets/run_GNN.py:318-319), kept as lookup tables.

Blocks: constant, attention, mixed, hard_attention; functions: laplacian, transformer, GAT (opt-in until its
kernel has run on hardware, see function_GAT_attention.py).  Keys the reference knows but
this package does not accelerate raise NotImplementedError naming the key; unknown keys raise the reference's
BlockNotDefined / FunctionNotDefined."""
from .function_transformer_attention import ODEFuncTransformerAtt
from .function_laplacian_diffusion import LaplacianODEFunc
from .function_GAT_attention import ODEFuncAtt
from .block_transformer_attention import AttODEblock
from .block_constant import ConstantODEblock
from .block_mixed import MixedODEblock
from .block_transformer_hard_attention import HardAttODEblock


class BlockNotDefined(Exception):
    pass


class FunctionNotDefined(Exception):
    pass


BLOCKS = {'constant': ConstantODEblock, 'attention': AttODEblock, 'mixed': MixedODEblock,
          'hard_attention': HardAttODEblock}
FUNCTIONS = {'laplacian': LaplacianODEFunc, 'transformer': ODEFuncTransformerAtt, 'GAT': ODEFuncAtt}
NOT_ACCELERATED = {
    'rewire_attention': "block 'rewire_attention' (block_transformer_rewiring.py: kNN / GDC rewiring inside the block) "
                        'is outside the accelerated path (SURVEY.md 2.1)',
}


def _lookup(table, key, missing):
    if key in table:
        return table[key]
    if key in NOT_ACCELERATED:
        raise NotImplementedError(NOT_ACCELERATED[key])
    raise missing


def set_block(opt):
    return _lookup(BLOCKS, opt['block'], BlockNotDefined)


def set_function(opt):
    return _lookup(FUNCTIONS, opt['function'], FunctionNotDefined)
