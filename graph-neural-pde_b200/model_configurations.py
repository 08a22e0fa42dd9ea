"""Mirror of the reference's src/model_configurations.py:17-44: the string -> class registry behind
the `--block` and `--function` flags (src/graph_datasets/run_GNN.py:318-319).  Only the blocks and
functions on the accelerated path are registered; the others raise, naming the key."""
from .function_transformer_attention import ODEFuncTransformerAtt
from .function_laplacian_diffusion import LaplacianODEFunc
from .block_transformer_attention import AttODEblock
from .block_constant import ConstantODEblock


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
    elif ode_str in ('mixed', 'hard_attention', 'rewire_attention'):
        raise NotImplementedError(f"block '{ode_str}' is outside the accelerated path (SURVEY.md 2.1); "
                                  "its ODE function is served by LaplacianODEFunc unchanged")
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
