"""grand_b200 - B200-native drop-in for the per-step graph-diffusion operator of
alimt1992/graph-neural-pde (GRAND / BLEND).  Import as `import grand_b200` (repo-root shim) - the
directory name `graph-neural-pde_b200` is not a Python identifier.

Module names mirror the reference's `src/` files so a maintainer can swap imports one for one:
  base_classes, function_laplacian_diffusion, function_transformer_attention, block_constant,
  block_transformer_attention, block_mixed, block_transformer_hard_attention, model_configurations, utils, GNN, GNN_early,
  early_stop_solver, CGNN (model classes), graph_rewiring (positional-encoding files only)  (+ integrators: torchdiffeq-compatible
  odeint / odeint_adjoint used when torchdiffeq is not installed).
All device work goes through libgrandb200.so (csrc/, C ABI in include/grandb200.h).
"""
from . import _lib
from .plan import GraphPlan, get_plan
from .utils import MaxNFEException, get_rw_adj, gcn_norm_fill_val, add_remaining_self_loops, Meter
from .base_classes import ODEFunc, ODEblock, GraphData
from .function_laplacian_diffusion import LaplacianODEFunc
from .function_transformer_attention import ODEFuncTransformerAtt, SpGraphTransAttentionLayer
from .function_GAT_attention import ODEFuncAtt, SpGraphAttentionLayer
from .block_constant import ConstantODEblock
from .block_transformer_attention import AttODEblock
from .block_mixed import MixedODEblock
from .block_transformer_hard_attention import HardAttODEblock
from .model_configurations import set_block, set_function
from . import integrators
from . import partition
from .graphed import GraphedPass
from .partition import RowPartition, Halo, LoopbackHalo, SparseHalo, CompactHalo, PushHalo, CopyHalo, CompactPushHalo
from .integrators import odeint, odeint_adjoint
from .early_stop_solver import EarlyStopInt
from .base_classes import BaseGNN
from .GNN import GNN
from .GNN_early import GNNEarly
from . import CGNN, graph_rewiring, ops

__all__ = ['GraphPlan', 'get_plan', 'MaxNFEException', 'get_rw_adj', 'gcn_norm_fill_val',
           'add_remaining_self_loops', 'Meter', 'ODEFunc', 'ODEblock', 'GraphData', 'LaplacianODEFunc',
           'ODEFuncTransformerAtt', 'SpGraphTransAttentionLayer', 'ODEFuncAtt', 'SpGraphAttentionLayer', 'ConstantODEblock', 'AttODEblock', 'MixedODEblock', 'HardAttODEblock',
           'set_block', 'set_function', 'partition', 'GraphedPass', 'RowPartition', 'Halo', 'LoopbackHalo', 'SparseHalo', 'CompactHalo', 'PushHalo', 'CopyHalo', 'CompactPushHalo', 'odeint', 'odeint_adjoint', 'integrators', 'EarlyStopInt', 'BaseGNN', 'GNN', 'GNNEarly']
