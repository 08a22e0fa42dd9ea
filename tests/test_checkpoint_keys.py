"""CPU: checkpoints of the reference load into the repo's models with strict=True.  tests/golden/ref_state_dict_keys.json
holds the names and shapes of `GNN(...).state_dict()` as the reference's own classes build it (oracle/make_state_keys.py,
run where /root/reference exists): `odeblock.odefunc.*`, `odeblock.reg_odefunc.odefunc.*` (the first function instance
kept by RegularizedODEfunc, base_classes.py:40-43), the block-level attention layer, encoder / decoder."""
import json
import os

import pytest
import torch

GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'ref_state_dict_keys.json')
OPT = {'hidden_dim': 8, 'max_nfe': 1000, 'no_alpha_sigmoid': False, 'add_source': True, 'augment': False,
       'adjoint': False, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0, 'data_norm': 'rw', 'self_loop_weight': 1.0,
       'method': 'rk4', 'step_size': 1.0, 'max_iters': 100, 'adjoint_method': 'rk4', 'adjoint_step_size': 1.0,
       'leaky_relu_slope': 0.2, 'heads': 2, 'attention_dim': 8, 'attention_norm_idx': 0, 'reweight_attention': False,
       'beltrami': False, 'attention_type': 'scaled_dot', 'square_plus': False, 'mix_features': False,
       'multi_modal': False, 'kinetic_energy': None, 'jacobian_norm2': None, 'total_deriv': None,
       'directional_penalty': None, 'use_labels': False, 'use_mlp': False, 'batch_norm': False, 'fc_out': False,
       'dropout': 0.0, 'input_dropout': 0.0, 'time': 2.0}


@pytest.mark.parametrize('case', ['laplacian/constant', 'transformer/attention', 'laplacian/attention'])
def test_reference_named_state_dict_loads_strictly(case):
    import grand_b200
    ref = json.load(open(GOLD))[case]
    function, block = case.split('/')
    model = grand_b200.GNN(dict(OPT, function=function, block=block), 3, 7, device=torch.device('cpu'))
    own = {k: list(v.shape) for k, v in model.state_dict().items()}
    assert own == ref, (sorted(set(ref) - set(own)), sorted(set(own) - set(ref)))
    gen = torch.Generator().manual_seed(0)
    sd = {k: torch.randn(shape, generator=gen) for k, shape in ref.items()}          # a checkpoint named like the reference's
    assert not model.load_state_dict(sd, strict=True).missing_keys
    for k, v in model.state_dict().items():
        assert torch.equal(v, sd[k]), k
    # same number of parameters as the reference model (its ray-tune (model, optimizer) checkpoints index them)
    assert sum(1 for _ in model.parameters()) == sum(1 for k in ref if 'running_' not in k and 'num_batches' not in k)


def test_getnfe_counts_each_evaluation_once():
    """base_classes.py:174-179: odefunc.nfe + reg_odefunc.odefunc.nfe over two different function instances"""
    import grand_b200
    model = grand_b200.GNN(dict(OPT, function='laplacian', block='constant'), 3, 7, device=torch.device('cpu'))
    blk = model.odeblock
    assert blk.reg_odefunc.odefunc is not blk.odefunc and isinstance(blk.reg_odefunc, torch.nn.Module)
    blk.odefunc.nfe = 5
    assert model.getNFE() == 5
    model.resetNFE()
    assert blk.odefunc.nfe == 0 and blk.reg_odefunc.odefunc.nfe == 0
