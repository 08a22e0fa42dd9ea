"""Write tests/golden/ref_state_dict_keys.json: the names and shapes of `GNN(opt, dataset).state_dict()` as the
REFERENCE's own classes build it (TEST INFRASTRUCTURE ONLY; run in the build container, where /root/reference exists).
The reference modules are imported unmodified behind the import-only stubs of oracle/ref_stubs plus an empty
`torchdiffeq` module (its integrators are only looked up, never called, while a model is constructed).
tests/test_checkpoint_keys.py loads a state_dict with exactly these keys into the repo's GNN with strict=True."""
import json
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = '/root/reference/src'


OPT = {'hidden_dim': 8, 'max_nfe': 1000, 'no_alpha_sigmoid': False, 'add_source': True, 'augment': False,
       'adjoint': False, 'tol_scale': 1.0, 'tol_scale_adjoint': 1.0, 'data_norm': 'rw', 'self_loop_weight': 1.0,
       'method': 'rk4', 'step_size': 1.0, 'max_iters': 100, 'adjoint_method': 'rk4', 'adjoint_step_size': 1.0,
       'leaky_relu_slope': 0.2, 'heads': 2, 'attention_dim': 8, 'attention_norm_idx': 0, 'reweight_attention': False,
       'beltrami': False, 'attention_type': 'scaled_dot', 'square_plus': False, 'mix_features': False,
       'multi_modal': False, 'kinetic_energy': None, 'jacobian_norm2': None, 'total_deriv': None,
       'directional_penalty': None, 'use_labels': False, 'use_mlp': False, 'batch_norm': False, 'fc_out': False,
       'dropout': 0.0, 'input_dropout': 0.0, 'time': 2.0, 'alpha_dim': 'sc', 'beta_dim': 'sc', 'dataset': 'Cora',
       'feat_hidden_dim': 6, 'pos_enc_hidden_dim': 2, 'pos_enc_type': 'GDC', 'edge_sampling': False}


def main():
    for p in (REF_SRC, os.path.join(HERE, 'ref_stubs')):
        if p not in sys.path:
            sys.path.insert(0, p)
    td = types.ModuleType('torchdiffeq')
    td.odeint = td.odeint_adjoint = lambda *a, **k: (_ for _ in ()).throw(RuntimeError('torchdiffeq is not installed'))
    sys.modules.setdefault('torchdiffeq', td)
    import GNN as ref_gnn                                   # noqa: E402  (the reference's own GNN.py)
    out = {}
    for function, block in (('laplacian', 'constant'), ('transformer', 'attention'), ('laplacian', 'attention')):
        opt = dict(OPT, function=function, block=block)
        model = ref_gnn.GNN(opt, 3, 7, device=torch.device('cpu'))
        out[f'{function}/{block}'] = {k: list(v.shape) for k, v in model.state_dict().items()}
    path = os.path.join(HERE, '..', 'tests', 'golden', 'ref_state_dict_keys.json')
    with open(path, 'w') as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    for k, v in out.items():
        print(k, len(v), 'keys')


if __name__ == '__main__':
    main()
