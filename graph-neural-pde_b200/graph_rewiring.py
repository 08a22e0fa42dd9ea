"""Positional-encoding files of the beltrami configurations - mirror of `apply_beltrami`
(src/graph_rewiring.py:277-313), the only part of that module on the data path of the hot path:
the per-node encodings p [N, P] that `GNN.forward` projects with `mp` and concatenates to the features.

    <data_dir>/pos_encodings/<dataset>_<pos_enc_type>.pkl      pickle of the [N, P] encoding;
                                                               DW* types: a dict whose 'data' entry is the encoding

Same file names, same pickle payloads, same `pos_enc_csv` side file as the reference, so encodings cached by the
reference are read unchanged and vice versa.  COMPUTING a missing GDC encoding needs torch_geometric.transforms.GDC
(`apply_gdc`, :52-81), which is not installable here: a missing file raises and names the path.  The rewiring
heuristics of the module (KNN, edge sampling, GDC sparsification) are out of scope (SURVEY.md 8 "out").
"""
import os
import pickle

import numpy as np
import torch

ROOT_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def pos_encoding_path(opt, data_dir=f'{ROOT_DIR}/data'):
    return os.path.join(f'{data_dir}', 'pos_encodings', f"{opt['dataset']}_{opt['pos_enc_type']}.pkl")


def save_pos_encoding(pos_encoding, opt, data_dir=f'{ROOT_DIR}/data'):
    """what the reference does after computing an encoding (:301-311): optional sparse CSV (row, col, value), pickle"""
    fname = pos_encoding_path(opt, data_dir)
    os.makedirs(os.path.dirname(fname), exist_ok=True)
    if opt.get('pos_enc_csv', False):
        sp = torch.as_tensor(pos_encoding).detach().cpu().to_sparse()
        table_mat = np.concatenate([sp.indices().numpy(), np.atleast_2d(sp.values().numpy())], axis=0).T
        np.savetxt(f'{fname[:-4]}.csv', table_mat, delimiter=',')
    payload = {'data': pos_encoding} if opt['pos_enc_type'].startswith('DW') else pos_encoding
    tmp = f'{fname}.tmp{os.getpid()}'
    with open(tmp, 'wb') as f:
        pickle.dump(payload, f)
    os.replace(tmp, fname)
    return fname


def apply_beltrami(data, opt, data_dir=f'{ROOT_DIR}/data'):
    """the cached positional encodings of opt['dataset'] / opt['pos_enc_type'] as a float tensor [N, P] on the CPU
    (the caller moves it to the device, as upstream's run_GNN does).  `data` is only used to check the node count."""
    fname = pos_encoding_path(opt, data_dir)
    if not os.path.exists(fname):
        if opt['pos_enc_type'] == 'GDC':
            raise NotImplementedError(f'{fname} does not exist and computing GDC encodings needs '
                                      'torch_geometric.transforms.GDC (graph_rewiring.py:52-81), which is not '
                                      'available: generate the file with the reference (or save_pos_encoding)')
        raise FileNotFoundError(f"[x] positional encodings {fname} not found "
                                f"(pos_enc_type {opt['pos_enc_type']!r} cannot be computed: graph_rewiring.py:299-300)")
    with open(fname, 'rb') as f:
        pos_encoding = pickle.load(f)
    if opt['pos_enc_type'].startswith('DW'):
        pos_encoding = pos_encoding['data']
    if isinstance(pos_encoding, np.ndarray):
        pos_encoding = torch.from_numpy(pos_encoding)
    if not torch.is_tensor(pos_encoding):
        raise TypeError(f'{fname}: expected a tensor / ndarray payload, got {type(pos_encoding).__name__}')
    if pos_encoding.is_sparse:
        pos_encoding = pos_encoding.to_dense()
    n = getattr(data, 'num_nodes', None)
    if n is not None and pos_encoding.shape[0] != int(n):
        raise ValueError(f'{fname}: {pos_encoding.shape[0]} rows for a graph of {int(n)} nodes')
    return pos_encoding.to(torch.float32)
