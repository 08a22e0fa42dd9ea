"""Import-only stand-in for the reference's `src/data_multi.py` (CLEVR ingestion: needs turtle,
h5py, skimage, torchvision and a network download).  `function_transformer_attention.py:4` imports
`get_dataset` from it but the hot path never calls it."""


def get_dataset(*args, **kwargs):
    raise RuntimeError("data_multi is stubbed: dataset ingestion is out of scope (SURVEY.md 2.1)")
