"""Pure-torch stand-in for the three torch_scatter entry points the reference calls
(`src/utils.py:9,122,125,138,191,230`).  TEST INFRASTRUCTURE ONLY: it exists so that
`oracle/make_golden.py` can import the reference's own modules in a container where
torch_scatter (pinned 2.0.5/2.0.6 by the reference) cannot be installed.  Semantics follow the
torch_scatter documentation: out[index[i]] (+)= src[i] along `dim`, `dim_size` rows, empty
groups are 0 for both add and max."""
import torch


def _expand_index(index, src, dim):
    if dim < 0:
        dim = src.dim() + dim
    if index.dim() == src.dim():
        return index.expand_as(src), dim
    shape = [1] * src.dim()
    # index covers the leading dims up to and including `dim` (torch_scatter broadcasting rule)
    for i in range(index.dim()):
        shape[i] = index.shape[i]
    return index.reshape(shape).expand_as(src), dim


def scatter_add(src, index, dim=-1, out=None, dim_size=None):
    idx, dim = _expand_index(index, src, dim)
    if dim_size is None:
        dim_size = int(index.max()) + 1
    shape = list(src.shape)
    shape[dim] = dim_size
    res = torch.zeros(shape, dtype=src.dtype, device=src.device)
    return res.scatter_add(dim, idx, src)


def scatter_max(src, index, dim=-1, out=None, dim_size=None):
    idx, dim = _expand_index(index, src, dim)
    if dim_size is None:
        dim_size = int(index.max()) + 1
    shape = list(src.shape)
    shape[dim] = dim_size
    res = torch.zeros(shape, dtype=src.dtype, device=src.device)
    res = res.scatter_reduce(dim, idx, src, reduce='amax', include_self=False)
    return res, None


def scatter(src, index, dim=-1, out=None, dim_size=None, reduce='sum'):
    if reduce in ('sum', 'add'):
        return scatter_add(src, index, dim, out, dim_size)
    if reduce == 'max':
        return scatter_max(src, index, dim, out, dim_size)[0]
    raise NotImplementedError(reduce)
