"""TEST-ONLY stand-in for torch_scatter 2.1.0 (see ../README.md): `scatter`, `scatter_mean`, `scatter_max`,
`scatter_min` along dim 0 with the library's documented semantics -- empty groups give 0 for every reduce,
`mean` divides by max(count, 1), `max` / `min` return (values, arg) with arg = the FIRST row attaining the
extremum (csrc/cpu/scatter_cpu.cpp scans rows in order and updates on strict > / <) and arg = src.size(0) for
empty groups.  Written with plain torch index ops, independently of oracle/hexgnn_oracle.py."""
from typing import Optional, Tuple

import torch

from . import composite  # noqa: F401


def _dim_size(index, dim_size):
    if dim_size is not None:
        return int(dim_size)
    return int(index.max()) + 1 if index.numel() else 0


def _expand(index, src):
    return index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)


def scatter_sum(src, index, dim=0, out=None, dim_size=None):
    assert dim in (0, -src.dim()) and out is None
    res = src.new_zeros((_dim_size(index, dim_size),) + tuple(src.shape[1:]))
    return res.index_add_(0, index, src)


scatter_add = scatter_sum


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    n = _dim_size(index, dim_size)
    total = scatter_sum(src, index, dim, out, n)
    count = torch.bincount(index, minlength=n).clamp_(min=1).to(src.dtype)
    return total / count.view(-1, *([1] * (src.dim() - 1)))


def _extremum(src, index, dim_size, largest) -> Tuple[torch.Tensor, torch.Tensor]:
    n = _dim_size(index, dim_size)
    idx = _expand(index, src)
    val = src.detach().new_zeros((n,) + tuple(src.shape[1:]))
    val = val.scatter_reduce(0, idx, src.detach(), reduce="amax" if largest else "amin", include_self=False)
    rows = _expand(torch.arange(src.shape[0], device=src.device), src)
    first = torch.where(src.detach() == val.index_select(0, index), rows, torch.full_like(rows, src.shape[0]))
    arg = torch.full(val.shape, src.shape[0], dtype=torch.long, device=src.device)
    arg = arg.scatter_reduce(0, idx, first, reduce="amin", include_self=True)
    # value = src at the arg row (0 for empty groups); gradient: the whole of it to the arg row
    # (torch_scatter/scatter.py: grad_in.scatter_(0, arg, grad_out))
    padded = torch.cat([src, src.new_zeros((1,) + tuple(src.shape[1:]))], 0)
    return padded.gather(0, arg), arg


def scatter_max(src, index, dim=0, out=None, dim_size=None):
    return _extremum(src, index, dim_size, True)


def scatter_min(src, index, dim=0, out=None, dim_size=None):
    return _extremum(src, index, dim_size, False)


def scatter(src, index, dim: int = 0, out: Optional[torch.Tensor] = None, dim_size: Optional[int] = None,
            reduce: str = "sum"):
    if reduce in ("sum", "add"):
        return scatter_sum(src, index, dim, out, dim_size)
    if reduce == "mean":
        return scatter_mean(src, index, dim, out, dim_size)
    if reduce == "max":
        return scatter_max(src, index, dim, out, dim_size)[0]
    if reduce == "min":
        return scatter_min(src, index, dim, out, dim_size)[0]
    raise ValueError(reduce)
