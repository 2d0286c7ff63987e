"""TEST-ONLY stand-in for torch_scatter.composite (see ../README.md)."""


def scatter_log_softmax(src, index, dim=0, eps=1e-12, dim_size=None):
    """torch_scatter/composite/softmax.py: x - max_g(x), then minus log(sum_g exp(.) + eps)."""
    from . import scatter_max, scatter_sum
    n = int(index.max()) + 1 if dim_size is None else int(dim_size)
    mx, _ = scatter_max(src, index, dim=0, dim_size=n)
    rec = src - mx.index_select(0, index)
    norm = scatter_sum(rec.exp(), index, dim=0, dim_size=n)
    return rec - (norm + eps).log().index_select(0, index)
