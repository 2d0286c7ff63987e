"""Message passing primitives (PyG 2.2.0 semantics, flow = source_to_target: messages travel
edge_index[0] -> edge_index[1]).  The reference's own restatement of SAGEConv.forward is
GN0/torch_script_models.py:52-73: `lin_l(propagate(x)) + lin_r(x)`."""
import torch
from torch.nn import Linear


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr="add", **kwargs):
        super().__init__()
        self.aggr = aggr

    def jittable(self, typing=None):
        return self

    def propagate(self, edge_index, x=None, size=None):
        """aggr over incoming edges of `message(x_j)`; a tuple x = (x_src, x_dst) uses x_src."""
        xs = x[0] if isinstance(x, (tuple, list)) else x
        src, dst = edge_index[0], edge_index[1]
        n = xs.size(0)
        msg = xs.index_select(0, src)
        out = xs.new_zeros((n,) + tuple(xs.shape[1:])).index_add_(0, dst, msg)
        if self.aggr == "mean":
            cnt = torch.bincount(dst, minlength=n).clamp_(min=1).to(xs.dtype)
            out = out / cnt.view(-1, *([1] * (xs.dim() - 1)))
        elif self.aggr not in ("add", "sum"):
            raise NotImplementedError(self.aggr)
        return out


class SAGEConv(MessagePassing):
    """PyG 2.2.0 `SAGEConv(in, out, aggr='mean', normalize=False, root_weight=True, project=False, bias=True)`:
    `lin_l` carries the bias, `lin_r` has none; both are created in this order (lin_l first)."""

    def __init__(self, in_channels, out_channels, aggr="mean", normalize=False, root_weight=True, project=False,
                 bias=True, **kwargs):
        super().__init__(aggr=aggr)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.normalize, self.root_weight, self.project = normalize, root_weight, project
        if isinstance(in_channels, int):
            in_channels = (in_channels, in_channels)
        if self.project:
            self.lin = Linear(in_channels[0], in_channels[0], bias=True)
        self.lin_l = Linear(in_channels[0], out_channels, bias=bias)
        if self.root_weight:
            self.lin_r = Linear(in_channels[1], out_channels, bias=False)

    def reset_parameters(self):
        self.lin_l.reset_parameters()
        if self.root_weight:
            self.lin_r.reset_parameters()

    def forward(self, x, edge_index, size=None):
        if torch.is_tensor(x):
            x = (x, x)
        out = self.lin_l(self.propagate(edge_index, x=x, size=size))
        if self.root_weight and x[1] is not None:
            out = out + self.lin_r(x[1])
        if self.normalize:
            out = torch.nn.functional.normalize(out, p=2.0, dim=-1)
        return out


class GCNConv(MessagePassing):      # named by the reference's imports only
    def __init__(self, *a, **k):
        raise NotImplementedError("pyg_stub: GCNConv is not on the pinned path")


class PNAConv(MessagePassing):      # named by the reference's imports only
    def __init__(self, *a, **k):
        raise NotImplementedError("pyg_stub: PNAConv is not on the pinned path")
