"""`LayerNorm` of PyG 2.2.0 (torch_geometric/nn/norm/layer_norm.py): default mode "graph"; WITHOUT a batch vector
(the reference calls `norm(x)`, GN0/models.py:285) the statistics run over ALL nodes and channels of the input:
    x = x - x.mean();  out = x / (x.std(unbiased=False) + eps);  out = out * weight + bias."""
import torch


class LayerNorm(torch.nn.Module):
    def __init__(self, in_channels, eps=1e-5, affine=True, mode="graph"):
        super().__init__()
        self.in_channels, self.eps, self.mode = in_channels, eps, mode
        if affine:
            self.weight = torch.nn.Parameter(torch.empty(in_channels))
            self.bias = torch.nn.Parameter(torch.empty(in_channels))
        else:
            self.register_parameter("weight", None)
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        if self.weight is not None:
            torch.nn.init.ones_(self.weight)
            torch.nn.init.zeros_(self.bias)

    def forward(self, x, batch=None):
        if self.mode == "graph":
            if batch is None:
                x = x - x.mean()
                out = x / (x.std(unbiased=False) + self.eps)
            else:
                from torch_scatter import scatter
                n = int(batch.max()) + 1
                norm = torch.bincount(batch, minlength=n).clamp_(min=1).to(x.dtype).mul_(x.size(-1)).view(-1, 1)
                mean = scatter(x, batch, dim=0, dim_size=n, reduce="sum").sum(dim=-1, keepdim=True) / norm
                x = x - mean.index_select(0, batch)
                var = scatter(x * x, batch, dim=0, dim_size=n, reduce="sum").sum(dim=-1, keepdim=True) / norm
                out = x / (var + self.eps).sqrt().index_select(0, batch)
            if self.weight is not None and self.bias is not None:
                out = out * self.weight + self.bias
            return out
        if self.mode == "node":
            return torch.nn.functional.layer_norm(x, (self.in_channels,), self.weight, self.bias, self.eps)
        raise ValueError(self.mode)


class GraphNorm(torch.nn.Module):   # subclassed by GN0/models.py:644 (CachedGraphNorm); not on the pinned path
    def __init__(self, in_channels, eps=1e-5):
        super().__init__()
        self.in_channels, self.eps = in_channels, eps
        self.weight = torch.nn.Parameter(torch.ones(in_channels))
        self.bias = torch.nn.Parameter(torch.zeros(in_channels))
        self.mean_scale = torch.nn.Parameter(torch.ones(in_channels))
