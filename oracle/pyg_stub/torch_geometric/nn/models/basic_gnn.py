"""`BasicGNN.__init__` of PyG 2.2.0 (torch_geometric/nn/models/basic_gnn.py): the layer and norm LISTS the
reference's `cachify_gnn` subclass walks (GN0/models.py:144-294).  The reference overrides `forward`, so only the
constructor, `init_conv` and `reset_parameters` matter here."""
import copy

import torch
from torch.nn import ModuleList

from ..conv import SAGEConv
from ..resolver import activation_resolver, normalization_resolver


class BasicGNN(torch.nn.Module):
    def __init__(self, in_channels, hidden_channels, num_layers, out_channels=None, dropout=0.0, act="relu",
                 act_first=False, act_kwargs=None, norm=None, norm_kwargs=None, jk=None, **kwargs):
        super().__init__()
        self.in_channels, self.hidden_channels, self.num_layers = in_channels, hidden_channels, num_layers
        self.dropout = dropout
        self.act = activation_resolver(act, **(act_kwargs or {}))
        self.jk_mode = jk
        self.act_first = act_first
        self.norm = norm if isinstance(norm, str) else None
        self.norm_kwargs = norm_kwargs
        self.out_channels = out_channels if out_channels is not None else hidden_channels

        self.convs = ModuleList()
        if num_layers > 1:
            self.convs.append(self.init_conv(in_channels, hidden_channels, **kwargs))
            in_channels = hidden_channels
        for _ in range(num_layers - 2):
            self.convs.append(self.init_conv(in_channels, hidden_channels, **kwargs))
            in_channels = hidden_channels
        if out_channels is not None and jk is None:
            self._is_conv_to_out = True
            self.convs.append(self.init_conv(in_channels, out_channels, **kwargs))
        else:
            self.convs.append(self.init_conv(in_channels, hidden_channels, **kwargs))

        self.norms = None
        if norm is not None:
            norm_layer = normalization_resolver(norm, hidden_channels, **(norm_kwargs or {}))
            self.norms = ModuleList()
            for _ in range(num_layers - 1):
                self.norms.append(copy.deepcopy(norm_layer))
            if jk is not None:
                self.norms.append(copy.deepcopy(norm_layer))
        if jk is not None:
            raise NotImplementedError("pyg_stub: jumping knowledge is not on the pinned path")

    def init_conv(self, in_channels, out_channels, **kwargs):
        raise NotImplementedError

    def reset_parameters(self):
        for conv in self.convs:
            conv.reset_parameters()
        for norm in self.norms or []:
            norm.reset_parameters()


class GraphSAGE(BasicGNN):
    supports_edge_weight = False
    supports_edge_attr = False

    def init_conv(self, in_channels, out_channels, **kwargs):
        return SAGEConv(in_channels, out_channels, **kwargs)


class PNA(BasicGNN):                # named by the reference's imports only
    def init_conv(self, in_channels, out_channels, **kwargs):
        raise NotImplementedError("pyg_stub: PNA is not on the pinned path")
