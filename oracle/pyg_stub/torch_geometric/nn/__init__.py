from .conv import MessagePassing, SAGEConv, GCNConv, PNAConv  # noqa: F401
from . import norm, models, resolver, aggr  # noqa: F401
