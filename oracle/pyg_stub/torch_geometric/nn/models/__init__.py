from .basic_gnn import BasicGNN, GraphSAGE, PNA  # noqa: F401
