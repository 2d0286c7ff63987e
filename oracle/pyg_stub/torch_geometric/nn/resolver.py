import torch


def activation_resolver(query="relu", *args, **kwargs):
    if query is None or isinstance(query, torch.nn.Module):
        return query
    table = {"relu": torch.nn.ReLU, "tanh": torch.nn.Tanh, "elu": torch.nn.ELU, "leaky_relu": torch.nn.LeakyReLU}
    return table[query.lower()](*args, **kwargs)


def normalization_resolver(query, *args, **kwargs):
    if isinstance(query, torch.nn.Module):
        return query
    from .norm import GraphNorm, LayerNorm
    table = {"layer_norm": LayerNorm, "layernorm": LayerNorm, "layer": LayerNorm, "graph_norm": GraphNorm,
             "graphnorm": GraphNorm, "graph": GraphNorm}
    return table[str(query).lower()](*args, **kwargs)
