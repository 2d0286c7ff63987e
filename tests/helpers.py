"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
import torch

from gnn_hex_b200 import ModernTwoHeaded, data
from oracle import hexgnn_oracle as orc

REL_TOL = 1e-5  # BASELINE.json north_star: 1e-5 relative (fp32) on Q-values and gradients


def make_pair(hidden, layers, head_layers=2, seed=0, dtype=torch.float32):
    """Oracle + product module with identical weights (default torch init, seeded)."""
    torch.manual_seed(seed)
    o = orc.ModernTwoHeadedOracle(hidden, layers, head_layers)
    m = ModernTwoHeaded(hidden, layers, head_layers)
    m.load_state_dict(o.state_dict())
    if dtype != torch.float32:
        o = o.to(dtype)
    return o, m


def loss_inputs(batch, seed=1):
    """a_g ~ randint(2, n_g), t_g ~ U(-1,1), w_g ~ U(0.5,1)  (SURVEY.md §8d config 1)."""
    g = torch.Generator().manual_seed(seed)
    n = (batch.ptr[1:] - batch.ptr[:-1])
    u = torch.rand(batch.num_graphs, generator=g)
    action = (2 + (u * (n - 2).clamp(min=1)).long()).clamp(max=n - 1).clamp(min=0)
    target = torch.rand(batch.num_graphs, generator=g) * 2 - 1
    weight = torch.rand(batch.num_graphs, generator=g) * 0.5 + 0.5
    return action, target, weight


def rel_err(a, b):
    """max|a-b| / max|b| (b = reference)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    denom = b.abs().max().clamp(min=1e-30)
    return float((a - b).abs().max() / denom)


def grad_errs(model, oracle):
    """Per-tensor max|d| / max(|ref tensor|_max, 1e-3*|all ref grads|_max)."""
    ref = dict(oracle.named_parameters())
    gmax = max(float(p.grad.abs().max()) for p in ref.values() if p.grad is not None)
    worst, name_worst = 0.0, None
    for name, p in model.named_parameters():
        rg = ref[name].grad
        rg = torch.zeros_like(ref[name]) if rg is None else rg
        g = torch.zeros_like(rg) if p.grad is None else p.grad.detach().cpu().to(rg.dtype)
        scale = max(float(rg.abs().max()), 1e-3 * gmax)
        e = float((g.double() - rg.double()).abs().max()) / scale
        if e > worst:
            worst, name_worst = e, name
    return worst, name_worst
