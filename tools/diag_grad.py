"""Diagnostic (not a test): kernel vs fp32 oracle vs fp64 oracle gradient errors per tensor."""
import sys
import torch

import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import data  # noqa: E402
import positions  # noqa: E402
from oracle import hexgnn_oracle as orc  # noqa: E402
from helpers import make_pair, loss_inputs  # noqa: E402

DEV = "cuda:0"


def run(hidden, layers, fixture, count, maker=1, path="auto"):
    o, m = make_pair(hidden, layers)
    o64 = orc.ModernTwoHeadedOracle(hidden, layers, 2).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = m.to(DEV)
    m.path = path
    b = data.collate(positions.load_positions(fixture, maker=maker)[:count])
    d = b.to(DEV)
    action, target, weight = loss_inputs(b)
    orc.rainbow_loss(o(b.x, b.edge_index, b.batch, b.ptr), b.ptr, action, target, weight).backward()
    orc.rainbow_loss(o64(b.x.double(), b.edge_index, b.batch, b.ptr), b.ptr, action, target.double(),
                     weight.double()).backward()
    q = m(d.x, d.edge_index, d.batch, d.ptr)
    orc.rainbow_loss(q, d.ptr, action.to(DEV), target.to(DEV), weight.to(DEV)).backward()
    r64 = dict(o64.named_parameters())
    r32 = dict(o.named_parameters())
    gmax = max(float(p.grad.abs().max()) for p in r64.values() if p.grad is not None)
    print(f"--- H={hidden} L={layers} {fixture} count={count} path={path} gmax={gmax:.3e}")
    for name, p in m.named_parameters():
        g64 = r64[name].grad
        if g64 is None:
            continue
        g32 = r32[name].grad.double()
        gk = p.grad.detach().cpu().double()
        scale = max(float(g64.abs().max()), 1e-3 * gmax)
        ek = (gk - g64).abs()
        e32 = (g32 - g64).abs()
        nbad = int((ek > 1e-5 * scale).sum())
        print(f"{name:44s} scale {scale:.2e} kern {float(ek.max())/scale:.2e} orc32 {float(e32.max())/scale:.2e} "
              f"n>tol {nbad}/{ek.numel()}")


if __name__ == "__main__":
    for path in ("simt", "auto"):
        run(35, 10, "pos_7x7_seed1234.npz", 256, path=path)
        run(110, 15, "pos_11x11_seed1234.npz", 48, path=path)
