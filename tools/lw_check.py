"""Dev check of the layer-at-a-time path (hexgnn_forward_lw / hexgnn_backward_lw) against the oracle; run on a GPU box."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import helpers  # noqa: E402
import positions  # noqa: E402
from gnn_hex_b200 import ModernTwoHeaded, data, ops  # noqa: E402
from oracle import hexgnn_oracle as orc  # noqa: E402


def norm_block(o, H, L, Lh):
    parts = [torch.cat([o.gnn.norms[s].weight, o.gnn.norms[s].bias]) for s in range(L)]
    parts.append(torch.cat([o.after_embed_norm.weight, o.after_embed_norm.bias]))
    for head in (o.maker_head, o.breaker_head):
        parts += [torch.cat([head.gnn.norms[j].weight, head.gnn.norms[j].bias]) for j in range(Lh)]
    return torch.cat(parts).detach().float().contiguous()


def run(tag, H, L, Lh, graphs, norm, maker, gain=1.5):
    dev = torch.device("cuda")
    o = orc.ModernTwoHeadedOracle(H, L, Lh, norm=norm)
    helpers.set_hash_weights(o, seed=3, gain=gain)
    m = ModernTwoHeaded(H, L, Lh)
    m.load_state_dict({k: v for k, v in o.state_dict().items() if "norm" not in k})
    m = m.to(dev)
    for g in graphs:
        g.x[:, 2] = maker
    b = data.collate(graphs)
    x, ei, bt, ptr = b.x.to(dev), b.edge_index.to(dev), b.batch.to(dev), b.ptr.to(dev)
    csr = m.build_csr(ei, bt, ptr, num_nodes=x.shape[0])
    npar = norm_block(o, H, L, Lh).to(dev) if norm else None
    flat = m.flat_parameters
    for mode, flags in (("q", 0), ("adv", ops.HEXGNN_ADV_ONLY), ("sep", ops.HEXGNN_SEPARATE)):
        out, val, _, _ = ops.lw_forward(m._desc, flat, npar, x, csr, flags)
        torch.cuda.synchronize()
        with torch.no_grad():
            ref = o(b.x, b.edge_index, b.batch, b.ptr, advantages_only=(mode == "adv"), seperate=(mode == "sep"))
        if mode == "sep":
            e = max(helpers.rel_err(val, ref[0]), helpers.rel_err(out, ref[1]))
        else:
            e = helpers.rel_err(out, ref.reshape(-1))
        print(f"{tag} norm={int(norm)} maker={maker} {mode}: rel err {e:.3e}  status {int(csr.status.item())}", flush=True)
    t0 = time.time()
    for _ in range(20):
        ops.lw_forward(m._desc, flat, npar, x, csr, 0)
    torch.cuda.synchronize()
    print(f"   lw forward: {(time.time() - t0) / 20 * 1e3:.3f} ms", flush=True)


def main_fwd():
    play7 = positions.load_positions("pos_7x7_seed1234.npz")
    play11 = positions.load_positions("pos_11x11_seed1234.npz")
    big = [positions.empty_board(s) for s in (5, 8, 11, 12, 13, 13, 12, 9)]
    for norm in (False, True):
        run("S b64", 35, 10, 2, play7[:64], norm, 1)
        run("S b64", 35, 10, 2, play7[:64], norm, 0)
        run("L b256", 110, 15, 2, play11, norm, 1, gain=1.7)
        run("big", 110, 15, 2, big, norm, 0, gain=1.7)
        run("H60", 60, 4, 2, play7[:16], norm, 1)


def norm_grad_block(o, H, L, Lh):
    def g(t):
        return torch.zeros_like(t) if t.grad is None else t.grad
    parts = [torch.cat([g(o.gnn.norms[s].weight), g(o.gnn.norms[s].bias)]) for s in range(L)]
    parts.append(torch.cat([g(o.after_embed_norm.weight), g(o.after_embed_norm.bias)]))
    for head in (o.maker_head, o.breaker_head):
        parts += [torch.cat([g(head.gnn.norms[j].weight), g(head.gnn.norms[j].bias)]) for j in range(Lh)]
    return torch.cat(parts).detach().double()


def run_bwd(tag, H, L, Lh, graphs, norm, maker, gain=1.5, mode="q"):
    dev = torch.device("cuda")
    o = orc.ModernTwoHeadedOracle(H, L, Lh, norm=norm)
    helpers.set_hash_weights(o, seed=3, gain=gain)
    o64 = orc.ModernTwoHeadedOracle(H, L, Lh, norm=norm).double()
    o64.load_state_dict({k: v.double() for k, v in o.state_dict().items()})
    m = ModernTwoHeaded(H, L, Lh)
    m.load_state_dict({k: v for k, v in o.state_dict().items() if "norm" not in k})
    m = m.to(dev)
    for g in graphs:
        g.x[:, 2] = maker
    b = data.collate(graphs)
    action, target, weight = helpers.loss_inputs(b)
    idx = action + b.ptr[:-1]
    B = b.num_graphs

    def loss_of(model, dt):
        x = b.x.to(dt)
        if mode == "q":
            q = model(x, b.edge_index, b.batch, b.ptr)
            return (weight.to(dt) * (q[idx] - target.to(dt)) ** 2).mean()
        if mode == "adv":
            a = model(x, b.edge_index, b.batch, b.ptr, advantages_only=True).reshape(-1)
            return (weight.to(dt) * (a[idx] - target.to(dt)) ** 2).mean()
        v, a = model(x, b.edge_index, b.batch, b.ptr, seperate=True)
        return (weight.to(dt) * (a.reshape(-1)[idx] - target.to(dt)) ** 2).mean() + (v.reshape(-1) * target.to(dt)).mean()

    loss_of(o, torch.float32).backward()
    loss_of(o64, torch.float64).backward()
    x, ei, bt, ptr = b.x.to(dev), b.edge_index.to(dev), b.batch.to(dev), b.ptr.to(dev)
    csr = m.build_csr(ei, bt, ptr, num_nodes=x.shape[0])
    npar = norm_block(o, H, L, Lh).to(dev) if norm else None
    flat = m.flat_parameters
    flags = {"q": 0, "adv": ops.HEXGNN_ADV_ONLY, "sep": ops.HEXGNN_SEPARATE}[mode]
    out, val, stash, sz = ops.lw_forward(m._desc, flat, npar, x, csr, flags, want_stash=True)
    d_nodes = torch.zeros(x.shape[0], device=dev)
    d_nodes[idx.to(dev)] = (2.0 * weight.to(dev) * (out[idx.to(dev)] - target.to(dev)) / B)
    d_value = (target.to(dev) / B) if mode == "sep" else None
    grad, gnorm, demb = ops.lw_backward(m._desc, flat, npar, x, csr, flags, d_nodes, d_value, stash, want_d_embeds=True)
    torch.cuda.synchronize()
    grad = grad.cpu().double()
    names = {id(p): n for n, p in m.named_parameters()}
    r64 = dict(o64.named_parameters())
    r32 = dict(o.named_parameters())
    gmax = max(float(p.grad.abs().max()) for p in r64.values() if p.grad is not None)
    worst, worst_name, glob, o32_worst = 0.0, None, 0.0, 0.0
    for p, (off, n, shp) in zip(m._params_cache, m._slices):
        name = names[id(p)]
        rg = r64[name].grad
        rg = torch.zeros(shp, dtype=torch.float64) if rg is None else rg
        g = grad[off:off + n].view(shp)
        d = float((g - rg).abs().max())
        sc = max(float(rg.abs().max()), 1e-3 * gmax)
        glob = max(glob, d / gmax)
        if d / sc > worst:
            worst, worst_name = d / sc, name
        if r32[name].grad is not None:
            o32_worst = max(o32_worst, float((r32[name].grad.double() - rg).abs().max()) / sc)
    msg = f"{tag} norm={int(norm)} maker={maker} {mode} BWD: global {glob:.3e}  worst tensor {worst:.3e} ({worst_name})  fp32-oracle worst {o32_worst:.3e}"
    if norm:
        rn = norm_grad_block(o64, H, L, Lh)
        e = float((gnorm.cpu().double() - rn).abs().max() / rn.abs().max().clamp(min=1e-30))
        msg += f"  norm-grad {e:.3e}"
    print(msg + f"  status {int(csr.status.item())}", flush=True)
    grad2, _, _ = ops.lw_backward(m._desc, flat, npar, x, csr, flags, d_nodes, d_value, stash)
    torch.cuda.synchronize()
    print("   bitwise repeat:", bool(torch.equal(grad2.cpu().double(), grad)), flush=True)
    t0 = time.time()
    for _ in range(10):
        ops.lw_backward(m._desc, flat, npar, x, csr, flags, d_nodes, d_value, stash)
    torch.cuda.synchronize()
    print(f"   lw backward: {(time.time() - t0) / 10 * 1e3:.3f} ms", flush=True)


def main_bwd():
    play7 = positions.load_positions("pos_7x7_seed1234.npz")
    play11 = positions.load_positions("pos_11x11_seed1234.npz")
    big = [positions.empty_board(s) for s in (5, 8, 11, 12, 13, 13, 12, 9)]
    for norm in (False, True):
        run_bwd("S b64", 35, 10, 2, play7[:64], norm, 1)
        run_bwd("S b64", 35, 10, 2, play7[:64], norm, 0, mode="sep")
        run_bwd("S b64", 35, 10, 2, play7[:64], norm, 1, mode="adv")
        run_bwd("big", 110, 15, 2, big, norm, 0, gain=1.7)
        run_bwd("L b256", 110, 15, 2, play11, norm, 1, gain=1.7)
        run_bwd("H60", 60, 4, 2, play7[:16], norm, 1)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "bwd":
        main_bwd()
    else:
        main_fwd()
