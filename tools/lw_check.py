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


if __name__ == "__main__":
    play7 = positions.load_positions("pos_7x7_seed1234.npz")
    play11 = positions.load_positions("pos_11x11_seed1234.npz")
    big = [positions.empty_board(s) for s in (5, 8, 11, 12, 13, 13, 12, 9)]
    for norm in (False, True):
        run("S b64", 35, 10, 2, play7[:64], norm, 1)
        run("S b64", 35, 10, 2, play7[:64], norm, 0)
        run("L b256", 110, 15, 2, play11, norm, 1, gain=1.7)
        run("big", 110, 15, 2, big, norm, 0, gain=1.7)
        run("H60", 60, 4, 2, play7[:16], norm, 1)
