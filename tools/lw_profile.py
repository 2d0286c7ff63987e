"""Profiling driver (not a test) of the layer-at-a-time path: GNN-L, 256 11x11 positions, train_step; and the ragged
5x5..13x13 transfer batch forward (BASELINE configs[3]).  Prints CUDA-event times; run under ncu for the launch list."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import helpers  # noqa: E402
import positions  # noqa: E402
from gnn_hex_b200 import ModernTwoHeaded, data  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
norm = len(sys.argv) > 2 and sys.argv[2] == "norm"
dev = "cuda:0"
torch.manual_seed(0)
m = ModernTwoHeaded(110, 15, 2, path="lw", norm=norm).to(dev)
hb = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1))
b = hb.to(dev)
action, target, weight = [t.to(dev) for t in helpers.loss_inputs(hb)]
csr = m.build_csr(b.edge_index, b.batch, b.ptr, num_nodes=b.x.shape[0])


def timed(fn, n):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n


def fwd():
    with torch.no_grad():
        m(b.x, b.edge_index, b.batch, b.ptr, csr=csr)


def step():
    m.zero_grad()
    m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=action, target=target, weight=weight, csr=csr)


print(f"lw GNN-L b256 norm={int(norm)}: forward {timed(fwd, reps):.3f} ms, train_step {timed(step, reps):.3f} ms", flush=True)
mix = data.collate(positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)).to(dev)
mr = ModernTwoHeaded(110, 15, 2, max_graph_nodes=256).to(dev)
csr2 = mr.build_csr(mix.edge_index, mix.batch, mix.ptr, num_nodes=mix.x.shape[0])


def ragged():
    with torch.no_grad():
        mr(mix.x, mix.edge_index, mix.batch, mix.ptr, csr=csr2)


print(f"ragged 5..13 forward ({mix.x.shape[0]} nodes): {timed(ragged, reps):.3f} ms", flush=True)
