"""Profiling driver (not a test): N forward-only calls of GNN-L on 256 positions."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, data
import positions  # noqa: E402
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
mode = sys.argv[2] if len(sys.argv) > 2 else "play"
torch.manual_seed(0)
m = ModernTwoHeaded(110, 15, 2).to("cuda:0")
gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1) if mode == "play" else [positions.empty_board(11, 1)] * 256
b = data.collate(gs).to("cuda:0")
with torch.no_grad():
    csr = m.build_csr(b.edge_index, b.batch, b.ptr, num_nodes=b.x.shape[0])
    for _ in range(reps):
        q = m(b.x, b.edge_index, b.batch, b.ptr, csr=csr)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(20):
        q = m(b.x, b.edge_index, b.batch, b.ptr, csr=csr)
    e.record(); torch.cuda.synchronize()
print("fwd ms/call", s.elapsed_time(e) / 20, "q[0]", float(q[0]))
from gnn_hex_b200 import _lib
ph = _lib.debug_phases(0)
import statistics
names = ["wait_Z", "phase1", "phase2", "load+l0", "head", "mma:wait_h", "mma:wait_w", "mma:issue", "total", "tiles"]
for k in (1, 2):
    rows = [r for r in ph if r[9] == k]
    if rows:
        print(f"CTAs with {k} tiles: {len(rows)};", ", ".join(f"{n}={statistics.mean(r[i] for r in rows)/k:.0f}" for i, n in enumerate(names[:9])), "(cycles per tile)")
