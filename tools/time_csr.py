"""Time hexgnn_build_csr alone on the bench batch with CUDA events (L2 flushed between calls, as in bench.py)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from gnn_hex_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
b = bench.make_batch("play", 0).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for mode in ("flush", "warm"):
    ts = []
    for i in range(60):
        if mode == "flush":
            flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        csr = ops.build_csr(b.edge_index, b.batch, b.ptr, num_nodes=b.x.shape[0], num_graphs=b.num_graphs)
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts = sorted(ts[10:])
    print(mode, "build_csr us: median %.1f min %.1f" % (ts[len(ts) // 2], ts[0]))
