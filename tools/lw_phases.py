"""Per-phase cycle profile of lw_sage_kernel (HEXGNN_PHASES build: HEXGNN_PHASES=1 HEXGNN_LIB=... python -m gnn_hex_b200.build)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import positions  # noqa: E402
from gnn_hex_b200 import ModernTwoHeaded, _lib, data  # noqa: E402

dev = "cuda:0"
m = ModernTwoHeaded(110, 15, 2, path="lw").to(dev)
names = ["setup+pdl_wait", "G1 stage rows+csr", "G2 gather", "A split->TMEM", "wait MMA", "E epilogue", "CTAs"]
for tag, fixture in (("play 11x11 b256", "pos_11x11_seed1234.npz"), ("ragged 5..13 b256", "pos_mix_5_13_seed1234.npz")):
    b = data.collate(positions.load_positions(fixture, maker=1)).to(dev)
    csr = m.build_csr(b.edge_index, b.batch, b.ptr, num_nodes=b.x.shape[0])
    with torch.no_grad():
        for _ in range(3):
            m(b.x, b.edge_index, b.batch, b.ptr, csr=csr)
        torch.cuda.synchronize()
        buf = (C.c_int64 * 8)()
        _lib.check(_lib.lib().hexgnn_debug_phases(2, buf, 8), "debug_phases")
        m(b.x, b.edge_index, b.batch, b.ptr, csr=csr)
        torch.cuda.synchronize()
        _lib.check(_lib.lib().hexgnn_debug_phases(2, buf, 8), "debug_phases")
    n = max(1, buf[6])
    print(tag, f"({b.x.shape[0]} nodes, {n} CTAs over 16 layers):", ", ".join(f"{nm}={buf[i] / n:.0f}" for i, nm in enumerate(names[:6])),
          "cycles per CTA", flush=True)
