"""Profiling driver (not a test): per-phase cycle counts of the fused forward / backward-chain kernels
(hexgnn_debug_phases) on the bench workload.  python tools/phase_profile.py [play|empty]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, data, _lib
import positions  # noqa: E402
mode = sys.argv[1] if len(sys.argv) > 1 else "play"
torch.manual_seed(0)
m = ModernTwoHeaded(110, 15, 2).to("cuda:0")
gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:256] if mode == "play" else [positions.empty_board(11, 1)] * 256
b = data.collate(gs).to("cuda:0")
for _ in range(5):
    m.zero_grad(set_to_none=True)
    q = m(b.x, b.edge_index, b.batch, b.ptr)
    (q * q).mean().backward()
torch.cuda.synchronize()
_lib.profile_enable(True); _lib.profile_read(0); _lib.profile_read(1)
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(20):
    m.zero_grad(set_to_none=True)
    q = m(b.x, b.edge_index, b.batch, b.ptr)
    (q * q).mean().backward()
e.record(); torch.cuda.synchronize()
f, fn = _lib.profile_read(0); bw, bn = _lib.profile_read(1)
print(f"step ms {s.elapsed_time(e)/20:.4f}  fwd kernel {f/max(fn,1):.4f}  bwd chain kernel {bw/max(bn,1):.4f}")
NAMES = {0: ["wait_Zl", "phase1", "gather", "load+l0", "head", "mma:wait_h", "mma:wait_w", "mma:issue", "total", "tiles",
             "wait_Zr", "phase3", "p3:ld", "p3:issued"],
         1: ["wait_dH", "E1", "E2_gather", "wait_r+A", "head+load", "layer0", "hl:load", "hl:+build+tanh", "total", "tiles",
             "hd:dq+adv", "hd:colpass", "hd:mlp_fwd", "hd:mlp_bwd", "hd:hrec+dp", "hd:rows"]}
for which in (0, 1):
    ph = _lib.debug_phases(which)
    print("forward" if which == 0 else "backward chain")
    for k in (1, 2, 3):
        rows = [r for r in ph if r[9] == k]
        if rows:
            print(f"  CTAs with {k} tiles: {len(rows)};", ", ".join(
                f"{n}={statistics.mean(r[i] for r in rows)/k:.0f}" for i, n in enumerate(NAMES[which]) if n not in ("-", "tiles")),
                "(cycles per tile)")
