"""Small end-to-end run for compute-sanitizer (memcheck / racecheck / synccheck): CSR build, train_step, autograd route,
oversize inference, learner step, device collate on a few 7x7 / mixed positions.
  compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, RainbowLearner, data, ops  # noqa: E402
import positions  # noqa: E402
from helpers import loss_inputs  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
m, t = ModernTwoHeaded(35, 10, 2).to(dev), ModernTwoHeaded(35, 10, 2).to(dev)
gs = positions.load_positions("pos_7x7_seed1234.npz", maker=1)
hb = data.collate(gs[:200])          # > 148 tiles? 7x7 graphs pack several per tile; still exercises two rounds of items
b = hb.to(dev)
a, tg, w = [u.to(dev) for u in loss_inputs(hb)]
loss, err, q = m.train_step(b.x, b.edge_index, b.batch, b.ptr, action=a, target=tg, weight=w)
m.zero_grad()
q2 = m(b.x, b.edge_index, b.batch, b.ptr)
ops.td_loss(q2, b.ptr, a, tg, w)[0].backward()
lrn = RainbowLearner(m, t, lr=1e-3)
hn = data.collate(gs[50:250]).to(dev)
out = lrn.train(b, hn, a, torch.zeros(200, device=dev), torch.zeros(200, device=dev), w)
col = data.DeviceCollator(dev, max_graphs=64)
cb = col([g.to(dev) for g in gs[:40]])
mix = data.collate(positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)[:40]).to(dev)
big = ModernTwoHeaded(110, 15, 2, max_graph_nodes=256).to(dev)
with torch.no_grad():
    qb = big(mix.x, mix.edge_index, mix.batch, mix.ptr)
big_l = ModernTwoHeaded(110, 15, 2).to(dev)
gl = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:180])   # 148+ tiles: both dW phases, 2 rounds
bl = gl.to(dev)
al, tl, wl = [u.to(dev) for u in loss_inputs(gl)]
l3 = big_l.train_step(bl.x, bl.edge_index, bl.batch, bl.ptr, action=al, target=tl, weight=wl)[0]
torch.cuda.synchronize()
print("ok", float(loss), float(out["loss"]), float(l3), int(cb.x.shape[0]), bool(torch.isfinite(qb).all()))
# round-1 additions: noisy-nets advantage layer (compose / chain-rule kernels) and the input gradient (hexgnn_backward_dx)
nz = ModernTwoHeaded(35, 10, 2, noisy_dqn=True).to(dev)
xg = b.x.clone().requires_grad_(True)
ops.td_loss(nz(xg, b.edge_index, b.batch, b.ptr), b.ptr, a, tg, w)[0].backward()
torch.cuda.synchronize()
print("ok noisy + dx", float(xg.grad.abs().max()), float(nz.maker_head.linear.weight_sigma.grad.abs().max()))
