"""Diagnostic: final_conv_grads vs oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import data
import positions  # noqa: E402
from helpers import make_pair
DEV = "cuda:0"
o, m = make_pair(35, 10)
m = m.to(DEV); m.capture_embed_grads = True
b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:40]); d = b.to(DEV)
gen = torch.Generator().manual_seed(3)
wn = torch.randn(b.x.shape[0], generator=gen); wv = torch.randn(b.num_graphs, generator=gen)
for mode in ("q", "sep", "adv"):
    kw = dict(seperate=True) if mode == "sep" else dict(advantages_only=True) if mode == "adv" else {}
    o.zero_grad(); m.zero_grad()
    r = o(b.x, b.edge_index, b.batch, **kw); o.final_conv_acts.retain_grad()
    k = m(d.x, d.edge_index, d.batch, d.ptr, **kw)
    if mode == "sep":
        ((r[1] * wn).sum() + (r[0] * wv).sum()).backward(); ((k[1] * wn.to(DEV)).sum() + (k[0] * wv.to(DEV)).sum()).backward()
    else:
        (r.reshape(-1) * wn).sum().backward(); (k.reshape(-1) * wn.to(DEV)).sum().backward()
    g, gr = m.final_conv_grads.cpu(), o.final_conv_acts.grad
    err = (g - gr).abs().max(dim=1).values
    bad = (err > 1e-5 * gr.abs().max()).nonzero().flatten()
    print(mode, "max err", float(err.max()), "ref max", float(gr.abs().max()), "bad rows", bad.numel(), "of", g.shape[0])
    print("  bad rows:", bad[:40].tolist()); print("  ptr:", b.ptr[:12].tolist())
    if bad.numel():
        i = int(bad[0]); print("  row", i, "deg", float(b.x[i, 0]), g[i, :6].tolist(), gr[i, :6].tolist())
# ---- more detail for mode q
o.zero_grad(); m.zero_grad()
r = o(b.x, b.edge_index, b.batch); o.final_conv_acts.retain_grad(); (r * wn).sum().backward()
gs = []
for rep in range(2):
    m.zero_grad()
    k = m(d.x, d.edge_index, d.batch, d.ptr); (k * wn.to(DEV)).sum().backward()
    gs.append(m.final_conv_grads.cpu().clone())
print("deterministic:", torch.equal(gs[0], gs[1]))
gr = o.final_conv_acts.grad
csr = m.last_csr
t = csr.tile_g0.cpu()[: int(csr.ntiles.item()) + 1].tolist()
print("tiles(graph idx):", t)
for ti in range(len(t) - 1):
    r0, r1 = int(b.ptr[t[ti]]), int(b.ptr[t[ti + 1]])
    e = (gs[0][r0:r1] - gr[r0:r1]).abs().max() / gr.abs().max()
    ne = int(((b.edge_index[1] >= r0) & (b.edge_index[1] < r1)).sum())
    print(f" tile {ti}: graphs {t[ti]}..{t[ti+1]} rows {r0}..{r1} n={r1-r0} ne={ne} relerr {float(e):.2e}")
acts = o.final_conv_acts.detach()
i = 388
print("row 388 embeds[:8]", acts[i, :8].tolist())
print("err cols where embeds>0:", float((gs[0][i] - gr[i])[acts[i] > 0].abs().max()), " ==0:", float((gs[0][i] - gr[i])[acts[i] == 0].abs().max()))
# ---- per-tensor param-grad errors (q mode, dense wn); m grads from last rep, o grads from r
ref = dict(o.named_parameters())
for name, p in m.named_parameters():
    rg = ref[name].grad
    if rg is None or "breaker" in name:
        continue
    e = float((p.grad.cpu() - rg).abs().max()) / float(rg.abs().max())
    print(f"  {name:40s} refmax {float(rg.abs().max()):.3e} relerr {e:.2e}")
# per-tile: oracle d_embeds restricted -> does masked version agree?
mask = (acts > 0).float()
for ti in range(len(t) - 1):
    r0, r1 = int(b.ptr[t[ti]]), int(b.ptr[t[ti + 1]])
    e1 = ((gs[0][r0:r1] - gr[r0:r1]) * mask[r0:r1]).abs().max() / gr.abs().max()
    print(f" tile {ti}: masked relerr {float(e1):.2e}; graph sizes {[int(b.ptr[g+1]-b.ptr[g]) for g in range(t[ti], t[ti+1])]}")
