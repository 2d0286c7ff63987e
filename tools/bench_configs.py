"""Secondary configurations of BASELINE.json (not the headline: bench.py measures configs[1]).  One JSON line each:
  gnn_s     configs[0]  GNN-S (hidden 35, 10+2 layers) fwd+bwd, 256 7x7 positions
  selfplay  configs[2]  GNN-L forward only, 128 11x11 positions, advantages_only + terminal mask + per-graph argmax
  ragged    configs[3]  GNN-L forward, 256 graphs of mixed 5x5..13x13 boards (ragged CSR; layer-at-a-time tensor-core path),
                        the same batch through train_step, and the mix with 14x14..19x19 boards added
  norm      SURVEY 8 f4: GNN-L with norm=True (PyG LayerNorm, whole-batch statistics) fwd+bwd, 256 11x11 positions
  hexara    SURVEY 8 f3: HexAra SAGE_torch_script (hidden 60, 15+2+2 convs) inference, 128 / 256 11x11 positions
  learner   SURVEY 8 f1+f2: the whole Rainbow learner step on GNN-L, 256 + 256 positions (RainbowLearner.train:
            2 CSR builds, 3 forwards, mask/argmax, TD target, train_step, clip + Adam)
CUDA events, inputs resident, 256 MiB L2 flush between iterations.   python tools/bench_configs.py [iters]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, data, ops
import positions  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    for _ in range(10):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    torch.cuda.synchronize()
    for s, e in evs:
        flush.fill_(1)
        s.record(); fn(); e.record()
    torch.cuda.synchronize()
    return sum(s.elapsed_time(e) for s, e in evs) / iters


def loss_inputs(b):
    g = torch.Generator().manual_seed(1)
    n = b.ptr[1:] - b.ptr[:-1]
    u = torch.rand(b.num_graphs, generator=g)
    action = (2 + (u * (n - 2).clamp(min=1)).long()).clamp(max=n - 1).clamp(min=0)
    return action.to(dev), (torch.rand(b.num_graphs, generator=g) * 2 - 1).to(dev), torch.ones(b.num_graphs, device=dev)


torch.manual_seed(0)
out = []
# configs[0]
m = ModernTwoHeaded(35, 10, 2).to(dev)
b = data.collate(positions.load_positions("pos_7x7_seed1234.npz", maker=1)[:256]); d = b.to(dev)
a, t, w = loss_inputs(b)
def step_s():
    m.zero_grad(set_to_none=True)
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=a, target=t, weight=w)
ms = timed(step_s)
out.append({"config": "gnn_s fwd+bwd, 256 7x7 positions", "nodes": int(b.x.shape[0]), "ms": ms, "graphs_per_s": 256 / ms * 1e3})
# configs[2]
m = ModernTwoHeaded(110, 15, 2).to(dev)
b = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:128]); d = b.to(dev)
def act():
    with torch.no_grad():
        adv = m(d.x, d.edge_index, d.batch, d.ptr, advantages_only=True).reshape(-1)
        return ops.mask_argmax(adv, d.ptr)[1]
ms = timed(act)
out.append({"config": "selfplay: GNN-L forward (advantages_only) + mask + argmax, 128 11x11 positions", "nodes": int(b.x.shape[0]),
            "ms": ms, "graphs_per_s": 128 / ms * 1e3})
# configs[3]
b = data.collate(positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)[:256]); d = b.to(dev)
m.max_graph_nodes = 363   # batches with graphs beyond one 128-row tile run the layer-at-a-time tensor-core path
def fwd():
    with torch.no_grad():
        return m(d.x, d.edge_index, d.batch, d.ptr)
ms = timed(fwd)
status = int(m.last_csr.status.item()) if getattr(m, "last_csr", None) is not None else 0
out.append({"config": "ragged: GNN-L forward, 256 graphs of 5x5..13x13 boards (hexgnn_forward_lw)", "nodes": int(b.x.shape[0]), "graphs": b.num_graphs,
            "ms": ms, "graphs_per_s": b.num_graphs / ms * 1e3, "device_status": status})
a, t, w = loss_inputs(b)
def step_ragged():
    m.zero_grad(set_to_none=True)
    m.train_step(d.x, d.edge_index, d.batch, d.ptr, action=a, target=t, weight=w)
ms = timed(step_ragged)
out.append({"config": "ragged: GNN-L fwd+bwd (train_step via hexgnn_forward_lw / hexgnn_backward_lw), same batch", "nodes": int(b.x.shape[0]),
            "graphs": b.num_graphs, "ms": ms, "graphs_per_s": b.num_graphs / ms * 1e3})
b = data.collate(positions.load_positions("pos_mix_5_13_seed1234.npz", maker=1)[:244] +
                 positions.load_positions("pos_transfer_14_19_seed1234.npz", maker=1)); d = b.to(dev)
ms = timed(fwd)
out.append({"config": "ragged + 14x14..19x19 boards (up to 363 nodes): GNN-L forward, 256 graphs", "nodes": int(b.x.shape[0]),
            "graphs": b.num_graphs, "ms": ms, "graphs_per_s": b.num_graphs / ms * 1e3})
# f4: norm=True
mn = ModernTwoHeaded(110, 15, 2, norm=True).to(dev)
b = data.collate(positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:256]); d = b.to(dev)
a, t, w = loss_inputs(b)
def step_norm():
    mn.zero_grad(set_to_none=True)
    mn.train_step(d.x, d.edge_index, d.batch, d.ptr, action=a, target=t, weight=w)
ms = timed(step_norm)
out.append({"config": "norm=True: GNN-L fwd+bwd (LayerNorm over the whole batch), 256 11x11 positions", "nodes": int(b.x.shape[0]),
            "ms": ms, "graphs_per_s": 256 / ms * 1e3})
def fwd_norm():
    with torch.no_grad():
        return mn(d.x, d.edge_index, d.batch, d.ptr)
ms = timed(fwd_norm)
out.append({"config": "norm=True: GNN-L forward, 256 11x11 positions", "nodes": int(b.x.shape[0]), "ms": ms, "graphs_per_s": 256 / ms * 1e3})
# f3: HexAra
from gnn_hex_b200 import get_current_model
ara = get_current_model("SAGE", swap_allowed=True).to(dev)
for count in (128, 256):
    b = data.collate(positions.load_positions_cpp("pos_11x11_seed1234.npz")[:count]); d = b.to(dev)
    def predict():
        with torch.no_grad():
            return ara(d.x, d.edge_index, d.batch, d.ptr)
    ms = timed(predict)
    out.append({"config": f"hexara: SAGE_torch_script (hidden 60, 15+2+2 convs, swap) inference incl. CSR build, {count} 11x11 positions",
                "nodes": int(b.x.shape[0]), "ms": ms, "graphs_per_s": count / ms * 1e3})
yv = torch.rand(256, device=dev) * 2 - 1
def ara_train():
    ara.zero_grad(set_to_none=True)
    pi, value, gi, bp = ara(d.x, d.edge_index, d.batch, d.ptr)
    loss = 0.01 * torch.nn.functional.mse_loss(value, yv) * 256 - 0.99 * pi.sum() / 256     # trainer_agent_pytorch.py:255-261 shape
    loss.backward()
ms = timed(ara_train)
out.append({"config": "hexara: SAGE_torch_script fwd + loss + bwd (autograd -> hexgnn_hexara_backward) incl. CSR build, 256 11x11 positions",
            "nodes": int(b.x.shape[0]), "ms": ms, "graphs_per_s": 256 / ms * 1e3})
# learner step (f1 + f2)
from gnn_hex_b200 import RainbowLearner
pol, tgt = ModernTwoHeaded(110, 15, 2).to(dev), ModernTwoHeaded(110, 15, 2).to(dev)
lrn = RainbowLearner(pol, tgt, lr=1e-4, adam_eps=1.5e-4, max_grad_norm=10.0, gamma=0.97, n_step=2)
gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1)
hs, hn = data.collate(gs[:256]), data.collate(gs[128:256] + gs[:128])
ds, dn = hs.to(dev), hn.to(dev)
a, _, w = loss_inputs(hs)
g = torch.Generator().manual_seed(2)
reward = (torch.rand(256, generator=g) < 0.1).float().to(dev)
done = (torch.rand(256, generator=g) < 0.1).float().to(dev)
ms = timed(lambda: lrn.train(ds, dn, a, reward, done, w))
out.append({"config": "learner: RainbowLearner.train, GNN-L, 256 state + 256 next_state 11x11 positions (double DQN, PER weights, "
                      "clip + Adam)", "nodes": int(hs.x.shape[0] + hn.x.shape[0]), "ms": ms, "steps_per_s": 1e3 / ms,
            "graphs_per_s": 256 / ms * 1e3})
for o in out:
    print(json.dumps(o))
