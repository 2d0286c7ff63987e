"""BASELINE.json configs[4]: GNN-L data-parallel Rainbow learner step, GLOBAL batch 2048 sharded over the ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 \
        tools/bench_dp_learner.py [iters] [global_batch]            (N = 1 works without torchrun)

Each rank takes global_batch / N positions (state and next_state: the committed 11x11 fixture, repeated and rotated per
rank), runs RainbowLearner.train = TD target (one CSR, two next-state forwards, mask + arg-max), train_step (forward +
PER-weighted loss + backward, loss scaled by 1/N), ONE all-reduce of the flat 1.95 MB gradient (own multimem kernel; HEXGNN_ALLREDUCE=nccl for NCCL), clip + Adam.
Strong scaling: the global batch is fixed.  CUDA events per step, L2 flushed between steps, max over ranks; rank 0 prints
one JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, RainbowLearner, data
import positions  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 50
global_batch = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
per = global_batch // world

gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1)


def shard(offset):
    k = (offset + 37 * rank) % len(gs)
    rot = gs[k:] + gs[:k]
    return data.collate((rot * ((per + len(rot) - 1) // len(rot)))[:per])


hs, hn = shard(0), shard(128)
ds, dn = hs.to(dev), hn.to(dev)
g = torch.Generator().manual_seed(1 + rank)
n = hs.ptr[1:] - hs.ptr[:-1]
action = (2 + (torch.rand(per, generator=g) * (n - 2).clamp(min=1)).long()).clamp(max=n - 1).clamp(min=0).to(dev)
weights = (torch.rand(per, generator=g) * 0.5 + 0.5).to(dev)
reward = (torch.rand(per, generator=g) < 0.1).float().to(dev)
done = (torch.rand(per, generator=g) < 0.1).float().to(dev)

torch.manual_seed(0)
pol, tgt = ModernTwoHeaded(110, 15, 2).to(dev), ModernTwoHeaded(110, 15, 2).to(dev)
collective = "none"
if world > 1:
    collective = "nccl"
    if os.environ.get("HEXGNN_ALLREDUCE", "mc") == "mc":
        try:
            if pol.enable_p2p_allreduce(mode="mc"):
                collective = "hexgnn_allreduce_mc"
        except Exception:  # noqa: BLE001 -- no multicast address: NCCL
            pass
lrn = RainbowLearner(pol, tgt, lr=1e-4, adam_eps=1.5e-4, max_grad_norm=10.0, gamma=0.97, n_step=2, world_size=world)
lrn.sync_Q_target()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


for _ in range(10):
    out = lrn.train(ds, dn, action, reward, done, weights)
evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
barrier()
for s, e in evs:
    flush.fill_(1)
    s.record()
    out = lrn.train(ds, dn, action, reward, done, weights)
    e.record()
barrier()
ms = torch.tensor([sum(s.elapsed_time(e) for s, e in evs) / iters], dtype=torch.float64, device=dev)
chk = pol.flat_parameters.double().sum().reshape(1)      # replicas must stay in lock-step
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    in_step = bool((lo == hi).item())
else:
    in_step = True
status = int(pol.last_csr.status.item())
if rank == 0:
    print(json.dumps({"config": f"dp learner: RainbowLearner.train, GNN-L, global batch {global_batch} (state + next_state 11x11 "
                                f"positions) over {world} B200, {collective} all-reduce of the flat gradient, clip + Adam",
                      "n_gpus": world, "graphs_per_gpu": per, "nodes_per_gpu": int(hs.x.shape[0] + hn.x.shape[0]), "ms_per_step": float(ms),
                      "graphs_per_s": global_batch / float(ms) * 1e3, "scaling": "strong", "replicas_identical": in_step,
                      "loss": float(out["loss"]), "device_status": status}), flush=True)
if world > 1:
    dist.destroy_process_group()
