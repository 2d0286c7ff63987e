"""torchrun check (N >= 2 GPUs): the library's one-shot all-reduce over NVLink peer memory (hexgnn_allreduce_p2p)
against NCCL on the same gradients.  Prints max |diff| per rank; exits non-zero on mismatch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gnn_hex_b200 import ModernTwoHeaded, data, ops
import positions  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1)
k = (32 * rank) % len(gs)
b = data.collate((gs[k:] + gs[:k])[:64]).to(dev)
action = torch.full((b.num_graphs,), 2, dtype=torch.long, device=dev)
target = torch.linspace(-1, 1, b.num_graphs, device=dev)
grads = {}
for mode in ("nccl", "p2p"):
    torch.manual_seed(0)
    m = ModernTwoHeaded(110, 15, 2).to(dev)
    if mode == "p2p":
        assert m.enable_p2p_allreduce(), "symmetric memory unavailable"
    for it in range(3):   # repeated steps exercise the barrier / buffer reuse protocol
        m.zero_grad(set_to_none=True)
        q = m(b.x, b.edge_index, b.batch, b.ptr)
        loss, _ = ops.td_loss(q, b.ptr, action, target, None, "mse", 1.0 / world)
        loss.backward()
        m.allreduce_gradients()
    torch.cuda.synchronize()
    grads[mode] = m.flat_gradient.clone()
diff = float((grads["nccl"] - grads["p2p"]).abs().max())
scale = float(grads["nccl"].abs().max())
gathered = [torch.zeros_like(grads["p2p"]) for _ in range(world)]
dist.all_gather(gathered, grads["p2p"])
same = all(torch.equal(gathered[0], g) for g in gathered)
print(f"rank {rank}: max|nccl - p2p| = {diff:.3e} (scale {scale:.3e}); p2p bit-identical across ranks: {same}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if (diff <= 1e-6 * scale and same) else 1)
