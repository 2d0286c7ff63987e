"""GPU, 2 ranks on 2 devices: data-parallel correctness of the fused path (no reference counterpart).

The sum over ranks of the shard gradients -- every rank runs `train_step` on its node-balanced contiguous shard
(data.shard_graphs) with the loss scaled by B_local / B_global, then `allreduce_gradients()` -- must equal the 1-GPU gradient
of the union batch to fp32 summation-order round-off, for NCCL and for the library's own all-reduce over NVLink peer memory
(`hexgnn_allreduce_p2p` one-shot pull, `hexgnn_allreduce_mc` two-shot through the NVSwitch), and every rank must end up with the same bits.  Skipped with fewer than two devices."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import helpers
        import positions
        from gnn_hex_b200 import ModernTwoHeaded, data
        gs = positions.load_positions("pos_11x11_seed1234.npz", maker=1)[:96]
        union = data.collate(gs)
        action, target, weight = helpers.loss_inputs(union)
        torch.manual_seed(0)
        m = ModernTwoHeaded(110, 15, 2).to(dev)          # same seed on every rank: identical replicas
        if mode in ("pull", "mc"):
            try:
                assert m.enable_p2p_allreduce(mode=mode), "symmetric memory unavailable"
            except Exception as e:  # noqa: BLE001
                if mode == "mc" and "multicast" in str(e):
                    out[rank] = ("skip", str(e), 0)
                    return
                raise
            assert m._p2p["mc"] == (mode == "mc")
        a, b = data.shard_by_nodes([int(g.x.shape[0]) for g in gs], world)[rank]
        shard = data.collate(gs[a:b]).to(dev)
        m.zero_grad(set_to_none=True)
        m.train_step(shard.x, shard.edge_index, shard.batch, shard.ptr, action=action[a:b].to(dev), target=target[a:b].to(dev),
                     weight=weight[a:b].to(dev), scale=(b - a) / len(gs))
        m.allreduce_gradients()
        torch.cuda.synchronize()
        g_dp = m.flat_gradient.clone()
        # a second step through the other peer-mapped buffer (they alternate) gives the same reduced gradient again
        m.zero_grad(set_to_none=True)
        m.train_step(shard.x, shard.edge_index, shard.batch, shard.ptr, action=action[a:b].to(dev), target=target[a:b].to(dev),
                     weight=weight[a:b].to(dev), scale=(b - a) / len(gs))
        m.allreduce_gradients()
        torch.cuda.synchronize()
        assert torch.equal(m.flat_gradient, g_dp)
        gathered = [torch.empty_like(g_dp) for _ in range(world)]
        dist.all_gather(gathered, g_dp)
        same_bits = all(torch.equal(gathered[0], t) for t in gathered)
        err = None
        if rank == 0:
            m1 = ModernTwoHeaded(110, 15, 2).to(dev)
            m1.load_state_dict(m.state_dict())
            d = union.to(dev)
            m1.train_step(d.x, d.edge_index, d.batch, d.ptr, action=action.to(dev), target=target.to(dev), weight=weight.to(dev))
            torch.cuda.synchronize()
            g1 = m1.flat_gradient
            err = float((g_dp - g1).abs().max() / g1.abs().max())
        out[rank] = (same_bits, err, b - a)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["nccl", "pull", "mc"])
def test_two_rank_gradient_equals_union_gradient(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), mode, out), nprocs=world, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    if res[0][0] == "skip":
        pytest.skip("no multicast (NVLS) address for symmetric memory on this box: " + res[0][1])
    assert res[0][0] and res[1][0], "ranks hold different reduced gradients"
    assert res[0][2] + res[1][2] == 96 and res[0][2] != 48      # node-balanced, not a graph-count split
    assert res[0][1] is not None and res[0][1] <= 1e-6, res[0][1]
