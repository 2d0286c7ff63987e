"""CPU, world_size 2 over gloo: the host-side data-parallel logic -- rank shards of the replay batch are
disjoint rotations of the fixture, and ModernTwoHeaded.allreduce_gradients() sums the ONE flat gradient every
p.grad aliases.  The kernels themselves need a GPU and are not called here."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gnn_hex_b200 import ModernTwoHeaded
        import bench
        torch.manual_seed(0)
        m = ModernTwoHeaded(35, 10, 2)
        m._flatten()                                   # CPU flat buffers: plumbing only, no kernel is launched
        fg = m._flat_grad_buffer()
        fg.copy_(torch.arange(fg.numel(), dtype=torch.float32) * (rank + 1))
        m._point_grads()
        m.allreduce_gradients()
        want = torch.arange(fg.numel(), dtype=torch.float32) * sum(r + 1 for r in range(world))
        ok_sum = torch.equal(m.flat_gradient, want)
        # every named grad aliases the reduced buffer
        w = m.gnn.convs[3].lin_l.weight
        ok_alias = w.grad.data_ptr() == m.flat_gradient.data_ptr() + 4 * m._slices[9][0] and torch.equal(
            w.grad.reshape(-1), want[m._slices[9][0]:m._slices[9][0] + w.numel()])
        # noisy nets: the advantage layers' mu / sigma gradients live outside the flat buffer and are reduced with it
        mn = ModernTwoHeaded(12, 3, 2, noisy_dqn=True)
        mn._flatten()
        mn._flat_grad_buffer().fill_(float(rank + 1))
        mn._point_grads()
        mn._noisy_grad = torch.full((2, 2, 13), float(10 * (rank + 1)))
        mn.allreduce_gradients()
        tot = float(sum(r + 1 for r in range(world)))
        ok_noisy = bool(torch.all(mn._noisy_grad == 10 * tot)) and bool(torch.all(mn.flat_gradient == tot))
        b = bench.make_batch("play", rank, world)
        out[rank] = (ok_sum, ok_alias and ok_noisy, int(b.num_graphs), int(b.x.shape[0]), float(b.x[:, 0].sum()))
    finally:
        dist.destroy_process_group()


def test_flat_gradient_allreduce_and_sharding_world2():
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
        res = dict(out)
    assert set(res) == {0, 1}
    for r in range(world):
        ok_sum, ok_alias, ng, nn, _ = res[r]
        assert ok_sum and ok_alias and ng == 256 and nn > 0
    # shards are different rotations of the fixture (same multiset of graphs, different order)
    assert res[0][3] == res[1][3] and res[0][4] == res[1][4]


def test_shard_by_nodes_is_a_balanced_contiguous_cover():
    """contiguous graph ranges per rank, balanced by node count."""
    import random
    sys.path.insert(0, ROOT)
    from gnn_hex_b200 import data
    rng = random.Random(3)
    for world in (1, 2, 3, 4, 8):
        for G in (0, 1, 5, 8, 256, 2048):
            ns = [rng.randint(3, 123) for _ in range(G)]
            r = data.shard_by_nodes(ns, world)
            assert len(r) == world and r[0][0] == 0 and r[-1][1] == G
            assert all(r[k][1] == r[k + 1][0] and r[k][0] <= r[k][1] for k in range(world - 1))
            if G >= world:
                assert all(b > a for a, b in r)
                loads = [sum(ns[a:b]) for a, b in r]
                assert max(loads) - min(loads) <= 2 * 123 + 1, (world, G, loads)
    # a node-balanced split is not a graph-count split: one heavy graph and many light ones
    assert data.shard_by_nodes([120] + [4] * 30, 2) == [(0, 1), (1, 31)]
