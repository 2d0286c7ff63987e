"""Host-side data plumbing of the hot path: Hex position graphs in the reference's tensor layout,
and the PyG-style collate (host and device).

Layout mirrors `convert_node_switching_game(..., old_style=True)` (GN0/util/convert_graph.py:77-85,
99-117): x[n,3] = [degree, is_terminal (rows 0,1), maker flag], edge_index = cat(E, flip(E)) int64,
and `Batch.from_data_list` / C++ `collate_batch` (cpp_hex/hex_graph_game/util.cpp:22-41).
"""
import os
from types import SimpleNamespace
from typing import List, Sequence

import numpy as np
import torch

class HexGraph(SimpleNamespace):
    """x [n,3] float32, edge_index [2,E] int64 (symmetric)."""

    def to(self, device, non_blocking=False):
        return HexGraph(x=self.x.to(device, non_blocking=non_blocking),
                        edge_index=self.edge_index.to(device, non_blocking=non_blocking))


class HexBatch(SimpleNamespace):
    """x, edge_index, batch, ptr, num_graphs -- the fields of a PyG `Batch` the model reads."""

    def to(self, device, non_blocking=False):
        buf = getattr(self, "_packed", None)
        if buf is not None and torch.device(device).type == "cuda":
            # a pinned, packed batch travels as ONE host-to-device copy; the fields are views of the device buffer
            return HexBatch._views(buf.to(device, non_blocking=non_blocking), self.x.shape[0], self.edge_index.shape[1],
                                   self.num_graphs)
        return HexBatch(x=self.x.to(device, non_blocking=non_blocking),
                        edge_index=self.edge_index.to(device, non_blocking=non_blocking),
                        batch=self.batch.to(device, non_blocking=non_blocking),
                        ptr=self.ptr.to(device, non_blocking=non_blocking), num_graphs=self.num_graphs)

    @staticmethod
    def _views(buf, N, E, B):
        """Fields as views of one byte buffer laid out [edge_index i64 2E | batch i64 N | ptr i64 B+1 | x f32 3N] (`_offsets`)."""
        o1, o2, o3, _ = HexBatch._offsets(N, E, B)
        out = HexBatch(x=buf[o3:o3 + 12 * N].view(torch.float32).view(N, 3),
                       edge_index=buf[:16 * E].view(torch.int64).view(2, E), batch=buf[o1:o1 + 8 * N].view(torch.int64),
                       ptr=buf[o2:o2 + 8 * (B + 1)].view(torch.int64), num_graphs=B)
        out._packed = buf
        return out

    @staticmethod
    def _offsets(N, E, B):
        """Byte offsets of batch, ptr, x and the total size; every field starts on a 256-byte boundary (the alignment a
        separate device allocation would have)."""
        up = lambda v: (v + 255) // 256 * 256
        o1 = up(16 * E)
        o2 = o1 + up(8 * N)
        o3 = o2 + up(8 * (B + 1))
        return o1, o2, o3, o3 + 12 * N

    def pin_memory(self):
        """Pinned host copy of the batch, packed into one buffer so that `.to(cuda)` is a single copy."""
        N, E, B = int(self.x.shape[0]), int(self.edge_index.shape[1]), int(self.num_graphs)
        if self.x.dtype != torch.float32 or self.edge_index.dtype != torch.int64 or int(self.ptr.shape[0]) != B + 1:
            return HexBatch(x=self.x.pin_memory(), edge_index=self.edge_index.pin_memory(),
                            batch=self.batch.pin_memory(), ptr=self.ptr.pin_memory(), num_graphs=self.num_graphs)
        buf = torch.empty(HexBatch._offsets(N, E, B)[3], dtype=torch.uint8).pin_memory()
        out = HexBatch._views(buf, N, E, B)
        out.x.copy_(self.x); out.edge_index.copy_(self.edge_index); out.batch.copy_(self.batch); out.ptr.copy_(self.ptr)
        return out


def graph_from_directed(n: int, src: np.ndarray, dst: np.ndarray, maker: int) -> HexGraph:
    """Symmetric directed edge list (as the C++ game stores it) -> reference Python layout."""
    src = np.asarray(src, dtype=np.int64)
    dst = np.asarray(dst, dtype=np.int64)
    keep = src < dst
    e = torch.from_numpy(np.stack([src[keep], dst[keep]]))
    edge_index = torch.cat((e, torch.flip(e, dims=(0,))), dim=1).contiguous()
    x = torch.zeros((n, 3), dtype=torch.float32)
    x[:, 0] = torch.bincount(edge_index[1], minlength=n).float()
    x[0, 1] = 1
    x[1, 1] = 1
    x[:, 2] = float(maker)
    return HexGraph(x=x, edge_index=edge_index)


def collate(graphs: Sequence[HexGraph]) -> HexBatch:
    """`Batch.from_data_list` for (x, edge_index): host-side, all int64 (util.cpp:22-41)."""
    ns = [int(g.x.shape[0]) for g in graphs]
    ptr = torch.zeros(len(graphs) + 1, dtype=torch.long)
    ptr[1:] = torch.cumsum(torch.tensor(ns, dtype=torch.long), 0)
    x = torch.cat([g.x for g in graphs], 0)
    edge_index = torch.cat([g.edge_index + int(o) for g, o in zip(graphs, ptr[:-1])], 1)
    batch = torch.repeat_interleave(torch.arange(len(graphs), dtype=torch.long), torch.tensor(ns, dtype=torch.long))
    return HexBatch(x=x, edge_index=edge_index, batch=batch, ptr=ptr, num_graphs=len(graphs))


def shard_by_nodes(num_nodes: Sequence[int], world: int):
    """Data-parallel partition of a replay batch: `world` CONTIGUOUS ranges of graphs, balanced by their
    node counts (positions have 3..123 nodes, so equal graph counts would not be equal work).  Boundary k is the graph
    index whose node prefix sum is nearest to k * total / world; every rank gets at least one graph when there are at
    least `world` of them.  Returns [(start, end)] * world.  Pure host arithmetic on the `ptr` / node-count vector; the
    same on every rank, no communication."""
    n = [int(v) for v in num_nodes]
    G = len(n)
    if world < 1:
        raise ValueError("world must be >= 1")
    pre = [0]
    for v in n:
        pre.append(pre[-1] + v)
    total = pre[-1]
    cuts = [0]
    for k in range(1, world):
        lo = min(cuts[-1] + 1, G) if G >= world else cuts[-1]          # keep ranks non-empty when possible
        hi = max(lo, G - (world - k)) if G >= world else G
        target = total * k / world
        best = lo
        for g in range(lo, hi + 1):                                   # prefix sums are monotone: first minimum wins
            if abs(pre[g] - target) < abs(pre[best] - target):
                best = g
            elif pre[g] > target:
                break
        cuts.append(best)
    cuts.append(G)
    return [(cuts[k], cuts[k + 1]) for k in range(world)]


def shard_graphs(graphs: Sequence[HexGraph], world: int, rank: int) -> List[HexGraph]:
    """This rank's contiguous share of `graphs` under `shard_by_nodes`."""
    a, b = shard_by_nodes([int(g.x.shape[0]) for g in graphs], world)[rank]
    return list(graphs[a:b])


class DeviceCollator:
    """`Batch.from_data_list` over graphs that already live on the GPU (the learner's replay-buffer samples,
    replay_buffer.py:116-117,215-216) as ONE kernel launch (hexgnn_collate) instead of 2 B small cat / add kernels
    issued from a Python loop.  Keeps one pinned table + one device table per slot and alternates between `slots` of them,
    so a table is not rewritten while an earlier call may still be reading it."""

    def __init__(self, device, max_graphs=4096, slots=4):
        import ctypes as C
        from . import _lib
        self._C, self._lib = C, _lib
        self.device = torch.device(device)
        self.max_graphs = int(max_graphs)
        nbytes = int(_lib.lib().hexgnn_collate_ws_bytes(self.max_graphs))
        self._host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(slots)]
        self._dev = [torch.empty(nbytes, dtype=torch.uint8, device=self.device) for _ in range(slots)]
        self._events = [None] * slots
        self._k = 0

    def __call__(self, graphs: Sequence[HexGraph]) -> HexBatch:
        C, _lib = self._C, self._lib
        B = len(graphs)
        if B < 1 or B > self.max_graphs:
            raise _lib.HexGnnError(f"collate: 1..{self.max_graphs} graphs expected, got {B}")
        xs, eis = [], []
        for g in graphs:
            x, ei = g.x, g.edge_index
            if not (x.is_cuda and ei.is_cuda) or x.dtype != torch.float32 or ei.dtype != torch.int64:
                raise _lib.HexGnnError("collate: every graph needs CUDA x float32 [n,3] and edge_index int64 [2,e]")
            xs.append(x.contiguous())
            eis.append(ei.contiguous())
        n_nodes = (C.c_int64 * B)(*[int(x.shape[0]) for x in xs])
        n_edges = (C.c_int64 * B)(*[int(e.shape[1]) for e in eis])
        x_ptrs = (C.c_void_p * B)(*[x.data_ptr() for x in xs])
        ei_ptrs = (C.c_void_p * B)(*[e.data_ptr() for e in eis])
        N, E = sum(n_nodes), sum(n_edges)
        dev = self.device
        x_out = torch.empty((N, 3), dtype=torch.float32, device=dev)
        ei_out = torch.empty((2, E), dtype=torch.int64, device=dev)
        batch = torch.empty(N, dtype=torch.int64, device=dev)
        ptr = torch.empty(B + 1, dtype=torch.int64, device=dev)
        k = self._k
        self._k = (k + 1) % len(self._host)
        if self._events[k] is not None:
            self._events[k].synchronize()          # the table of this slot has been consumed
        stream = torch.cuda.current_stream(dev)
        with torch.cuda.device(dev):
            rc = _lib.lib().hexgnn_collate(x_ptrs, ei_ptrs, n_nodes, n_edges, B, C.c_void_p(x_out.data_ptr()),
                                           C.c_void_p(ei_out.data_ptr()), C.c_void_p(batch.data_ptr()),
                                           C.c_void_p(ptr.data_ptr()), C.c_void_p(self._dev[k].data_ptr()),
                                           self._dev[k].numel(), C.c_void_p(self._host[k].data_ptr()), self._host[k].numel(),
                                           C.c_void_p(stream.cuda_stream))
        _lib.check(rc, "hexgnn_collate")
        ev = torch.cuda.Event()
        ev.record(stream)
        self._events[k] = ev
        return HexBatch(x=x_out, edge_index=ei_out, batch=batch, ptr=ptr, num_graphs=B)
