"""Host-side data plumbing of the hot path: Hex position graphs in the reference's tensor layout,
the PyG-style collate, and the committed position fixtures.

Layout mirrors `convert_node_switching_game(..., old_style=True)` (GN0/util/convert_graph.py:77-85,
99-117): x[n,3] = [degree, is_terminal (rows 0,1), maker flag], edge_index = cat(E, flip(E)) int64,
and `Batch.from_data_list` / C++ `collate_batch` (cpp_hex/hex_graph_game/util.cpp:22-41).
"""
import os
from types import SimpleNamespace
from typing import List, Sequence

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


class HexGraph(SimpleNamespace):
    """x [n,3] float32, edge_index [2,E] int64 (symmetric)."""

    def to(self, device, non_blocking=False):
        return HexGraph(x=self.x.to(device, non_blocking=non_blocking),
                        edge_index=self.edge_index.to(device, non_blocking=non_blocking))


class HexBatch(SimpleNamespace):
    """x, edge_index, batch, ptr, num_graphs -- the fields of a PyG `Batch` the model reads."""

    def to(self, device, non_blocking=False):
        return HexBatch(x=self.x.to(device, non_blocking=non_blocking),
                        edge_index=self.edge_index.to(device, non_blocking=non_blocking),
                        batch=self.batch.to(device, non_blocking=non_blocking),
                        ptr=self.ptr.to(device, non_blocking=non_blocking), num_graphs=self.num_graphs)

    def pin_memory(self):
        return HexBatch(x=self.x.pin_memory(), edge_index=self.edge_index.pin_memory(),
                        batch=self.batch.pin_memory(), ptr=self.ptr.pin_memory(), num_graphs=self.num_graphs)


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


def load_positions(name: str, maker=None) -> List[HexGraph]:
    """Load tests/golden/<name>.npz (positions played by the reference's own C++ game)."""
    path = name if os.path.exists(name) else os.path.join(GOLDEN_DIR, name)
    z = np.load(path)
    out, eo = [], 0
    for g in range(len(z["n"])):
        ne = int(z["ne"][g])
        m = int(z["maker"][g]) if maker is None else int(maker)
        out.append(graph_from_directed(int(z["n"][g]), z["src"][eo:eo + ne], z["dst"][eo:eo + ne], m))
        eo += ne
    return out


def empty_board(size: int, maker: int = 1) -> HexGraph:
    """Empty size x size board (graph_game/hex_board_game.py:214-232, red graph), from the
    committed reference dump for sizes 3..13."""
    z = np.load(os.path.join(GOLDEN_DIR, "empty_boards_3_13.npz"))
    sizes = list(z["size"])
    if size not in sizes:
        raise ValueError(f"no committed empty board of size {size}")
    g = sizes.index(size)
    eo = int(z["ne"][:g].sum())
    ne = int(z["ne"][g])
    return graph_from_directed(int(z["n"][g]), z["src"][eo:eo + ne], z["dst"][eo:eo + ne], maker)


def collate(graphs: Sequence[HexGraph]) -> HexBatch:
    """`Batch.from_data_list` for (x, edge_index): host-side, all int64 (util.cpp:22-41)."""
    ns = [int(g.x.shape[0]) for g in graphs]
    ptr = torch.zeros(len(graphs) + 1, dtype=torch.long)
    ptr[1:] = torch.cumsum(torch.tensor(ns, dtype=torch.long), 0)
    x = torch.cat([g.x for g in graphs], 0)
    edge_index = torch.cat([g.edge_index + int(o) for g, o in zip(graphs, ptr[:-1])], 1)
    batch = torch.repeat_interleave(torch.arange(len(graphs), dtype=torch.long), torch.tensor(ns, dtype=torch.long))
    return HexBatch(x=x, edge_index=edge_index, batch=batch, ptr=ptr, num_graphs=len(graphs))
