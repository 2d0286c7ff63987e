"""TEST / BENCH INPUTS (not part of the product package): the committed position fixtures under tests/golden/, played by
the reference's own C++ game compiled unchanged (tests/golden/make_fixtures.py, oracle/refgame/)."""
import os
from typing import List

import numpy as np
import torch

from gnn_hex_b200.data import HexGraph, graph_from_directed

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


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


def load_positions_cpp(name: str, swap_pattern: bool = True) -> List[HexGraph]:
    """The same committed positions in the C++ feature convention the HexAra network reads
    (cpp_hex/hex_graph_game/shannon_node_switching_game.cpp:690-726): x[:,0] = IS_TERMINAL exactly as the compiled reference
    game emitted it (`term`), x[:,1] / x[:,2] = the swap flags of move 0 / move 1.  The flags follow a fixed pattern over the
    graph index (g % 3 == 0: move 1 -> x[:,2] = 1; g % 3 == 1: move 0 -> x[:,1] = 1; else none) when `swap_pattern`."""
    path = name if os.path.exists(name) else os.path.join(GOLDEN_DIR, name)
    z = np.load(path)
    out, eo, no = [], 0, 0
    for g in range(len(z["n"])):
        ne, n = int(z["ne"][g]), int(z["n"][g])
        gr = graph_from_directed(n, z["src"][eo:eo + ne], z["dst"][eo:eo + ne], 0)
        gr.x.zero_()
        gr.x[:, 0] = torch.from_numpy(z["term"][no:no + n].astype(np.float32))
        if swap_pattern and g % 3 == 0:
            gr.x[:, 2] = 1.0
        elif swap_pattern and g % 3 == 1:
            gr.x[:, 1] = 1.0
        out.append(gr)
        eo += ne
        no += n
    return out
