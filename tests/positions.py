"""TEST / BENCH INPUTS (not part of the product package): the committed position fixtures under tests/golden/, played by
the reference's own C++ game compiled unchanged (tests/golden/make_fixtures.py, oracle/refgame/)."""
import os
from typing import List

import numpy as np

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
