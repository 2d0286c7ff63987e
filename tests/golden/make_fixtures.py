"""Generates tests/golden/pos_*.npz from the reference's own C++ game (oracle/_ref/hexgen,
built by oracle/refgame/Makefile from /root/reference/cpp_hex/hex_graph_game/*.cpp unchanged).

Run in the authoring container only (the GPU box has no /root/reference):
    make -C oracle/refgame && python tests/golden/make_fixtures.py

Each fixture holds G positions of seeded uniform-random playouts with dead/captured removal
(the loop of cpp_hex/hex_graph_game/tests/speedtest.cpp:33-49, srand(1234)), sampled uniformly
without replacement (numpy RandomState(1234)) from all non-terminal positions:
    n[G]        nodes per position (terminals are rows 0,1)
    ne[G]       directed edges per position, exactly as convert_graph emitted them
    maker[G]    1 if the maker is on turn
    src, dst    concatenated LOCAL node ids (uint16) of every directed edge, in the reference's
                CSR order (source-major, target ascending: graph.cpp:16-35)
    term        concatenated IS_TERMINAL column (uint8) as emitted (x[:,0] of the C++ convention)
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
HEXGEN = os.path.join(ROOT, "oracle", "_ref", "hexgen")


def play(size, games, seed):
    out = subprocess.run([HEXGEN, str(size), str(games), str(seed)], check=True, capture_output=True, text=True).stdout
    lines = out.split("\n")
    pos = []
    i = 0
    while i < len(lines):
        if not lines[i].startswith("P "):
            i += 1
            continue
        _, g, mv, n, E, maker = lines[i].split()
        n, E = int(n), int(E)
        term = np.array(lines[i + 1].split(), dtype=np.uint8)
        src = np.array(lines[i + 2].split(), dtype=np.int64)
        dst = np.array(lines[i + 3].split(), dtype=np.int64)
        assert term.size == n and src.size == E and dst.size == E
        pos.append(dict(size=size, n=n, maker=int(maker), term=term, src=src, dst=dst))
        i += 4
    return pos


def save(name, positions):
    np.savez_compressed(
        os.path.join(HERE, name),
        n=np.array([p["n"] for p in positions], dtype=np.int32),
        ne=np.array([p["src"].size for p in positions], dtype=np.int32),
        maker=np.array([p["maker"] for p in positions], dtype=np.uint8),
        size=np.array([p["size"] for p in positions], dtype=np.uint8),
        src=np.concatenate([p["src"] for p in positions]).astype(np.uint16),
        dst=np.concatenate([p["dst"] for p in positions]).astype(np.uint16),
        term=np.concatenate([p["term"] for p in positions]).astype(np.uint8),
    )
    ns = np.array([p["n"] for p in positions])
    es = np.array([p["src"].size for p in positions])
    print(f"{name}: G={len(positions)} nodes mean {ns.mean():.1f} max {ns.max()} edges mean {es.mean():.1f} max {es.max()}")


def sample(pos, G, rng):
    idx = rng.choice(len(pos), size=G, replace=False)
    return [pos[i] for i in idx]


def main():
    rng = np.random.RandomState(1234)
    p11 = play(11, 100, 1234)
    save("pos_11x11_seed1234.npz", sample(p11, 256, rng))
    p7 = play(7, 100, 1234)
    save("pos_7x7_seed1234.npz", sample(p7, 256, rng))
    # empty boards straight from the reference game: pins hex_board_undirected_edges()
    empties = [play(s, 1, 1)[0] for s in range(3, 14)]
    save("empty_boards_3_13.npz", empties)
    # ragged transfer mix (BASELINE config 4): board size uniform in 5..13, then a position
    per_size = {s: play(s, 40, 1234 + s) for s in range(5, 14)}
    mix = []
    for _ in range(256):
        s = int(rng.randint(5, 14))
        mix.append(per_size[s][int(rng.randint(len(per_size[s])))])
    save("pos_mix_5_13_seed1234.npz", mix)


if __name__ == "__main__":
    sys.exit(main())
