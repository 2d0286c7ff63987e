"""Generates tests/golden/pos_transfer_14_19_seed1234.npz: positions on 14x14 .. 19x19 boards (the board-size transfer
evaluation of the reference, GN0/RainbowDQN/evaluate_elo.py:228-251, GN0/util/plotting.py:61-80), played by the reference's
own C++ game (oracle/_ref/hexgen) -- the empty board (first position of a game) plus seeded mid-game positions per size.
Same layout as make_fixtures.py.  Authoring container only:
    make -C oracle/refgame && python tests/golden/make_transfer_fixtures.py
"""
import sys

import numpy as np

import make_fixtures as mf


def main():
    rng = np.random.RandomState(1234)
    out = []
    for size in (14, 16, 19):
        pos = mf.play(size, 3, 1234 + size)
        empty = max(pos, key=lambda p: p["n"])
        assert empty["n"] == size * size + 2
        out.append(empty)
        rest = [p for p in pos if p["n"] > 128 and p is not empty]
        out += [rest[int(i)] for i in rng.choice(len(rest), size=3, replace=False)]
    mf.save("pos_transfer_14_19_seed1234.npz", out)


if __name__ == "__main__":
    sys.exit(main())
