"""A few search waves in which EVERY game has a non-terminal leaf (two stones on the board, so no playout can end a
game): each launch of the net kernels then evaluates exactly --games leaves, which is what tools/ncu_traffic.py needs
to turn an ncu capture of trunk_tc_kernel into DRAM bytes per leaf on the SEARCH path (leaves staged from the leaf
bitboards, not from fp32 planes).
usage: ncu --set full --clock-control none -k regex:trunk_tc -s 2 -c 1 -o rep python tools/leaf_wave.py [--games 16384]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from alphazero_gomoku_b200 import Engine  # noqa: E402
from helpers import load_golden_weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--games", type=int, default=16384)
ap.add_argument("--waves", type=int, default=4)
a = ap.parse_args()
e = Engine(width=8, height=8, n_in_row=5, n_games=a.games, n_playout=a.waves)
e.set_weights(load_golden_weights("weights_8_8_5"))
for g in range(a.games):
    e.boards_set(g, [27 + (g % 3), 36 - (g % 5)])
e.search(a.waves)
e.sync()
c = e.counters()
print("leaf_evals", c["leaf_evals"], "=", a.waves, "x", a.games)
assert c["leaf_evals"] == a.waves * a.games
e.close()
