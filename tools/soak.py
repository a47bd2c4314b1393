"""Soak run: full-size self-play for many plies with slot recycling; checks every drained game
against the oracle rules and the engine's own counters.  usage: python tools/soak.py [plies] [games]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from alphazero_gomoku_b200 import Engine  # noqa: E402
from helpers import load_golden_weights   # noqa: E402
from oracle import orc                    # noqa: E402

plies = int(sys.argv[1]) if len(sys.argv) > 1 else 200
G = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
w, n, P = 8, 5, 400
e = Engine(width=w, height=w, n_in_row=n, n_games=G, n_playout=P, seed=11)
e.set_weights(load_golden_weights("weights_8_8_5"))
e.selfplay_begin(temp=1.0, recycle=True)
t0 = time.time()
n_games = n_samples = checked = 0
for step in range(0, plies, 8):
    e.selfplay_step(min(8, plies - step))
    games = e.selfplay_drain_games()
    data = e.selfplay_drain()
    n_games += len(games)
    n_samples += len(data["zs"])
    assert len(data["zs"]) == sum(len(m) for _, _, m in games)
    assert np.isfinite(data["pis"]).all() and abs(data["pis"].sum(axis=1) - 1).max() < 1e-4 if len(data["zs"]) else True
    for gid, win, mv in games[:40]:
        b = orc.Board(w, w, n)
        b.init_board(0)
        for m in mv:
            assert not b.game_end()[0]
            b.do_move(m)
        assert b.game_end() == (True, win)
        checked += 1
st = e.selfplay_stats()
dt = time.time() - t0
print("soak ok: %d plies x %d slots in %.1f s, %d games (%d replayed through the oracle), %d samples, "
      "pool_overflows %d, max_tree_nodes %d of %d, %.1f M playouts/s"
      % (plies, G, dt, n_games, checked, n_samples, st["pool_overflows"], st["max_tree_nodes"], e.nodes_per_game,
         st["n_playouts"] / dt / 1e6))
assert st["pool_overflows"] == 0 and st["n_plies"] == plies * G
e.close()
