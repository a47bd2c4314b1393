"""Latency of one search for few games (the reference's interactive use: MCTSPlayer.get_action for ONE game,
human_play.py:98-140; the arena's 64 games), sequential and with K leaves in flight (gz_search_vl).
usage: python tools/search_latency.py [board] [n_in_row] [playouts]   -> one JSON line"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from alphazero_gomoku_b200 import Engine  # noqa: E402
from helpers import load_golden_weights  # noqa: E402

w = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 5
P = int(sys.argv[3]) if len(sys.argv) > 3 else 400
arrs = load_golden_weights("weights_%d_%d_%d" % (w, w, n))
out = {"config": "%dx%d/%d, %d playouts per search, shipped weights, fresh tree every search" % (w, w, n, P), "ms_per_search": {}}
for G in (1, 64, 1024):
    e = Engine(width=w, height=w, n_in_row=n, n_games=G, n_playout=P)
    e.set_weights(arrs)
    for g in range(G):
        e.boards_set(g, [27, 28])
    row = {}
    for K in (1, 2, 4, 8, 16):
        ts = []
        for rep in range(4):
            e.update_with_move(np.full(G, -1, np.int32))
            e.sync()
            t0 = time.perf_counter()
            e.search(P, leaves_in_flight=K)
            e.sync()
            ts.append((time.perf_counter() - t0) * 1e3)
        row["K=%d" % K] = round(min(ts[1:]), 3)
    out["ms_per_search"]["%d games" % G] = row
    e.close()
print(json.dumps(out))
