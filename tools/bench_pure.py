"""Pure-MCTS (mcts_pure.py) throughput: BASELINE config C1's geometry - 6x6 four-in-a-row, 1000
random-rollout playouts per move - on the GPU over many concurrent games, with the CPU oracle port of
the same search on one host thread beside it (the reference runs C1 as a single game on the CPU).
usage: python tools/bench_pure.py [games] [playouts] [board] [n_in_row]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphazero_gomoku_b200 import Engine  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
w = int(sys.argv[3]) if len(sys.argv) > 3 else 6
n = int(sys.argv[4]) if len(sys.argv) > 4 else 4

stream = torch.cuda.Stream()
e = Engine(width=w, height=w, n_in_row=n, n_games=G, n_playout=P, seed=0, stream=stream.cuda_stream)
# mid-game positions: 4 seeded random plies in every game (rollouts from the empty board take the
# reference's hard-coded rows/cols 5..9 candidate box, which does not exist on a 6x6 board)
rng = np.random.RandomState(0)
for g in range(G):
    e.boards_set(g, [int(x) for x in rng.permutation(w * w)[:4]])
e.pure_search(50, search_id=0)   # warm-up
e.sync()
e.counters(reset=True)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ev0.record(stream)
e.pure_search(P, search_id=1)
ev1.record(stream)
torch.cuda.synchronize()
ms = ev0.elapsed_time(ev1)
c = e.counters(reset=True)
out = {"config": "pure MCTS %dx%d %d-in-a-row, %d playouts/move, %d concurrent games" % (w, w, n, P, G),
       "ms": ms, "playouts_per_s": c["playouts"] / ms * 1e3, "rollout_plies_per_s": c["rollout_plies"] / ms * 1e3,
       "rollout_plies_per_playout": c["rollout_plies"] / max(c["playouts"], 1)}
e.close()

try:  # CPU side: the oracle port of mcts_pure on one thread, one game, same position recipe
    from oracle import orc
    b = orc.Board(w, w, n)
    b.init_board(0)
    for m in rng.permutation(w * w)[:4]:
        b.do_move(int(m))
    m = orc.MCTS(5, P, orc.EVAL_PURE)
    m.set_stream(0, 1)
    t0 = time.perf_counter()
    m.pure_get_move(b)
    dt = time.perf_counter() - t0
    out["cpu_port_playouts_per_s"] = P / dt
    out["cpu_port_cores"] = 1
    del m
except Exception as ex:  # the tool is about the GPU number; say why the CPU leg is missing
    out["cpu_port_error"] = repr(ex)
print(json.dumps(out))
