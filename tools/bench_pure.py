"""Pure-MCTS (mcts_pure.py) throughput: BASELINE config C1's geometry - 6x6 four-in-a-row, 1000
random-rollout playouts per move - on the GPU over many concurrent games, with the CPU oracle port of
the same search beside it (the reference runs C1 as a single game on the CPU: BASELINE.md B1).
Also times the arena's use of it: 64 games x 3000 rollouts per move (train.py:249, 341-342).
usage: python tools/bench_pure.py [games] [playouts] [board] [n_in_row]   -> one JSON line"""
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


def timed_search(games, playouts, warm=50):
    stream = torch.cuda.Stream()
    e = Engine(width=w, height=w, n_in_row=n, n_games=games, n_playout=playouts, seed=0, stream=stream.cuda_stream)
    # mid-game positions: 4 seeded random plies in every game (rollouts from the empty board take the
    # reference's hard-coded rows/cols 5..9 candidate box, which does not exist on a 6x6 board)
    rng = np.random.RandomState(0)
    for g in range(games):
        e.boards_set(g, [int(x) for x in rng.permutation(w * w)[:4]])
    e.pure_search(warm, search_id=0)
    e.update_with_move(np.full(games, -1, np.int32))
    e.sync()
    e.counters(reset=True)
    e.profile_enable(True, stride=8)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    e.pure_search(playouts, search_id=1)
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    c = e.counters(reset=True)
    prof = {k: v for k, v in e.profile_read().items() if v[1]}
    e.close()
    return ms, c, prof


ms, c, prof = timed_search(G, P)
n_sm = torch.cuda.get_device_properties(0).multi_processor_count
out = {"metric": "pure_mcts_playouts_per_sec", "unit": "playouts/s",
       "config": {"workload": "pure MCTS %dx%d %d-in-a-row, %d playouts/move, %d concurrent games" % (w, w, n, P, G)},
       "ms": ms, "value": c["playouts"] / ms * 1e3, "rollout_plies_per_s": c["rollout_plies"] / ms * 1e3,
       "rollout_plies_per_playout": c["rollout_plies"] / max(c["playouts"], 1),
       "kernel_ms": {k: round(v[0], 3) for k, v in prof.items()},
       "kernel_launches": {k: v[1] for k, v in prof.items()}}
if "rollout" in prof and prof["rollout"][0] > 0:
    # the rollout kernel alone: plies/s per SM against the SM clock = cycles one SM spends per rollout ply
    rp = c["rollout_plies"] / (prof["rollout"][0] * 1e-3)
    out["rollout_kernel"] = {"plies_per_s": rp, "ns_per_ply_per_sm": 1e9 * n_sm / rp}

# the arena's shape: few games, many rollouts per move
ams, ac, _ = timed_search(64, 3000, warm=20)
out["arena_64x3000"] = {"ms_per_move": ams, "playouts_per_s": ac["playouts"] / ams * 1e3}

try:  # CPU side: the oracle port of mcts_pure, one game per host thread, same position recipe
    from concurrent.futures import ThreadPoolExecutor
    from oracle import orc
    cores = os.cpu_count() or 1
    rng = np.random.RandomState(0)
    b = orc.Board(w, w, n)
    b.init_board(0)
    for m in rng.permutation(w * w)[:4]:
        b.do_move(int(m))

    def one(i):
        m = orc.MCTS(5, P, orc.EVAL_PURE)
        m.set_stream(i, 1)
        t0 = time.perf_counter()
        m.pure_get_move(b)
        return time.perf_counter() - t0
    single = one(0)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as pool:  # ctypes releases the GIL
        list(pool.map(one, range(4 * cores)))
    dt = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": 4 * cores * P / dt, "unit": "playouts/s", "cores": cores, "kind": "port",
                           "sample": "%d searches of %d playouts, one game per thread" % (4 * cores, P),
                           "single_thread": P / single}
except Exception as ex:  # the tool is about the GPU number; say why the CPU leg is missing
    out["cpu_baseline_error"] = repr(ex)
print(json.dumps(out))
