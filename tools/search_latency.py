import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np
from helpers import load_golden_weights
from alphazero_gomoku_b200 import Engine
for G in (1, 64, 1024):
    e = Engine(width=8, height=8, n_in_row=5, n_games=G, n_playout=400, seed=1)
    e.set_weights(load_golden_weights("weights_8_8_5"))
    e.search(400); e.sync()
    e.update_with_move(np.full(G, -1, np.int32))
    t0 = time.perf_counter()
    for _ in range(5):
        e.search(400); e.sync()
        e.update_with_move(np.full(G, -1, np.int32))
    dt = (time.perf_counter() - t0) / 5
    print("G=%d: %.2f ms per 400-playout search (%.1f us per wave), %.0f playouts/s" % (G, dt * 1e3, dt / 400 * 1e6, G * 400 / dt))
    e.close()
# device time per kernel class at G = 1 (CUDA events around every launch; the events themselves add gaps)
e = Engine(width=8, height=8, n_in_row=5, n_games=1, n_playout=400, seed=1)
e.set_weights(load_golden_weights("weights_8_8_5"))
e.search(400); e.sync()
e.update_with_move(np.full(1, -1, np.int32))
e.profile_read(reset=True)
e.profile_enable(True, stride=1)
e.search(400); e.sync()
e.profile_enable(False)
print({k: (round(v[0] / max(v[1], 1) * 1e3, 2), v[1]) for k, v in e.profile_read().items() if v[1]}, "(us per launch, launches)")
e.close()
