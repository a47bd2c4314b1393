"""HBM roofline of the stateless rules kernel (gz_rules_eval: Board.do_move / game_end / availables /
current_state / get_move_target over positions given as move lists), device-resident tensors.
usage: python tools/bench_rules.py [n_positions] [board] [n_in_row]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphazero_gomoku_b200 import Engine  # noqa: E402
from alphazero_gomoku_b200.engine import _ptr  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
w = int(sys.argv[2]) if len(sys.argv) > 2 else 8
k = int(sys.argv[3]) if len(sys.argv) > 3 else 5
S = w * w
stream = torch.cuda.Stream()
e = Engine(width=w, height=w, n_in_row=k, n_games=1, n_playout=1, nodes_per_game=4, stream=stream.cuda_stream)
rng = np.random.RandomState(0)
lens = rng.randint(0, S // 2 + 1, size=n)                       # 0 .. S/2 random legal plies per position
offs = np.zeros(n + 1, np.int64)
offs[1:] = np.cumsum(lens)
perm = np.argsort(rng.rand(n, S), axis=1).astype(np.int32)      # a random permutation of the cells per position
flat = perm[np.arange(S)[None, :] < lens[:, None]]              # its first lens[i] entries
dev = "cuda"
d_moves, d_offs = torch.from_numpy(flat).to(dev), torch.from_numpy(offs).to(dev)
out = {"status": torch.empty(n, dtype=torch.int32, device=dev), "ended": torch.empty(n, dtype=torch.int8, device=dev),
       "winner": torch.empty(n, dtype=torch.int8, device=dev), "has_win": torch.empty(n, dtype=torch.int8, device=dev),
       "legal": torch.empty((n, S), dtype=torch.uint8, device=dev),
       "planes": torch.empty((n, 4, w, w), dtype=torch.float32, device=dev),
       "target": torch.empty((n, S), dtype=torch.uint8, device=dev), "cur": torch.empty(n, dtype=torch.int8, device=dev)}


def call():
    e._ck(e.L.gz_rules_eval(e._h, _ptr(d_moves), _ptr(d_offs), n, None, _ptr(out["status"]), _ptr(out["ended"]),
                            _ptr(out["winner"]), _ptr(out["has_win"]), _ptr(out["legal"]), _ptr(out["planes"]),
                            _ptr(out["target"]), _ptr(out["cur"])))


for _ in range(3):
    call()
torch.cuda.synchronize()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e.profile_read(reset=True)
e.profile_enable(True)
ev0.record(stream)
for _ in range(reps):
    call()
ev1.record(stream)
torch.cuda.synchronize()
e.profile_enable(False)
prof = e.profile_read(reset=True)
call_ms = ev0.elapsed_time(ev1) / reps            # whole call (reads offsets[n] back first, stream sync at the end)
ms = prof["other"][0] / max(prof["other"][1], 1)  # the kernel alone (CUDA events around the launch)
byts = float(flat.size * 4 + (n + 1) * 8 + n * (4 + 1 + 1 + 1 + 1) + n * S * (1 + 16 + 1))
peak = 6457.4
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak)
except Exception:
    pass
assert int(out["status"].abs().sum()) == 0
print(json.dumps({"kernel": "rules_eval_kernel", "positions": n, "board": w, "mean_plies": float(lens.mean()),
                  "kernel_ms": ms, "call_ms": call_ms, "algorithmic_bytes": byts, "achieved_gbs": byts / ms / 1e6, "peak_gbs": peak,
                  "frac": byts / ms / 1e6 / peak, "positions_per_s": n / ms * 1e3,
                  "finished_positions": int(out["ended"].sum())}))
