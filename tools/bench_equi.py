"""HBM roofline of the dihedral augmentation kernel (gz_equi_data) with device-resident tensors.
usage: python tools/bench_equi.py [n_samples] [board]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from alphazero_gomoku_b200 import Engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
w = int(sys.argv[2]) if len(sys.argv) > 2 else 8
S = w * w
e = Engine(width=w, height=w, n_in_row=3, n_games=1, n_playout=1, nodes_per_game=4)
st = (torch.rand((n, 4, w, w), device="cuda") < 0.3).float()
pi = torch.rand((n, S), device="cuda")
z = torch.zeros((n,), device="cuda")
for _ in range(3):
    e.equi_data(st, pi, z)
torch.cuda.synchronize()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e.profile_read(reset=True)
e.profile_enable(True)
ev0.record()
for _ in range(reps):
    out = e.equi_data(st, pi, z)   # allocates its outputs: includes torch.empty
ev1.record()
torch.cuda.synchronize()
e.profile_enable(False)
prof = e.profile_read(reset=True)
call_ms = ev0.elapsed_time(ev1) / reps       # whole call: output allocation, table upload, launch, stream sync
ms = prof["other"][0] / max(prof["other"][1], 1)  # the kernel alone (CUDA events around the launch)
byts = 9.0 * (5 * S + 1) * 4 * n
peak = 6554.2
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs_sustained", peak)
except Exception:
    pass
print(json.dumps({"kernel": "equi_gather_kernel", "samples": n, "board": w, "kernel_ms": ms, "call_ms": call_ms, "algorithmic_bytes": byts,
                  "achieved_gbs": byts / ms / 1e6, "peak_gbs": peak, "frac": byts / ms / 1e6 / peak,
                  "samples_per_s": n / ms * 1e3}))
