"""Profiling aid: per-phase clock64 timeline of CTA 0 of the tcgen05 trunk (GZ_TC_TIMELINE)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/tc_timeline.txt"
w = int(sys.argv[2]) if len(sys.argv) > 2 else 8
B = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
if os.path.exists(out):
    os.remove(out)
os.environ["GZ_TC_TIMELINE"] = out
from alphazero_gomoku_b200 import Engine
from alphazero_gomoku_b200.engine import weight_shapes

rng = np.random.RandomState(0)
arrs = [(rng.randn(*s) * 0.05).astype(np.float32) for s in weight_shapes(w, w)]
e = Engine(width=w, height=w, n_in_row=3, net_impl=2)
e.set_weights(arrs)
planes = (rng.rand(B, 4, w, w) < 0.3).astype(np.float32)
import torch
xp = torch.from_numpy(planes).cuda()
for _ in range(3):
    e.net_forward(xp)
e.sync()
names = ["conv2_mma", "ep2+stage_next", "conv1_next+ep1", "wait_conv3_issue"]
rows = []
for blk in open(out).read().split("launch")[1:]:
    lines = [l for l in blk.strip().splitlines()[1:] if l.strip()]
    t = np.array([[int(x) for x in l.split()] for l in lines if len(l.split()) == 4])
    rows.append(t)
t = rows[-1]
d = np.diff(t.reshape(-1))
d = np.append(d, 0).reshape(-1, 4)[:-1]  # drop the last (partial) group
print("groups", len(t), "cycles/group", (t[-1, 0] - t[0, 0]) / (len(t) - 1))
for i, n in enumerate(names):
    print("%-10s mean %8.0f  min %8.0f  max %8.0f" % (n, d[1:, i].mean(), d[1:, i].min(), d[1:, i].max()))
