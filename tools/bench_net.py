"""Time / profile the net forward alone: python tools/bench_net.py --board 15 --batch 2048 --trunk 64 128 256
(CUDA events around `iters` forwards of a device-resident batch; used under ncu for the per-kernel view)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from alphazero_gomoku_b200 import Engine
    from alphazero_gomoku_b200.engine import weight_shapes
    ap = argparse.ArgumentParser()
    ap.add_argument("--board", type=int, default=15)
    ap.add_argument("--batch", type=int, default=2048)
    ap.add_argument("--trunk", type=int, nargs=3, default=[64, 128, 256])
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    w, B, trunk = a.board, a.batch, tuple(a.trunk)
    rng = np.random.RandomState(0)
    arrs = [(rng.randn(*s) * (np.sqrt(2.0 / int(np.prod(s[1:]))) if len(s) > 1 else 0.1)).astype(np.float32)
            for s in weight_shapes(w, w, trunk)]
    stream = torch.cuda.Stream()
    e = Engine(width=w, height=w, n_in_row=min(w, 5), n_games=1, trunk=trunk, stream=stream.cuda_stream)
    e.set_weights(arrs)
    planes = torch.from_numpy((rng.rand(B, 4, w, w) < 0.3).astype(np.float32)).cuda()
    probs = torch.empty((B, w * w), device="cuda")
    vals = torch.empty((B,), device="cuda")
    for _ in range(3):
        e.net_forward(planes, probs, vals)
    e.sync()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(a.iters):
        e.net_forward(planes, probs, vals)
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / a.iters
    c1, c2, c3 = trunk
    S = w * w
    fl = 2.0 * S * (36 * c1 + 9 * c1 * c2 + 9 * c2 * c3 + 6 * c3 + 4 * S + 128) + 128
    print("net forward %dx%d trunk %s batch %d: %.3f ms, %.1f TFLOP/s algorithmic, %.2f M leaves/s"
          % (w, w, trunk, B, ms, B * fl / ms / 1e9, B / ms / 1e3))


if __name__ == "__main__":
    main()
