"""Does the tree kernel of one engine overlap the trunk kernel of another?  One engine of G games on
one stream against K engines of G / K games on K streams of the same GPU, the waves of the K engines
interleaved on the host in chunks (the launch queue is finite).  The trunk kernel keeps every SM's
shared memory; a 128-thread tree block (16 B static) fits beside it only just.
usage: python tools/overlap_engines.py [games] [playouts] [chunk]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import numpy as np
import torch
from helpers import load_golden_weights
from alphazero_gomoku_b200 import Engine

G = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
P = int(sys.argv[2]) if len(sys.argv) > 2 else 400
CH = int(sys.argv[3]) if len(sys.argv) > 3 else 10
wts = load_golden_weights("weights_8_8_5")


def run(K):
    streams = [torch.cuda.Stream() for _ in range(K)]
    es = [Engine(width=8, height=8, n_in_row=5, n_games=G // K, n_playout=P, seed=1 + k, stream=s.cuda_stream)
          for k, s in enumerate(streams)]
    for e in es:
        e.set_weights(wts)
    res = []
    for it in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for c in range(0, P, CH):
            for e in es:
                e.search(min(CH, P - c))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        res.append(G * P / dt / 1e6)
        for e in es:
            e.update_with_move(np.full(G // K, -1, np.int32))
    for e in es:
        e.close()
    return res


for K in (1, 2, 3, 4):
    print("engines=%d x %d games: %s M playouts/s" % (K, G // K, ["%.2f" % r for r in run(K)]), flush=True)
