"""Shared test helpers (test infrastructure; may import oracle/)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def hash_eval_batch(cells, cur_player, salt=0):
    """numpy twin of oracle/orc_mcts.c:orc_hash_eval for a batch.
    cells int8 [G,S] (0 empty, 1, 2), cur_player int8 [G] -> priors f32 [G,S], values f32 [G]."""
    cells = np.asarray(cells)
    G, S = cells.shape
    salt = np.uint64(salt)
    with np.errstate(over="ignore"):
        key = splitmix64(salt + np.uint64(1000) + cur_player.astype(np.uint64))
        m = np.arange(S, dtype=np.uint64)[None, :]
        stone = cells > 0
        h = splitmix64(salt + (m * np.uint64(2) + (cells.astype(np.int64) - 1).clip(0).astype(np.uint64)) + np.uint64(1))
        h = np.where(stone, h, np.uint64(0))
        key = key ^ np.bitwise_xor.reduce(h, axis=1)
        hw = splitmix64(key[:, None] + m + np.uint64(1))
        t = (hw >> np.uint64(56)).astype(np.uint32) + np.uint32(1)
        w = (t * t * t).astype(np.uint32)
        w = np.where(stone, np.uint32(0), w)
        s = w.sum(axis=1, dtype=np.uint64).astype(np.uint32)
        priors = w.astype(np.float32) / s.astype(np.float32)[:, None]
        vi = (splitmix64(key ^ np.uint64(0x5555555555555555)) % np.uint64(2001)).astype(np.int64) - 1000
        values = vi.astype(np.float32) / np.float32(1000.0)
    return priors.astype(np.float32), values.astype(np.float32)


def random_moves(rng, S, k):
    return [int(x) for x in rng.permutation(S)[:k]]


def load_golden_weights(name):
    """tests/golden/<name>.npz: 16 arrays in torch state_dict order/layout."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return [z["p%02d" % i] for i in range(16)]
