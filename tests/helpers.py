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


# ---- round-2 fixtures --------------------------------------------------------------------------------------
def seed0_net_arrays(width, height=None, trunk=(32, 64, 128)):
    """The reference Net's parameters under torch.manual_seed(0) (policy_value_net_pytorch.py:25-41; SURVEY
    8d C4/C5: no 9x9 / 15x15 weights ship), rebuilt WITHOUT the reference: the same layers constructed in
    the same order consume the same generator stream.  tests check them against the checksums that
    tests/golden/make_golden2.py stored from the live reference (`<tag>_wsum` in mcts2.npz)."""
    import torch
    from torch import nn
    height = width if height is None else height
    S = width * height
    c1, c2, c3 = trunk
    torch.manual_seed(0)
    layers = [nn.Conv2d(4, c1, kernel_size=3, padding=1), nn.Conv2d(c1, c2, kernel_size=3, padding=1),
              nn.Conv2d(c2, c3, kernel_size=3, padding=1), nn.Conv2d(c3, 4, kernel_size=1), nn.Linear(4 * S, S),
              nn.Conv2d(c3, 2, kernel_size=1), nn.Linear(2 * S, 64), nn.Linear(64, 1)]
    out = []
    for l in layers:
        out += [l.weight.detach().numpy().astype(np.float32).copy(), l.bias.detach().numpy().astype(np.float32).copy()]
    return out


def golden_net_arrays(tag, z=None):
    """Weights of an mcts2.npz configuration: shipped 6x6 / 8x8 fixtures, seed-0 reference Net for 9x9 / 15x15
    (checked against the stored checksums)."""
    if tag in ("6_6_4", "8_8_5"):
        return load_golden_weights("weights_" + tag)
    w = int(tag.split("_")[0])
    arrs = seed0_net_arrays(w)
    if z is None:
        z = np.load(os.path.join(GOLDEN, "mcts2.npz"))
    ws = np.array([[a.astype(np.float64).sum(), np.abs(a.astype(np.float64)).sum()] for a in arrs])
    assert np.allclose(ws, z[tag + "_wsum"], rtol=0, atol=1e-9), "seed-0 Net differs from the reference's (torch version?)"
    return arrs


class Mcts2:
    """Accessor for tests/golden/mcts2.npz (reference searches with per-playout decision traces)."""

    def __init__(self, tag):
        with np.load(os.path.join(GOLDEN, "mcts2.npz")) as z:  # read eagerly: NpzFile is not thread-safe
            self.z = {k: z[k] for k in z.files if k.startswith(tag + "_")}
        self.tag = tag
        self.w, _, self.n = [int(x) for x in tag.split("_")]
        self.S = self.w * self.w
        self.n_playout = int(self.z[tag + "_n_playout"])
        self.offsets = self.z[tag + "_offsets"]
        self.n_pos = len(self.offsets) - 1

    def moves(self, i):
        return [int(m) for m in self.z[self.tag + "_moves"][self.offsets[i]:self.offsets[i + 1]]]

    def visits(self, s, i):
        return self.z["%s_visits%d" % (self.tag, s)][i]

    def best_move(self, i):
        return int(self.z[self.tag + "_move"][i])

    def gap(self, s, i):
        """smallest (best - second) / max(1, |best|) over the reference's select calls"""
        return float(self.z["%s_gap%d" % (self.tag, s)][i])

    def margin(self, s, i):
        return float(self.z["%s_margin%d" % (self.tag, s)][i])

    def trace(self, s, i):
        """(tlen uint8 [n_playout], moves uint8 flat) of search s (1 or 2) at position i"""
        to = self.z["%s_troff%d" % (self.tag, s)]
        return self.z["%s_tlen%d" % (self.tag, s)][i], self.z["%s_trace%d" % (self.tag, s)][to[i]:to[i + 1]]

    def paths(self, s, i):
        tl, tr = self.trace(s, i)
        out, o = [], 0
        for d in tl:
            out.append([int(x) for x in tr[o:o + int(d)]])
            o += int(d)
        return out

    def has_search2(self, i):
        return bool((self.visits(2, i) >= 0).any())


def tv_distance(a, b):
    a = np.maximum(np.asarray(a, np.float64), 0)
    b = np.maximum(np.asarray(b, np.float64), 0)
    return 0.5 * np.abs(a / a.sum() - b / b.sum()).sum()


def oracle_arena(orc, net, w, n, az_playout, pure_playout, start_player, n_games, seed, c_puct=5):
    """The reference arena (train.py:216-276 -> game.py:211-290) on the oracle: MCTSPlayer (net, fresh tree
    every move, move ~ softmax(log(visits)/1e-3), mcts_alphaZero.py:244-255; off 15x15 the A-13 opening
    override is skipped) as player 1 vs mcts_pure.MCTSPlayer (most visited child, mcts_pure.py:261-270) as
    player 2.  Returns winners (1, 2, -1) per game."""
    rng = np.random.RandomState(seed)
    winners = []
    for g in range(n_games):
        b = orc.Board(w, w, n)
        b.init_board(start_player)
        ply = 0
        while True:
            if b.current_player == 1:
                m = orc.MCTS(c_puct, az_playout, orc.EVAL_NET, net=net)
                acts, probs = m.get_move_probs(b, temp=1e-3)
                move = int(rng.choice(acts, p=probs))
            else:
                m = orc.MCTS(c_puct, pure_playout, orc.EVAL_PURE, salt=seed)
                m.set_stream(g, 1000 * start_player + ply)
                move = m.pure_get_move(b)
            b.do_move(move)
            ply += 1
            end, winner = b.game_end()
            if end:
                winners.append(winner)
                break
    return np.array(winners, np.int8)


def binomial_band_ok(k1, n1, k2, n2, sigmas=4.0):
    """two-sample test of proportions with a pooled, Laplace-smoothed variance (robust at p = 0 or 1)"""
    p = (k1 + k2 + 1.0) / (n1 + n2 + 2.0)
    sd = np.sqrt(p * (1 - p) * (1.0 / n1 + 1.0 / n2))
    return abs(k1 / n1 - k2 / n2) <= sigmas * sd + 1e-12


def chi_square_two_sample(a, b, what, min_expected=8.0):
    """chi-square homogeneity test of two histograms; bins are pooled until both expect >= min_expected"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    order = np.argsort(-(a + b))
    bins, cur = [], np.zeros(2)
    for i in order:
        cur += (a[i], b[i])
        tot = cur.sum()
        if min(tot * a.sum(), tot * b.sum()) / (a.sum() + b.sum()) >= min_expected:
            bins.append(cur)
            cur = np.zeros(2)
    if cur.sum() > 0:
        if bins:
            bins[-1] = bins[-1] + cur
        else:
            bins.append(cur)
    t = np.array(bins)
    dof = len(t) - 1
    if dof < 1:
        return
    exp = t.sum(axis=1, keepdims=True) * t.sum(axis=0, keepdims=True) / t.sum()
    chi = ((t - exp) ** 2 / exp).sum()
    # p ~ 1e-5 bound: chi < dof + 4.3 * sqrt(2 dof) + 12 (Wilson-Hilferty-ish, loose on purpose)
    assert chi < dof + 4.3 * np.sqrt(2 * dof) + 12, (what, chi, dof, t.tolist())
    print("%s: chi2 = %.1f at %d dof" % (what, chi, dof))


def dihedral_images(state, pi, z, width, height):
    """Numpy checker for the 8-fold augmentation (reference train.py:116-135, SURVEY A-19), test infrastructure:
    for k = 1..4 quarter turns, without and then with a left-right mirror, the image T of every state plane, and
    for pi - which is indexed by raw move id while the planes are stored row-reversed (A-3) - the same T applied
    between two up-down flips.  Pinned against the live reference's output in tests/golden/equi.npz."""
    out = []
    grid = np.asarray(pi).reshape(height, width)[::-1]
    for k in (1, 2, 3, 4):
        for mirror in (False, True):
            def T(a):
                a = np.rot90(a, k)
                return a[:, ::-1] if mirror else a
            out.append((np.array([T(p) for p in state]), T(grid)[::-1].flatten(), z))
    return out
