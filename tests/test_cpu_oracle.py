"""CPU tests (no GPU): the oracle against the committed golden vectors generated from the live
reference (tests/golden/make_golden.py), and against the reference itself when /root/reference
is present (build container only)."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, chi_square_two_sample, load_golden_weights

from oracle import orc
from oracle import ref_harness as rh

RULE_TAGS = [(6, 6, 4), (8, 8, 5), (9, 9, 5), (15, 15, 5), (3, 3, 3), (7, 7, 3), (8, 6, 4), (11, 11, 5)]


def _play(w, h, n, moves, start=0):
    b = orc.Board(w, h, n)
    b.init_board(int(start))
    for m in moves:
        b.do_move(int(m))
    return b


@pytest.mark.parametrize("w,h,n", RULE_TAGS)
def test_oracle_rules_vs_golden(w, h, n):
    z = np.load(os.path.join(GOLDEN, "rules.npz"))
    tag = "%dx%dx%d" % (w, h, n)
    moves, offs = z[tag + "_moves"], z[tag + "_offsets"]
    S = w * h
    for i in range(len(offs) - 1):
        b = _play(w, h, n, moves[offs[i]:offs[i + 1]], z[tag + "_start"][i])
        end, winner = b.game_end()
        assert (int(end), winner) == (int(z[tag + "_ended"][i]), int(z[tag + "_winner"][i]))
        assert int(b.has_a_winner()[0]) == int(z[tag + "_haswin"][i])
        legal = np.zeros(S, np.uint8); legal[b.availables] = 1
        assert np.array_equal(legal, z[tag + "_legal"][i])
        b.get_move_target()
        tg = np.zeros(S, np.uint8); tg[b.move_target] = 1
        assert np.array_equal(tg, z[tag + "_target"][i])
        assert b.get_current_player() == int(z[tag + "_cur"][i])
        if w == h:
            assert np.array_equal(b.current_state().astype(np.float32), z[tag + "_planes"][i])


@pytest.mark.parametrize("tag,w", [("g15", 15), ("g9", 9)])
def test_oracle_corpus_replay(tag, w):
    """Every ply of recorded reference games (logs/steps.log, human_gomoku.txt, run.log)."""
    z = np.load(os.path.join(GOLDEN, "corpus.npz"))
    moves, offs = z[tag + "_moves"], z[tag + "_offsets"]
    for g in range(len(offs) - 1):
        b = orc.Board(w, w, 5)
        b.init_board(0)
        for k in range(offs[g], offs[g + 1]):
            b.do_move(int(moves[k]))
            end, winner = b.game_end()
            assert (int(end), winner) == (int(z[tag + "_ended"][k]), int(z[tag + "_winner"][k]))


@pytest.mark.parametrize("tag,w", [("6_6_4", 6), ("8_8_5", 8)])
def test_oracle_net_vs_golden(tag, w):
    """C forward vs the reference PyTorch fp32 Net on the shipped weights: 1e-4 absolute."""
    z = np.load(os.path.join(GOLDEN, "net.npz"))
    net = orc.Net(w, w, load_golden_weights("weights_" + tag))
    lp, v = net.forward(z[tag + "_planes"])
    assert np.abs(np.exp(lp) - z[tag + "_probs"]).max() < 1e-4
    assert np.abs(v - z[tag + "_values"]).max() < 1e-4


# positions of tests/golden/mcts2.npz whose every decision is replayed on the CPU here (the C net costs 1-5 ms per
# leaf); GZ_FULL_GOLDEN=1 replays every position of the fixture (64 / 64 / 32 / 16).  The free-running comparison
# below always uses the whole fixture.
MCTS2_REPLAY = {"6_6_4": 8, "8_8_5": 8, "9_9_5": 4, "15_15_5": 2}
NEAR_TIE = 1e-5  # |score1 - score2| <= 1e-5 * max(1, |score|) on the reference's path


def _threads(fn, items):
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 4)) as pool:  # ctypes calls release the GIL
        return list(pool.map(fn, items))


@pytest.mark.parametrize("tag", ["6_6_4", "8_8_5", "9_9_5", "15_15_5"])
def test_oracle_replays_reference_decisions(tag):
    """Decision-level pin of the oracle's tree arithmetic: EVERY select call the reference made in two
    consecutive searches per position (400-800 playouts each, tree reuse, mcts_alphaZero.py:154-192) is
    replayed through the oracle tree.  The recorded move must be the oracle's own argmax or lose to it by
    no more than the App. B near-tie bound (the C net is ~1e-6 from torch), an exact score tie must have
    been broken towards the first child, and the replayed tree must end with the reference's visit counts."""
    from helpers import Mcts2, golden_net_arrays
    fx = Mcts2(tag)
    net = orc.Net(fx.w, fx.w, golden_net_arrays(tag, fx.z))

    def run(i):
        b = _play(fx.w, fx.w, fx.n, fx.moves(i))
        m = orc.MCTS(5, fx.n_playout, orc.EVAL_NET, net=net)
        out = []
        for s in (1, 2):
            if s == 2:
                m.update_with_move(fx.best_move(i))
                b.do_move(fx.best_move(i))
                if not fx.has_search2(i):
                    assert b.game_end()[0]
                    break
            st = m.replay(b, *fx.trace(s, i))
            assert st["n_tie_order"] == 0, (tag, i, s, st)
            assert st["max_regret_rel"] <= NEAR_TIE, (tag, i, s, st)
            acts, vis, _, _ = m.root_children()
            ref = fx.visits(s, i)
            assert np.array_equal(np.nonzero(ref >= 0)[0], acts)
            assert np.array_equal(ref[acts], vis), (tag, i, s)
            # the gap the generator measured inside the reference agrees with the oracle's view of the same path
            assert abs(st["min_gap_rel"] - fx.gap(s, i)) <= 2e-6 + 0.05 * fx.gap(s, i), (tag, i, s, st, fx.gap(s, i))
            out.append(st)
        return out

    n = fx.n_pos if os.environ.get("GZ_FULL_GOLDEN") else min(fx.n_pos, MCTS2_REPLAY[tag])
    stats = [st for out in _threads(run, range(n)) for st in out]
    decisions = sum(st["n_decisions"] for st in stats)
    print("replayed %d reference decisions at %s, %d within-near-tie deviations"
          % (decisions, tag, sum(st["n_deviations"] for st in stats)))
    assert decisions > 1000


@pytest.mark.parametrize("tag", ["6_6_4", "8_8_5", "9_9_5", "15_15_5"])
def test_oracle_search_vs_golden(tag):
    """Free-running oracle MCTS vs the reference's visit counts (north_star search tier) on EVERY fixture position
    (64 / 64 / 32 / 16), both searches of a tree-reuse pair: identical visit vectors wherever no near-tie (App. B
    rule, measured inside the reference run) lies on the reference's path, total variation <= 0.02 elsewhere."""
    from helpers import Mcts2, golden_net_arrays, tv_distance
    fx = Mcts2(tag)
    net = orc.Net(fx.w, fx.w, golden_net_arrays(tag, fx.z))

    def run(i):
        b = _play(fx.w, fx.w, fx.n, fx.moves(i))
        m = orc.MCTS(5, fx.n_playout, orc.EVAL_NET, net=net)
        same_tree, out = True, []
        for s in (1, 2):
            if s == 2:
                m.update_with_move(fx.best_move(i))
                b.do_move(fx.best_move(i))
                if not fx.has_search2(i):
                    break
            acts, vis, _, _ = m.search(b)
            ref = fx.visits(s, i)
            assert np.array_equal(np.nonzero(ref >= 0)[0], acts)
            assert vis.sum() == ref[acts].sum()
            assert tv_distance(vis, ref[acts]) <= 0.02, (tag, i, s)
            exact = np.array_equal(vis, ref[acts])
            required = same_tree and fx.gap(s, i) > NEAR_TIE
            if required:
                assert exact, (tag, i, s, fx.gap(s, i))
            same_tree = same_tree and exact  # after a near-tie flip the second search starts from another tree
            out.append((exact, required))
        return out

    res = [r for out in _threads(run, range(fx.n_pos)) for r in out]
    print("%s: %d searches, %d identical to the reference, %d of them required (no near-tie)"
          % (tag, len(res), sum(r[0] for r in res), sum(r[1] for r in res)))


def test_oracle_pure_get_move_vs_reference_histogram():
    """mcts_pure.MCTS.get_move (mcts_pure.py:214-231) above the rollout: the move the reference chose in 200
    runs per position (unseeded numpy RNG) against 800 runs of the oracle's ORC_EVAL_PURE search on its Philox
    stream - two-sample chi-square on the chosen-move histogram (rare moves pooled) and the mean root visit
    share of every child within 5 standard errors."""
    z = np.load(os.path.join(GOLDEN, "pure.npz"))
    for pi in range(3):
        w, n, P = [int(x) for x in z["p%d_board" % pi]]
        moves = z["p%d_moves" % pi]
        S = w * w
        ref_chosen = np.bincount(z["p%d_chosen" % pi], minlength=S).astype(np.float64)
        R = int(ref_chosen.sum())
        N = 800
        got_chosen, vsum, vsq = np.zeros(S), np.zeros(S), np.zeros(S)
        b = _play(w, w, n, moves)
        for r in range(N):
            m = orc.MCTS(5, P, orc.EVAL_PURE, salt=1234)
            m.set_stream(r, pi)
            acts, vis, _, _ = m.search(b)
            got_chosen[int(acts[np.argmax(vis)])] += 1
            v = np.zeros(S); v[acts] = vis
            vsum += v; vsq += v * v
        chi_square_two_sample(ref_chosen, got_chosen, "pure get_move p%d" % pi)
        ref_mean, ref_var = z["p%d_visits_sum" % pi] / R, z["p%d_visits_sqsum" % pi] / R - (z["p%d_visits_sum" % pi] / R) ** 2
        got_mean, got_var = vsum / N, vsq / N - (vsum / N) ** 2
        se = np.sqrt(np.maximum(ref_var, 0) / R + np.maximum(got_var, 0) / N) + 1e-9
        assert (np.abs(ref_mean - got_mean) <= 5 * se + 0.5).all(), (pi, ref_mean, got_mean)


@pytest.mark.parametrize("key,games", [("", 24), ("b_", 64)])
def test_oracle_arena_vs_reference_fixture(key, games):
    """The arena (train.py:216-276 policy_evaluate -> game.py:211-290 start_play -> non-self-play get_action,
    mcts_alphaZero.py:244-255, vs mcts_pure.MCTSPlayer.get_action :261-270) on the oracle against the
    reference's win / lose / tie counts over 200 games per start order (tests/golden/arena.npz; A-13: off
    15x15 the net player's opening override is skipped).  Net-player win frequency and tie frequency inside
    the 4-sigma band of a two-sample test of proportions."""
    from helpers import binomial_band_ok, oracle_arena
    z = np.load(os.path.join(GOLDEN, "arena.npz"))
    if key + "config" not in z:
        pytest.skip("arena fixture %r not generated" % key)
    assert int(z["a13_unshimmed_raises"]) == 1
    w, n, az, pure = [int(x) for x in z[key + "config"]]
    net = orc.Net(w, w, load_golden_weights("weights_%d_%d_%d" % (w, w, n)))
    for sp in (0, 1):
        ref = z["%ssp%d_winners" % (key, sp)]
        got = oracle_arena(orc, net, w, n, az, pure, sp, games, seed=11 + sp)
        for who in (1, -1):
            assert binomial_band_ok((ref == who).sum(), len(ref), (got == who).sum(), len(got)), \
                (key, sp, who, (ref == who).sum(), len(ref), (got == who).sum(), len(got))
        print("arena %r start_player %d: reference W/L/T %d/%d/%d of %d, oracle %d/%d/%d of %d" % (
            key, sp, (ref == 1).sum(), (ref == 2).sum(), (ref == -1).sum(), len(ref),
            (got == 1).sum(), (got == 2).sum(), (got == -1).sum(), len(got)))


def test_oracle_philox_known_answers():
    """Philox4x32-10 known-answer vectors (Random123 kat_vectors)."""
    assert tuple(orc.philox(0, 0, 0, 0, 0, 0)) == (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)
    assert tuple(orc.philox(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff)) == \
        (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)
    assert tuple(orc.philox(0xa4093822, 0x299f31d0, 0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344)) == \
        (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)


def test_oracle_rollout_distribution_vs_golden():
    """Philox rollouts from fixed positions vs the reference's _evaluate_rollout outcome counts
    (unseeded numpy RNG there): chi-square on (loss, tie, win) frequencies."""
    z = np.load(os.path.join(GOLDEN, "rollout.npz"))
    for tag, (w, n) in (("6x6x4_a", (6, 4)), ("6x6x4_b", (6, 4)), ("8x8x5_a", (8, 5)), ("8x8x5_b", (8, 5))):
        ref = z[tag + "_counts"].astype(np.float64)
        got = np.zeros(3)
        N = 3000
        for r in range(N):
            b = _play(w, w, n, z[tag + "_moves"])
            got[orc.rollout_once(b, 7, 0, 1, r) + 1] += 1
        p = (ref + got) / (ref.sum() + N)
        chi = sum(((c - tot * pi) ** 2 / (tot * pi)) for cnt, tot in ((ref, ref.sum()), (got, N))
                  for c, pi in zip(cnt, p) if pi > 0)
        assert chi < 20.0, (tag, ref, got, chi)  # 2 dof; 20 ~ p = 5e-5


@pytest.mark.reference
@pytest.mark.skipif(not rh.available(), reason="/root/reference not present (GPU box)")
def test_oracle_vs_live_reference_rules_and_net():
    """Seeded random positions through the unmodified reference Board / PyTorch Net."""
    G = rh.game()
    rng = np.random.RandomState(3)
    for (w, n) in ((6, 4), (8, 5), (9, 5), (15, 5)):
        for _ in range(40):
            rb = G.Board(width=w, height=w, n_in_row=n)
            rb.init_board(0)
            ob = orc.Board(w, w, n)
            ob.init_board(0)
            for _k in range(int(rng.randint(0, w * w // 2))):
                if rb.game_end()[0]:
                    break
                m = rb.availables[rng.randint(len(rb.availables))]
                rb.do_move(m); ob.do_move(m)
                assert rb.game_end() == ob.game_end()
            assert rb.availables == ob.availables and rb.states == ob.states
            assert np.array_equal(np.ascontiguousarray(rb.current_state()), ob.current_state())
            rb.get_move_target(); ob.get_move_target()
            assert list(rb.move_target) == list(ob.move_target)
    for tag, w in (("6_6_4", 6), ("8_8_5", 8)):
        arrs = load_golden_weights("weights_" + tag)
        pv = rh.make_ref_net(w, w, arrs)
        x = (rng.rand(16, 4, w, w) < 0.3).astype(np.float32)
        rp, rv = pv.policy_value(x)
        lp, v = orc.Net(w, w, arrs).forward(x)
        assert np.abs(np.exp(lp) - rp).max() < 1e-4 and np.abs(v - rv.reshape(-1)).max() < 1e-4
