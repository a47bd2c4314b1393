"""CPU tests (no GPU): the oracle against the committed golden vectors generated from the live
reference (tests/golden/make_golden.py), and against the reference itself when /root/reference
is present (build container only)."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, load_golden_weights

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


@pytest.mark.parametrize("tag,w,n", [("6_6_4", 6, 4), ("8_8_5", 8, 5)])
def test_oracle_search_vs_golden(tag, w, n):
    """Oracle MCTS (400 playouts, c_puct 5, shipped weights) vs the reference's visit counts:
    total variation <= 0.02 (north_star), both searches of a tree-reuse pair."""
    z = np.load(os.path.join(GOLDEN, "mcts.npz"))
    moves, offs = z[tag + "_moves"], z[tag + "_offsets"]
    net = orc.Net(w, w, load_golden_weights("weights_" + tag))
    exact = 0
    for g in range(len(offs) - 1):
        b = _play(w, w, n, moves[offs[g]:offs[g + 1]])
        m = orc.MCTS(5, 400, orc.EVAL_NET, net=net)
        acts, vis, _, _ = m.search(b)
        ref = z[tag + "_visits1"][g]
        assert np.array_equal(np.nonzero(ref >= 0)[0], np.asarray(acts))
        a, r = np.asarray(vis, np.float64), ref[acts].astype(np.float64)
        assert a.sum() == r.sum() == 399
        assert 0.5 * np.abs(a / a.sum() - r / r.sum()).sum() <= 0.02
        exact += int(np.array_equal(a, r))
        mv = int(z[tag + "_move"][g])
        m.update_with_move(mv)
        b.do_move(mv)
        ref2 = z[tag + "_visits2"][g]
        if b.game_end()[0] or (ref2 < 0).all():
            continue
        acts2, vis2, _, _ = m.search(b)
        a2, r2 = np.asarray(vis2, np.float64), ref2[acts2].astype(np.float64)
        assert a2.sum() == r2.sum()
        assert 0.5 * np.abs(a2 / a2.sum() - r2 / r2.sum()).sum() <= 0.02
    assert exact >= 1  # off near-ties the counts are identical


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
