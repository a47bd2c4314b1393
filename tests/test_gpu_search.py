"""GPU parity tests for the tree kernels, the search loop, rollouts and self-play.

Tiers (BASELINE.json north_star): tree statistics are BIT-EXACT against the oracle when both
sides see the same evaluator outputs (synthetic integer-hash evaluator, injected through
gz_search_select / gz_search_expand_backup); with the real net the visit distributions must
be within total-variation 0.02 of the reference's (golden mcts.npz), exact off near-ties."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, hash_eval_batch, load_golden_weights

pytestmark = pytest.mark.gpu


def _engine(w, h, n, **kw):
    from alphazero_gomoku_b200 import Engine
    return Engine(width=w, height=h, n_in_row=n, **kw)


def _random_positions(rng, w, n, G, max_plies, min_plies=0):
    from oracle import orc
    boards, lists = [], []
    while len(boards) < G:
        b = orc.Board(w, w, n)
        b.init_board(0)
        mv = []
        for _ in range(int(rng.randint(min_plies, max_plies + 1))):
            av = b.availables
            m = av[rng.randint(len(av))]
            b.do_move(m)
            mv.append(m)
            if b.game_end()[0]:
                break
        if b.game_end()[0]:
            continue
        boards.append(b)
        lists.append(mv)
    return boards, lists


def _device_hash_search(e, n_playout, salt):
    for _ in range(n_playout):
        st, planes, cells, player = e.search_select()
        with np.errstate(all="ignore"):
            priors, values = hash_eval_batch(cells, player, salt)
        priors = np.nan_to_num(priors)
        e.search_expand_backup(priors, values)


@pytest.mark.parametrize("w,n,G,n_playout,c_puct", [(6, 4, 48, 200, 5.0), (8, 5, 96, 400, 5.0),
                                                    (9, 5, 24, 150, 3.0), (15, 5, 8, 120, 5.0)])
def test_tree_bit_exact_with_hash_evaluator(w, n, G, n_playout, c_puct):
    """select / expand / backup / re-root vs the oracle's pointer tree: identical visit
    counts AND identical float64 Q for every root child, over two searches with tree reuse."""
    from oracle import orc
    rng = np.random.RandomState(w)
    salt = 1234 + w
    boards, lists = _random_positions(rng, w, n, G, max_plies=w * w // 2)
    e = _engine(w, w, n, n_games=G, n_playout=n_playout, c_puct=c_puct)
    for g, mv in enumerate(lists):
        e.boards_set(g, mv)
    _device_hash_search(e, n_playout, salt)
    vis, q, rn = e.root_visits(want_q=True)
    trees = []
    for g, b in enumerate(boards):
        m = orc.MCTS(c_puct, n_playout, orc.EVAL_HASH, salt=salt)
        acts, ovis, oq, _ = m.search(b)
        trees.append(m)
        assert np.nonzero(vis[g] >= 0)[0].tolist() == acts.tolist()
        assert np.array_equal(vis[g][acts], ovis), (g, vis[g][acts], ovis)
        assert np.array_equal(q[g][acts], oq)  # bit-exact float64
        assert int(rn[g]) == m.root_visits()
    # tree reuse: play the most visited move everywhere, search again
    best = np.array([int(np.argmax(vis[g])) for g in range(G)], np.int32)
    e.update_with_move(best)
    e.boards_do_move(best)
    ended = e.boards_status()["ended"]
    _device_hash_search(e, n_playout // 2, salt)
    vis2, q2, rn2 = e.root_visits(want_q=True)
    for g, b in enumerate(boards):
        trees[g].update_with_move(int(best[g]))
        b.do_move(int(best[g]))
        assert bool(ended[g]) == b.game_end()[0]
        acts, ovis, oq, _ = trees[g].search(b, n_playout // 2)
        if b.game_end()[0]:
            # a finished root never gets children (terminal leaves are not expanded)
            assert (vis2[g] < 0).all() and len(acts) == 0
        else:
            assert np.array_equal(vis2[g][acts], ovis)
            assert np.array_equal(q2[g][acts], oq)
        assert int(rn2[g]) == trees[g].root_visits()
    # reset (update_with_move(-1)) gives fresh roots
    e.update_with_move(np.full(G, -1, np.int32))
    vis3, rn3 = e.root_visits()
    assert (vis3 < 0).all() and (rn3 == 0).all()
    assert e.selfplay_stats()["pool_overflows"] == 0
    e.close()


@pytest.mark.parametrize("tag,w,n", [("6_6_4", 6, 4), ("8_8_5", 8, 5)])
def test_search_with_net_vs_reference_golden(tag, w, n):
    """Full device search (net on device) vs the reference's visit counts (400 playouts,
    shipped weights): TV <= 0.02, and the second search reuses the subtree."""
    z = np.load(os.path.join(GOLDEN, "mcts.npz"))
    moves, offs = z[tag + "_moves"], z[tag + "_offsets"]
    G = len(offs) - 1
    e = _engine(w, w, n, n_games=G, n_playout=400)
    e.set_weights(load_golden_weights("weights_" + tag))
    for g in range(G):
        e.boards_set(g, moves[offs[g]:offs[g + 1]])
    e.search(400)
    vis, rn = e.root_visits()
    exact = 0
    for g in range(G):
        ref = z[tag + "_visits1"][g]
        assert np.array_equal(vis[g] >= 0, ref >= 0)
        a, b = np.maximum(vis[g], 0), np.maximum(ref, 0)
        assert a.sum() == b.sum() == 399
        tv = 0.5 * np.abs(a / a.sum() - b / b.sum()).sum()
        exact += int(np.array_equal(a, b))
        assert tv <= 0.02, (g, tv)
    mv = z[tag + "_move"].astype(np.int32)
    e.update_with_move(mv)
    e.boards_do_move(mv)
    e.search(400)
    vis2, _ = e.root_visits()
    for g in range(G):
        ref = z[tag + "_visits2"][g]
        if (ref < 0).all():
            continue
        a, b = np.maximum(vis2[g], 0), np.maximum(ref, 0)
        assert a.sum() == b.sum()
        tv = 0.5 * np.abs(a / a.sum() - b / b.sum()).sum()
        exact += int(np.array_equal(a, b))
        assert tv <= 0.02, (g, tv)
    print("search vs reference", tag, "exact visit vectors:", exact, "of", 2 * G)
    e.close()


def test_search_with_net_vs_oracle_many_games():
    """Same as above against the C oracle (its own fp32 net) on more positions, 8x8."""
    from oracle import orc
    w, n, G, P = 8, 5, 32, 200
    arrs = load_golden_weights("weights_8_8_5")
    rng = np.random.RandomState(11)
    boards, lists = _random_positions(rng, w, n, G, max_plies=20)
    e = _engine(w, w, n, n_games=G, n_playout=P)
    e.set_weights(arrs)
    for g, mv in enumerate(lists):
        e.boards_set(g, mv)
    e.search(P)
    vis, _ = e.root_visits()
    onet = orc.Net(w, w, arrs)
    exact, worst = 0, 0.0
    for g, b in enumerate(boards):
        m = orc.MCTS(5, P, orc.EVAL_NET, net=onet)
        acts, ovis, _, _ = m.search(b)
        a = np.maximum(vis[g], 0)[acts]
        tv = 0.5 * np.abs(a / a.sum() - ovis / ovis.sum()).sum()
        worst = max(worst, tv)
        exact += int(np.array_equal(a, ovis))
    print("search vs oracle: exact", exact, "of", G, "worst TV", worst)
    assert worst <= 0.02
    assert exact >= G // 2
    e.close()


def test_search_with_wide_trunk_vs_oracle():
    """BASELINE config C4's 'wider conv trunk' (64/128/256) goes through the layer-wise tcgen05 kernels
    (gz_net_tcw.cu) inside the search as well: visit distributions against the C oracle's MCTS with its
    own fp32 net of the same widths, and self-play on top of it produces legal finished games."""
    from oracle import orc
    from alphazero_gomoku_b200.engine import weight_shapes
    w, n, G, P, trunk = 9, 5, 24, 120, (64, 128, 256)
    wr = np.random.RandomState(21)
    arrs = [(wr.randn(*sh) * (np.sqrt(2.0 / int(np.prod(sh[1:]))) if len(sh) > 1 else 0.1)).astype(np.float32)
            for sh in weight_shapes(w, w, trunk)]
    rng = np.random.RandomState(12)
    boards, lists = _random_positions(rng, w, n, G, max_plies=16)
    e = _engine(w, w, n, n_games=G, n_playout=P, trunk=trunk, net_impl=2)
    e.set_weights(arrs)
    for g, mv in enumerate(lists):
        e.boards_set(g, mv)
    e.search(P)
    vis, _ = e.root_visits()
    onet = orc.Net(w, w, arrs)
    exact, worst = 0, 0.0
    for g, b in enumerate(boards):
        m = orc.MCTS(5, P, orc.EVAL_NET, net=onet)
        acts, ovis, _, _ = m.search(b)
        a = np.maximum(vis[g], 0)[acts]
        tv = 0.5 * np.abs(a / a.sum() - ovis / ovis.sum()).sum()
        worst = max(worst, tv)
        exact += int(np.array_equal(a, ovis))
    print("wide-trunk search vs oracle: exact", exact, "of", G, "worst TV", worst)
    assert worst <= 0.02
    assert exact >= G // 2
    e.close()
    e = _engine(w, w, n, n_games=16, n_playout=40, trunk=trunk, net_impl=2, seed=4)
    e.set_weights(arrs)
    e.selfplay_begin(temp=1.0, recycle=False)
    for _ in range(w * w):
        e.selfplay_step(1)
        if e.selfplay_stats()["n_games"] >= 16:
            break
    games = e.selfplay_drain_games()
    assert len(games) == 16
    for gid, win, mv in games:
        b = orc.Board(w, w, n)
        b.init_board(0)
        for m_ in mv:
            assert not b.game_end()[0]
            b.do_move(m_)
        assert b.game_end() == (True, win)
    e.close()


@pytest.mark.parametrize("w,n,G,P", [(8, 5, 200, 150), (6, 4, 96, 120), (5, 4, 64, 80), (9, 5, 40, 60), (11, 5, 24, 40),
                                     (15, 5, 12, 40), (13, 5, 8, 30), (7, 4, 32, 60)])
def test_fused_wave_kernel_matches_split_kernels(w, n, G, P):
    """gz_search (fused expand+backup+select kernel across the wave boundary, its low-latency
    variant on boards of up to 64 cells) against the same search driven wave by wave through
    gz_search_select / gz_net_forward / gz_search_expand_backup: identical trees, bit for bit."""
    rng = np.random.RandomState(7)
    if w == 8:
        arrs = load_golden_weights("weights_8_8_5")
    else:
        from alphazero_gomoku_b200.engine import weight_shapes
        wr = np.random.RandomState(5)
        arrs = [(wr.randn(*sh) * (np.sqrt(2.0 / int(np.prod(sh[1:]))) if len(sh) > 1 else 0.1)).astype(np.float32)
                for sh in weight_shapes(w, w)]
    _, lists = _random_positions(rng, w, n, G, max_plies=min(w * w - 2, 30))
    res = []
    for fused in (True, False):
        e = _engine(w, w, n, n_games=G, n_playout=P, seed=3)
        e.set_weights(arrs)
        for g, mv in enumerate(lists):
            e.boards_set(g, mv)
        for rounds in (P, P // 3):  # the second search continues on the grown trees
            if fused:
                e.search(rounds)
            else:
                for _ in range(rounds):
                    st, planes, _, _ = e.search_select(want_cells=False)
                    probs, values = e.net_forward(planes)
                    e.search_expand_backup(probs, values)
        vis, q, rn = e.root_visits(want_q=True)
        res.append((vis.copy(), q.copy(), rn.copy()))
        e.close()
    for a, b in zip(res[0], res[1]):
        assert np.array_equal(a, b)
    assert (res[0][2] == P + P // 3).all()


def test_rollouts_bit_exact_vs_oracle():
    from oracle import orc
    for (w, n) in [(6, 4), (8, 5), (15, 5), (9, 5), (7, 3), (5, 4), (8, 4), (11, 5)]:
        rng = np.random.RandomState(w)
        G, R, seed, sid = 24, 16, 0x1234567890ABCDEF, 7
        # (an empty board narrower than 6 has no candidate box, game.py:41: see test_gpu_golden2)
        boards, lists = _random_positions(rng, w, n, G, max_plies=w * w // 3, min_plies=1 if w < 6 else 0)
        e = _engine(w, w, n, n_games=G, seed=seed, game_id_base=100)
        for g, mv in enumerate(lists):
            e.boards_set(g, mv)
        res, plies = e.rollouts(R, search_id=sid)
        for g, b in enumerate(boards):
            for r in range(R):
                assert int(res[g, r]) == orc.rollout_once(b, seed, 100 + g, sid, r)
        assert plies > 0
        e.close()


def test_rollout_distribution_vs_reference_golden():
    """Outcome frequencies of the Philox rollouts vs 3000 reference rollouts per position."""
    z = np.load(os.path.join(GOLDEN, "rollout.npz"))
    for tag, w, n in [("6x6x4_a", 6, 4), ("6x6x4_b", 6, 4), ("8x8x5_a", 8, 5), ("8x8x5_b", 8, 5)]:
        e = _engine(w, w, n, n_games=1, seed=99)
        e.boards_set(0, z[tag + "_moves"])
        R = 20000
        res, _ = e.rollouts(R, search_id=1)
        got = np.array([(res == -1).sum(), (res == 0).sum(), (res == 1).sum()], np.float64) / R
        ref = z[tag + "_counts"].astype(np.float64)
        nref = ref.sum()
        ref /= nref
        sd = np.sqrt(ref * (1 - ref) / nref + got * (1 - got) / R) + 1e-9
        print("rollout dist", tag, got, ref)
        assert (np.abs(got - ref) < 5 * sd + 0.005).all(), (tag, got, ref)
        e.close()


@pytest.mark.parametrize("w,n,P", [(6, 4, 300), (8, 5, 200), (9, 5, 150), (15, 5, 60), (7, 3, 120), (5, 4, 100), (11, 5, 80)])
def test_pure_mcts_bit_exact_vs_oracle(w, n, P):
    from oracle import orc
    rng = np.random.RandomState(3)
    G, seed, sid = 16, 42, 5
    boards, lists = _random_positions(rng, w, n, G, max_plies=10, min_plies=1 if w < 6 else 0)
    e = _engine(w, w, n, n_games=G, n_playout=P, seed=seed, game_id_base=0)
    for g, mv in enumerate(lists):
        e.boards_set(g, mv)
    e.pure_search(P, search_id=sid)
    vis, rn = e.root_visits()
    best = e.pure_best_move()
    for g, b in enumerate(boards):
        m = orc.MCTS(5, P, orc.EVAL_PURE, salt=seed)
        m.set_stream(g, sid)
        acts, ovis, _, _ = m.search(b)
        assert np.nonzero(vis[g] >= 0)[0].tolist() == acts.tolist()
        assert np.array_equal(vis[g][acts], ovis)
        assert int(best[g]) == int(acts[np.argmax(ovis)])
    e.close()


def test_root_select_move_distribution():
    """pi = softmax(log(v)/temp) exact vs numpy float64; sampled moves follow
    0.75*pi + 0.25*Dir(0.1) (self-play) resp. pi (match play)."""
    w, n, G = 6, 4, 4096
    e = _engine(w, w, n, n_games=G, n_playout=60, seed=7)
    e.set_weights(load_golden_weights("weights_6_6_4"))
    e.boards_set(0, [14, 15])
    for g in range(1, G):
        e.boards_set(g, [14, 15])
    e.search(60)
    vis, _ = e.root_visits()
    assert (vis == vis[0]).all()  # identical games, deterministic search
    v = vis[0]
    acts = np.nonzero(v >= 0)[0]
    for temp in (1.0, 0.5, 1e-3):
        moves, probs = e.root_select_move(temp=temp, selfplay=False)
        x = 1.0 / temp * np.log(v[acts] + 1e-10)
        p = np.exp(x - x.max())
        p /= p.sum()
        assert np.abs(probs[0][acts] - p).max() < 1e-6
        assert np.abs(probs[0].sum() - 1) < 1e-5
        cnt = np.bincount(moves, minlength=w * w)[acts]
        sd = np.sqrt(G * p * (1 - p)) + 1.0
        assert (np.abs(cnt - G * p) < 5 * sd).all(), (temp, cnt, G * p)
    moves, probs = e.root_select_move(temp=1.0, selfplay=True)
    x = np.log(v[acts] + 1e-10)
    p = np.exp(x - x.max()); p /= p.sum()
    mix = 0.75 * p + 0.25 / len(acts)
    cnt = np.bincount(moves, minlength=w * w)[acts]
    sd = np.sqrt(G * mix * (1 - mix)) + 1.0
    assert (np.abs(cnt - G * mix) < 5 * sd).all(), (cnt, G * mix)
    assert np.abs(probs[0][acts] - p).max() < 1e-6  # returned pi is noise-free (A-10)
    e.close()


def test_selfplay_records_match_oracle_replay():
    """Run device self-play to completion (one game per slot) and re-derive every record
    from the move lists with the oracle: states, z, pi normalisation, legality, winner."""
    from oracle import orc
    w, n, G, P = 6, 4, 48, 48
    e = _engine(w, w, n, n_games=G, n_playout=P, seed=3)
    e.set_weights(load_golden_weights("weights_6_6_4"))
    e.selfplay_begin(temp=1.0, recycle=False)
    for _ in range(w * w):
        e.selfplay_step(1)
        if e.selfplay_pending()[2] == 0:
            break
    ns, ng, nact = e.selfplay_pending()
    assert nact == 0 and ng == G
    games = e.selfplay_drain_games()
    data = e.selfplay_drain()
    assert len(games) == G and len(data["zs"]) == ns == sum(len(m) for _, _, m in games)
    stats = e.selfplay_stats()
    assert stats["pool_overflows"] == 0 and stats["n_games"] == G and stats["n_samples"] == ns
    by_gid = {gid: (win, mv) for gid, win, mv in games}
    assert sorted(by_gid) == list(range(G))
    order = np.lexsort((data["plies"], data["game_ids"]))
    k = 0
    for gid in sorted(by_gid):
        win, mv = by_gid[gid]
        b = orc.Board(w, w, n)
        b.init_board(0)
        for t, m in enumerate(mv):
            i = order[k]; k += 1
            assert data["game_ids"][i] == gid and data["plies"][i] == t
            assert np.array_equal(data["states"][i], b.current_state().astype(np.float32))
            pi = data["pis"][i]
            assert abs(pi.sum() - 1) < 1e-4
            assert (pi[[c for c in range(w * w) if c not in b.availables]] == 0).all()
            z = 0.0 if win == -1 else (1.0 if b.current_player == win else -1.0)
            assert data["zs"][i] == z
            assert not b.game_end()[0]
            b.do_move(m)  # raises if the engine played an illegal move
        assert b.game_end() == (True, win)
    e.close()


def test_selfplay_pi_matches_oracle_search_with_tree_reuse():
    """Follow the device's sampled moves with the oracle MCTS (tree reuse) and compare the
    recorded pi (temp = 1 -> visit fractions) ply by ply: TV <= 0.02."""
    from oracle import orc
    w, n, G, P = 6, 4, 6, 120
    arrs = load_golden_weights("weights_6_6_4")
    e = _engine(w, w, n, n_games=G, n_playout=P, seed=17)
    e.set_weights(arrs)
    e.selfplay_begin(temp=1.0, recycle=False)
    while e.selfplay_pending()[2] > 0:
        e.selfplay_step(1)
    games = e.selfplay_drain_games()
    data = e.selfplay_drain()
    onet = orc.Net(w, w, arrs)
    worst, exact, total = 0.0, 0, 0
    for gid, win, mv in games:
        b = orc.Board(w, w, n)
        b.init_board(0)
        m = orc.MCTS(5, P, orc.EVAL_NET, net=onet)
        for t, move in enumerate(mv):
            acts, ovis, _, _ = m.search(b)
            i = np.nonzero((data["game_ids"] == gid) & (data["plies"] == t))[0][0]
            pi = data["pis"][i][acts].astype(np.float64)
            ref = ovis / ovis.sum()
            tv = 0.5 * np.abs(pi - ref).sum()
            worst = max(worst, tv)
            exact += int(tv < 1e-6)
            total += 1
            m.update_with_move(move)
            b.do_move(move)
    print("selfplay pi vs oracle: exact", exact, "of", total, "worst TV", worst)
    assert worst <= 0.02
    e.close()


def test_full_size_search_equals_small_engine():
    """BASELINE configs[2] per-GPU size (16384 games, 8x8, 400 playouts): 64 distinct positions,
    256 copies of each spread over the slots.  Size-independent properties: every copy of a
    position ends with the identical tree statistics (games never interact), those statistics equal
    the same search in a 64-game engine (whose kernels the oracle tests pin), and the per-game
    visit invariants hold."""
    w, n, G, P, K = 8, 5, 16384, 400, 64
    arrs = load_golden_weights("weights_8_8_5")
    rng = np.random.RandomState(21)
    _, lists = _random_positions(rng, w, n, K, max_plies=24)
    small = _engine(w, w, n, n_games=K, n_playout=P)
    small.set_weights(arrs)
    for g, mv in enumerate(lists):
        small.boards_set(g, mv)
    small.search(P)
    svis, sq, srn = small.root_visits(want_q=True)
    small.close()
    big = _engine(w, w, n, n_games=G, n_playout=P)
    big.set_weights(arrs)
    # copy c of position k sits in slot c * K + k: neighbouring slots hold different positions
    for g in range(G):
        big.boards_set(g, lists[g % K])
    big.search(P)
    vis, q, rn = big.root_visits(want_q=True)
    ctr = big.counters()
    overflows = big.selfplay_stats()["pool_overflows"]
    big.close()
    assert (rn == P).all()
    assert (np.maximum(vis, 0).sum(axis=1) == P - 1).all()
    v3, q3 = vis.reshape(G // K, K, -1), q.reshape(G // K, K, -1)
    assert (v3 == v3[0]).all() and (q3 == q3[0]).all()         # replica invariance, bit for bit
    assert np.array_equal(v3[0], svis) and np.array_equal(q3[0], sq)  # == the small engine
    for k, mv in enumerate(lists):                             # children are exactly the empty cells
        legal = np.ones(w * w, bool); legal[mv] = False
        assert ((vis[k] >= 0) == legal).all()
    assert ctr["playouts"] == G * P and overflows == 0


def test_full_size_selfplay_records_are_legal_games():
    """16384 concurrent self-play games (8x8, 40 playouts to keep it short): every finished game
    replays legally through the oracle rules to the recorded winner, z follows game.py:396-399,
    sample count = sum of game lengths, game ids are unique."""
    from oracle import orc
    w, n, G, P = 8, 5, 16384, 40
    e = _engine(w, w, n, n_games=G, n_playout=P, seed=9)
    e.set_weights(load_golden_weights("weights_8_8_5"))
    e.selfplay_begin(temp=1.0, recycle=True)
    e.selfplay_step(14)
    games = e.selfplay_drain_games()
    data = e.selfplay_drain()
    stats = e.selfplay_stats()
    e.close()
    assert stats["pool_overflows"] == 0 and stats["n_plies"] == 14 * G
    assert len(games) == stats["n_games"] > 0
    assert len(data["zs"]) == stats["n_samples"] == sum(len(m) for _, _, m in games)
    gids = [gid for gid, _, _ in games]
    assert len(set(gids)) == len(gids)
    zs = {}
    for gid, ply, z in zip(data["game_ids"], data["plies"], data["zs"]):
        zs[(int(gid), int(ply))] = float(z)
    assert abs(data["pis"].sum(axis=1) - 1).max() < 1e-4
    for gid, win, mv in games[:600]:
        b = orc.Board(w, w, n)
        b.init_board(0)
        for t, m in enumerate(mv):
            assert not b.game_end()[0]
            want = 0.0 if win == -1 else (1.0 if b.current_player == win else -1.0)
            assert zs[(gid, t)] == want
            b.do_move(m)
        assert b.game_end() == (True, win)


def test_degenerate_batches_do_not_disturb_other_games():
    """Edge cases of the wave machinery: an empty net batch, slots whose root is already terminal
    (won / full board: every playout ends at the root and no leaf reaches the net) next to a live
    game, and a wave in which no game needs the net at all."""
    w, n, P = 6, 4, 50
    arrs = load_golden_weights("weights_6_6_4")
    e = _engine(w, w, n, n_games=4, n_playout=P)
    e.set_weights(arrs)
    probs, vals = e.net_forward(np.zeros((0, 4, w, w), np.float32))
    assert probs.shape == (0, w * w) and vals.shape == (0,)
    won = [0, 6, 1, 7, 2, 8, 3]                      # player 1 completes 0-1-2-3
    live = [14, 15, 20]
    e.boards_set(0, won)
    e.boards_set(1, live)
    e.boards_set(2, won)
    e.boards_set(3, live)
    st = e.boards_status()
    assert list(st["ended"]) == [1, 0, 1, 0] and st["winner"][0] == 1
    e.search(P)
    vis, rn = e.root_visits()
    assert (vis[0] == -1).all() and (vis[2] == -1).all()          # terminal roots never expand
    assert rn[0] == P and rn[2] == P
    assert np.array_equal(vis[1], vis[3]) and np.maximum(vis[1], 0).sum() == P - 1
    alone = _engine(w, w, n, n_games=1, n_playout=P)
    alone.set_weights(arrs)
    alone.boards_set(0, live)
    alone.search(P)
    v1, _ = alone.root_visits()
    assert np.array_equal(v1[0], vis[1])
    alone.close()
    # now only terminal roots are left active: waves with an empty leaf batch
    e.boards_set(1, won)
    e.boards_set(3, won)
    e.update_with_move(np.full(4, -1, np.int32))
    e.search(7)
    vis, rn = e.root_visits()
    assert (vis == -1).all() and (rn == 7).all()
    e.close()


_PDL_SCRIPT = r"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {tests!r})
import numpy as np
from helpers import load_golden_weights
from alphazero_gomoku_b200 import Engine
from alphazero_gomoku_b200.engine import weight_shapes
out = {{}}
# shipped weights at 6x6 / 8x8; seeded He weights at 9x9 / 15x15 (other tree-kernel instantiations, the
# large-board heads path) and for a wide trunk at 8x8 (layer-wise kernels, launched outside the chain)
for name, w, n, trunk in (("6_6_4", 6, 4, None), ("8_8_5", 8, 5, None), ("9", 9, 5, None), ("15", 15, 5, None),
                          ("8w", 8, 5, (64, 64, 128))):
    if name in ("6_6_4", "8_8_5"):
        wts = load_golden_weights("weights_" + name)
    else:
        rng = np.random.RandomState(w)
        wts = [(rng.randn(*s) * (np.sqrt(2.0 / int(np.prod(s[1:]))) if len(s) > 1 else 0.1)).astype(np.float32)
               for s in weight_shapes(w, w, trunk or (32, 64, 128))]
    for G in (1, 3, 64, 700):
        kw = dict(trunk=trunk) if trunk else dict()
        e = Engine(width=w, height=w, n_in_row=n, n_games=G, n_playout=150, seed=5, **kw)
        e.set_weights(wts)
        rng = np.random.RandomState(G)
        for g in range(G):
            e.boards_set(g, [int(x) for x in rng.permutation(w * w)[:int(rng.randint(0, 5))]])
        for rep in range(2):  # the second search reuses the tree after a move in every game
            e.search(150)
            vis, rn = e.root_visits()
            out["%s_%d_%d" % (name, G, rep)] = vis
            e.update_with_move(vis.argmax(axis=1).astype(np.int32))
        e.close()
np.savez(sys.argv[1], **out)
"""


def test_dependent_launch_chain_matches_serial_launches(tmp_path):
    """The wave chain launched with programmatic dependent launch (kernels start before their
    predecessor has drained) must build the same trees as plain stream-ordered launches
    (GZ_NO_PDL=1): identical visit counts, small grids included - those are the ones where a
    successor can become resident early.  Third arm, GZ_LEAF_FLOAT=1: the tree kernels write fp32
    leaf planes for the trunk instead of the trunk forming them from the leaf bitboards of the slot's
    game (the shipped path) - the same planes either way, so again identical trees (the wide trunk at
    8x8 takes fp32 planes in both arms)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "pdl_case.py"
    script.write_text(_PDL_SCRIPT.format(root=root, tests=os.path.join(root, "tests")))
    res = {}
    for tag, extra in (("pdl", {}), ("serial", {"GZ_NO_PDL": "1"}), ("float_planes", {"GZ_LEAF_FLOAT": "1"})):
        env = {k: v for k, v in os.environ.items() if k not in ("GZ_NO_PDL", "GZ_LEAF_FLOAT")}
        env.update(extra)
        out = tmp_path / (tag + ".npz")
        subprocess.check_call([sys.executable, str(script), str(out)], env=env, timeout=600)
        res[tag] = np.load(out)
    assert sorted(res["pdl"].files) == sorted(res["serial"].files) and len(res["pdl"].files) == 40
    assert sorted(res["pdl"].files) == sorted(res["float_planes"].files)
    for k in res["pdl"].files:
        assert res["pdl"][k].sum() > 0, k
        assert np.array_equal(res["pdl"][k], res["serial"][k]), k
        assert np.array_equal(res["pdl"][k], res["float_planes"][k]), k
