"""GPU tests of the drop-in Python API (Board / Game / MCTS / MCTSPlayer / PolicyValueNet /
TrainPipeline.collect_selfplay_data) against the oracle and the committed reference goldens.
These read like the calls a user of the reference makes."""
import copy
import os

import numpy as np
import pytest

from helpers import GOLDEN, load_golden_weights

pytestmark = pytest.mark.gpu


def _save_state_dict(tmp_path, tag):
    import torch
    from alphazero_gomoku_b200.engine import STATE_DICT_KEYS
    arrs = load_golden_weights("weights_" + tag)
    path = str(tmp_path / ("w_%s.model" % tag))
    torch.save({k: torch.from_numpy(a.copy()) for k, a in zip(STATE_DICT_KEYS, arrs)}, path)
    return path, arrs


def test_board_protocol_matches_oracle():
    from alphazero_gomoku_b200 import Board
    from oracle import orc
    rng = np.random.RandomState(11)
    for (w, n) in ((6, 4), (8, 5), (15, 5)):
        for trial in range(12):
            b = Board(width=w, height=w, n_in_row=n)
            b.init_board(trial % 2)
            ob = orc.Board(w, w, n)
            ob.init_board(trial % 2)
            for _ in range(int(rng.randint(0, w * w))):
                if ob.game_end()[0]:
                    break
                m = ob.availables[rng.randint(len(ob.availables))]
                b.do_move(m); ob.do_move(m)
            b2 = copy.deepcopy(b)  # mcts_alphaZero.py:162 deep-copies boards
            assert b2.availables == ob.availables and b2.states == ob.states
            assert b2.last_move == ob.last_move and b2.current_player == ob.current_player
            assert b2.game_end() == ob.game_end() and b2.has_a_winner() == ob.has_a_winner()
            cs = b2.current_state()
            assert cs.dtype == np.float64 and np.array_equal(cs, ob.current_state())
            b2.get_move_target(); ob.get_move_target()
            assert b2.move_target == list(ob.move_target)
            assert b.location_to_move(b.move_to_location(7)) == 7 and b.location_to_move([w, 0]) == -1
    with pytest.raises(Exception):
        Board(width=4, height=4, n_in_row=5).init_board()


@pytest.mark.parametrize("tag,w,n", [("6_6_4", 6, 4), ("8_8_5", 8, 5)])
def test_policy_value_net_and_mcts_vs_reference_golden(tag, w, n, tmp_path):
    from alphazero_gomoku_b200 import Board, MCTS, PolicyValueNet
    path, arrs = _save_state_dict(tmp_path, tag)
    net = PolicyValueNet(w, w, model_file=path)
    z = np.load(os.path.join(GOLDEN, "net.npz"))
    probs, vals = net.policy_value(z[tag + "_planes"])
    assert vals.shape == (len(probs), 1)
    assert np.abs(probs - z[tag + "_probs"]).max() < 1e-4 and np.abs(vals[:, 0] - z[tag + "_values"]).max() < 1e-4
    for a, b in zip(net.get_policy_param(), arrs):
        assert np.array_equal(a, b)
    # MCTS.get_move_probs vs the reference's visit counts (400 playouts, c_puct 5)
    g = np.load(os.path.join(GOLDEN, "mcts.npz"))
    moves, offs = g[tag + "_moves"], g[tag + "_offsets"]
    for k in range(3):
        b = Board(width=w, height=w, n_in_row=n)
        b.init_board()
        for m in moves[offs[k]:offs[k + 1]]:
            b.do_move(int(m))
        mcts = MCTS(net.policy_value_fn, c_puct=5, n_playout=400)
        acts, pr = mcts.get_move_probs(b, temp=1.0)
        ref = g[tag + "_visits1"][k]
        assert list(acts) == np.nonzero(ref >= 0)[0].tolist() == b.availables
        want = ref[list(acts)] / ref[list(acts)].sum()
        assert pr.dtype == np.float64 and abs(pr.sum() - 1) < 1e-9
        assert 0.5 * np.abs(pr - want).sum() <= 0.02
    # policy_value_fn protocol: ascending legal moves, probabilities of exactly those moves
    ap, v = net.policy_value_fn(b)
    ap = list(ap)
    assert [a for a, _ in ap] == b.availables and -1.0 <= v <= 1.0 and isinstance(v, float)
    # save_model round trip (torch state_dict, loadable by the reference)
    out = str(tmp_path / "saved.model")
    net.save_model(out)
    net2 = PolicyValueNet(w, w, model_file=out)
    assert all(np.array_equal(a, b) for a, b in zip(net2.get_policy_param(), arrs))


def test_external_evaluator_protocol_matches_oracle():
    """Any callable honouring policy_value_fn works: the tree must be bit-identical to the
    oracle's MCTS driven by the same callable."""
    from alphazero_gomoku_b200 import Board, MCTS
    from oracle import orc

    def evaluator(board):
        av = list(board.availables)
        s = sum(board.states.keys()) + 3 * len(board.states) + board.current_player
        w = np.array([((m * 7 + s) % 13) + 1 for m in av], dtype=np.float32)
        return zip(av, w / w.sum()), float(((s * 37) % 21 - 10) / 10.0)

    w, n, P = 6, 4, 120
    b = Board(width=w, height=w, n_in_row=n)
    b.init_board()
    ob = orc.Board(w, w, n)
    ob.init_board(0)
    for m in (14, 15, 20, 21):
        b.do_move(m); ob.do_move(m)
    acts, pr = MCTS(evaluator, c_puct=5, n_playout=P).get_move_probs(b, temp=1.0)

    # oracle side: the C restatement of the reference loop calling the same evaluator
    om = orc.MCTS(5, P, orc.EVAL_CALLBACK, callback=_oracle_callback(evaluator, w, n))
    oacts, ovis, _, _ = om.search(ob)
    assert list(acts) == list(oacts)
    from alphazero_gomoku_b200.mcts_alphaZero import softmax
    want = softmax(np.log(np.asarray(ovis, np.float64) + 1e-10))  # mcts_alphaZero.py:179 at temp 1
    assert np.array_equal(pr, want)  # identical visit counts -> identical probabilities


def _oracle_callback(evaluator, w, n):
    """Adapts a python evaluator to the oracle's C callback (CBoard*, float* priors, float* value)."""
    class _B(object):
        pass

    def cb(cboard, priors, value):
        c = cboard.contents
        bb = _B()
        S = c.width * c.height
        bb.width, bb.height, bb.n_in_row = c.width, c.height, c.n_in_row
        bb.states = {m: c.states[m] for m in range(S) if c.states[m] > 0}
        bb.availables = [c.availables[i] for i in range(c.n_avail)]
        bb.current_player = c.current_player
        ap, v = evaluator(bb)
        for i, (a, p) in enumerate(ap):  # probs[i] belongs to availables[i] (ascending moves)
            assert a == bb.availables[i]
            priors[i] = p
        value[0] = v
    return cb


def test_selfplay_game_and_player_protocol(tmp_path):
    """Game.start_self_play with the drop-in MCTSPlayer: a legal finished game, winners_z per
    game.py:396-399, states equal to replaying the moves, noise-free pi that sums to 1."""
    from alphazero_gomoku_b200 import Board, Game, MCTSPlayer, PolicyValueNet
    from oracle import orc
    path, _ = _save_state_dict(tmp_path, "6_6_4")
    net = PolicyValueNet(6, 6, model_file=path)
    player = MCTSPlayer(net.policy_value_fn, c_puct=5, n_playout=60, is_selfplay=1)
    np.random.seed(0)
    game = Game(Board(width=6, height=6, n_in_row=4))
    winner, data = game.start_self_play(player, temp=1.0)
    data = list(data)
    assert winner in (1, 2, -1) and 7 <= len(data) <= 36
    ob = orc.Board(6, 6, 4)
    ob.init_board(0)
    moves = list(game.board.states.keys())
    for (state, pi, z), mv in zip(data, moves):
        assert np.array_equal(state, ob.current_state())
        assert abs(pi.sum() - 1) < 1e-9 and pi[mv] > 0 and all(pi[m] == 0 for m in ob.states)
        assert z == (0.0 if winner == -1 else (1.0 if ob.current_player == winner else -1.0))
        ob.do_move(mv)
    assert ob.game_end() == (True, winner)
    # match play: AlphaZero player vs the pure-MCTS rollout player (train.py:216-276)
    from alphazero_gomoku_b200 import MCTS_Pure
    az = MCTSPlayer(net.policy_value_fn, c_puct=5, n_playout=80)
    pure = MCTS_Pure(c_puct=5, n_playout=200, seed=3)
    w2 = Game(Board(width=6, height=6, n_in_row=4)).start_play(az, pure, start_player=1, is_shown=0)
    assert w2 in (1, 2, -1)


def test_collect_selfplay_data_batched():
    """TrainPipeline.collect_selfplay_data(n_games): games played concurrently on the device,
    8x augmentation, reference buffer item types; then one policy_update."""
    from alphazero_gomoku_b200 import TrainPipeline
    conf = {"board_width": 6, "board_height": 6, "n_in_row": 4, "n_playout": 40, "buffer_size": 10000,
            "batch_size": 64, "epochs": 2}
    tp = TrainPipeline(conf, seed=1)
    tp.collect_selfplay_data(n_games=12)
    assert len(tp.last_games) == 12
    plies = sum(len(m) for _, _, m in tp.last_games)
    assert tp.episode_len == plies and len(tp.data_buffer) == 8 * plies
    s, p, z = tp.data_buffer[0]
    assert s.shape == (4, 6, 6) and s.dtype == np.float64 and p.shape == (36,) and z in (-1.0, 0.0, 1.0)
    from oracle import orc
    for gid, winner, moves in tp.last_games:  # every recorded game is a legal, finished game
        ob = orc.Board(6, 6, 4)
        ob.init_board(0)
        for m in moves[:-1]:
            ob.do_move(m)
            assert not ob.game_end()[0]
        ob.do_move(moves[-1])
        assert ob.game_end() == (True, winner)
    before = [a.copy() for a in tp.policy_value_net.get_policy_param()]
    loss, entropy = tp.policy_update()
    after = tp.policy_value_net.get_policy_param()
    assert np.isfinite(loss) and np.isfinite(entropy)
    assert any(not np.array_equal(a, b) for a, b in zip(before, after))
    tp.collect_selfplay_data(n_games=4)  # the engine picks up the updated weights
    assert len(tp.last_games) == 4


def test_batched_arena_policy_evaluate():
    """TrainPipeline.policy_evaluate (train.py:216-276): every evaluation game runs concurrently on
    the device, net player vs pure-MCTS rollouts, alternating first move.  The shipped 6x6 policy
    must beat a weak rollout player clearly; a stronger rollout player must do no worse against
    the same net than a weaker one does."""
    from alphazero_gomoku_b200 import Engine, TrainPipeline
    from alphazero_gomoku_b200.train import arena_play
    from oracle import orc
    w, n, G = 6, 4, 64
    e = Engine(width=w, height=w, n_in_row=n, n_games=G, n_playout=400, seed=5)
    e.set_weights(load_golden_weights("weights_6_6_4"))
    start = np.arange(G) % 2
    winners = arena_play(e, n_playout=200, pure_playouts=100, start_players=start)
    assert set(np.unique(winners)) <= {1, 2, -1}
    ratio_weak = ((winners == 1).sum() + 0.5 * (winners == -1).sum()) / G
    # the boards hold finished, legal games whose winner is what the arena reported
    st = e.boards_status()
    assert st["ended"].all() and np.array_equal(st["winner"], winners)
    cells = e.boards_cells()
    for g in range(0, G, 7):
        n1, n2 = int((cells[g] == 1).sum()), int((cells[g] == 2).sum())
        first, second = (n1, n2) if start[g] == 0 else (n2, n1)
        assert first - second in (0, 1)           # players alternated, the chosen one moved first
    winners2 = arena_play(e, n_playout=200, pure_playouts=400, start_players=start, search_salt=77)
    ratio_strong = ((winners2 == 1).sum() + 0.5 * (winners2 == -1).sum()) / G
    e.close()
    print("arena: net(200) vs pure(100): %.3f, vs pure(400): %.3f" % (ratio_weak, ratio_strong))
    assert ratio_weak >= 0.75
    assert ratio_strong <= ratio_weak + 0.1
    # through the reference-shaped entry point
    tp = TrainPipeline({"board_width": 6, "board_height": 6, "n_in_row": 4, "n_playout": 100,
                        "pure_mcts_playout_num": 60}, seed=2)
    tp.policy_value_net.set_policy_param(load_golden_weights("weights_6_6_4"))
    r = tp.policy_evaluate(n_games=16)
    assert 0.5 <= r <= 1.0
    with pytest.raises(NotImplementedError):
        tp.policy_evaluate(n_games=2, player=1)


def test_device_equi_data_vs_reference_golden():
    """gz_equi_data (the 8 dihedral images on the device) against the live reference's
    get_equi_data fixtures: a pure permutation, so bit-exact; host and device (torch) pointers."""
    import torch
    from alphazero_gomoku_b200 import Engine
    z = np.load(os.path.join(GOLDEN, "equi.npz"))
    for w in (6, 8, 9, 15):
        e = Engine(width=w, height=w, n_in_row=3, n_games=1, n_playout=1, nodes_per_game=4)
        st, pi, zs = z["w%d_states" % w], z["w%d_pis" % w], z["w%d_zs" % w]
        os_, op, oz = e.equi_data(st, pi, zs)
        assert np.array_equal(os_, z["w%d_out_states" % w].astype(np.float32))
        assert np.array_equal(op, z["w%d_out_pis" % w].astype(np.float32))
        assert np.array_equal(oz, z["w%d_out_zs" % w].astype(np.float32))
        ts, tp_, tz = (torch.from_numpy(a.astype(np.float32)).cuda() for a in (st, pi, zs))
        ds, dp, dz = e.equi_data(ts, tp_, tz)
        assert np.array_equal(ds.cpu().numpy(), os_) and np.array_equal(dp.cpu().numpy(), op)
        assert np.array_equal(dz.cpu().numpy(), oz)
        # a large ragged batch through the grid-stride path
        rng = np.random.RandomState(w)
        n = 3001
        bs = (rng.rand(n, 4, w, w) < 0.3).astype(np.float32)
        bp = rng.rand(n, w * w).astype(np.float32)
        bz = rng.choice([-1.0, 0.0, 1.0], size=n).astype(np.float32)
        gs, gp, gz_ = e.equi_data(bs, bp, bz)
        from helpers import dihedral_images
        for i in (0, 1499, 3000):
            ref = dihedral_images(bs[i], bp[i], bz[i], w, w)
            assert np.array_equal(gs[8 * i:8 * i + 8], np.array([r[0] for r in ref]))
            assert np.array_equal(gp[8 * i:8 * i + 8], np.array([r[1] for r in ref]))
            assert (gz_[8 * i:8 * i + 8] == bz[i]).all()
        assert e.equi_data(bs[:0], bp[:0], bz[:0])[0].shape[0] == 0
        e.close()


def test_start_from_train_data_vs_reference_golden():
    """recorded games -> supervised samples (game.py:292-358), one game through the Game
    API and all games through the batched rules call, against the live reference's outputs."""
    from alphazero_gomoku_b200 import Board, Game, TrainPipeline
    from alphazero_gomoku_b200.engine import Engine
    from alphazero_gomoku_b200.train import samples_from_train_data
    z = np.load(os.path.join(GOLDEN, "traindata.npz"))

    class _P(object):
        def reset_player(self):
            self.was_reset = True

    for tag, w, n in (("15", 15, 5), ("6", 6, 4)):
        offs = z["td%s_offsets" % tag]
        games = [[int(z["td%s_winners" % tag][i]), z["td%s_moves" % tag][offs[i]:offs[i + 1]].tolist()]
                 for i in range(len(offs) - 1)]
        want_st, want_pi, want_z = z["td%s_states" % tag].astype(np.float64), z["td%s_pis" % tag], z["td%s_zs" % tag]
        e = Engine(width=w, height=w, n_in_row=n, n_games=1, n_playout=1, nodes_per_game=4)
        got = samples_from_train_data(e, games)
        assert [len(g) for g in got] == [len(g[1]) for g in games]
        flat = [s for g in got for s in g]
        assert np.array_equal(np.array([s[0] for s in flat]), want_st)
        assert np.array_equal(np.array([s[1] for s in flat]), want_pi)
        assert np.array_equal(np.array([s[2] for s in flat]), want_z)
        p = _P()
        winner, data = Game(Board(width=w, height=w, n_in_row=n)).start_from_train_data(p, games[0])
        data = list(data)
        assert winner == games[0][0] and p.was_reset and len(data) == len(games[0][1])
        assert all(np.array_equal(d[0], want_st[i]) and np.array_equal(d[1], want_pi[i]) and d[2] == want_z[i]
                   for i, d in enumerate(data))
        with pytest.raises(ValueError):
            samples_from_train_data(e, [[1, [3, 4, 3]]])  # occupied cell
    tp = TrainPipeline({"board_width": 6, "board_height": 6, "n_in_row": 4, "n_playout": 10, "buffer_size": 500})
    tp.collect_selfplay_data(2, train_data=games)  # games = the 6x6 records
    assert tp.episode_len == 13 and len(tp.data_buffer) == 8 * 13
    assert tp.data_buffer[0][0].shape == (4, 6, 6) and abs(tp.data_buffer[0][1].sum() - (0.99999 + 35e-6)) < 1e-6


@pytest.mark.parametrize("device_buffer", [False, True])
def test_continuous_collect_matches_batch_mode(device_buffer):
    """TrainPipeline(continuous=True): slots never idle, games in flight survive across collect_selfplay_data calls,
    and a call hands out the games with the next n START serials - every game once, in start order, so short games
    are not over-represented.  Checked against the reference-shaped batch mode on the same weights: the game-length
    histograms agree (two-sample chi-square), serials are released consecutively, the buffer receives 8 images per
    recorded ply, and later calls need fewer plies than the first because they find games already in progress."""
    from helpers import chi_square_two_sample
    from alphazero_gomoku_b200.train import TrainPipeline
    conf = {"board_width": 6, "board_height": 6, "n_in_row": 4, "n_playout": 40, "buffer_size": 200000}
    n, calls = 128, 4
    model = os.path.join(GOLDEN, "weights_6_6_4.npz")
    arrs = load_golden_weights("weights_6_6_4")

    def lengths(tp):
        out, plies_per_call = [], []
        for k in range(calls):
            before = tp._cont["plies"] if tp._cont else 0
            tp.collect_selfplay_data(n)
            assert len(tp.last_games) == n
            out += [len(m) for _, _, m in tp.last_games]
            assert tp.episode_len == sum(len(m) for _, _, m in tp.last_games)
            if tp._cont:
                plies_per_call.append(tp._cont["plies"] - before)
                ser = [gid - tp._engine.game_id_base for gid, _, _ in tp.last_games]
                assert ser == list(range(ser[0], ser[0] + n)) and (k == 0 or ser[0] == prev_end)
                prev_end = ser[-1] + 1
        assert len(tp.data_buffer) == 8 * sum(out)
        return np.array(out), plies_per_call

    tp_b = TrainPipeline(conf, seed=3, device_buffer=device_buffer)
    tp_b.policy_value_net.set_policy_param(arrs)
    len_b, _ = lengths(tp_b)
    tp_c = TrainPipeline(conf, seed=4, device_buffer=device_buffer, continuous=True)
    tp_c.policy_value_net.set_policy_param(arrs)
    len_c, plies = lengths(tp_c)
    hist = lambda a: np.bincount(a, minlength=37).astype(np.float64)
    chi_square_two_sample(hist(len_b), hist(len_c), "game lengths, batch vs continuous")
    assert abs(len_b.mean() - len_c.mean()) < 4 * np.sqrt(len_b.var() / len(len_b) + len_c.var() / len(len_c)) + 1e-9
    assert max(plies[1:]) < plies[0], plies  # games were already in progress
    # taking the first n games to FINISH would be biased: the continuous mode's first call is not shorter on average
    first_call = len_c[:n]
    assert first_call.mean() > len_b.mean() - 4 * np.sqrt(len_b.var() / n)
    print("batch mean length %.2f, continuous %.2f (first call %.2f); plies stepped per call %s"
          % (len_b.mean(), len_c.mean(), first_call.mean(), plies))
