"""Generate the golden fixtures in tests/golden/ by IMPORTING the unmodified Python reference
from /root/reference (build container only).  Run:  python tests/golden/make_golden.py

Fixtures (all small):
  weights_6_6_4.npz / weights_8_8_5.npz  shipped Theano pickles converted to the torch
        state_dict layout — data fixtures, not reference source.
  rules.npz   seeded random positions for 6 board shapes with the reference Board's
        availables / current_state / has_a_winner / game_end / move_target / current player.
  corpus.npz  the first recorded 15x15 games of logs/steps.log + human_gomoku.txt and 9x9
        games of run.log with the reference's per-ply game_end results.
  net.npz     planes and the reference PyTorch fp32 Net outputs (exp(log_softmax), value).
  mcts.npz    reference MCTS.get_move_probs visit counts (400 playouts, c_puct 5, shipped
        weights, value wrapped as float — SURVEY 8c-3) on seeded positions, with tree reuse
        over two consecutive searches.
  equi.npz    get_equi_data (train.py:116-135) outputs on random samples.
  train_step.npz  three reference train_step calls (loss, entropy, final parameters) from the shipped
        weights on seeded batches.
  traindata.npz  Game.start_from_train_data samples of recorded games (game.py:292-358).
  rollout.npz reference _evaluate_rollout outcome counts from fixed positions (distribution
        check for the Philox rollout kernel) and pure-MCTS get_move visit statistics.
"""
import ast
import os
import random
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402

assert rh.available(), "needs /root/reference"
G = rh.game()
AZ = rh.mcts_alphazero()
PURE = rh.mcts_pure()


def save(name, **kw):
    np.savez_compressed(os.path.join(HERE, name), **kw)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in kw.items()})


def make_weights():
    out = {}
    for name, tag in (("best_policy_6_6_4.model", "6_6_4"), ("best_policy_8_8_5.model", "8_8_5")):
        arrs = rh.load_theano_weights(name)
        save("weights_%s.npz" % tag, **{"p%02d" % i: a for i, a in enumerate(arrs)})
        out[tag] = arrs
    return out


def ref_board_outputs(b):
    end, winner = b.game_end()
    hw, _ = b.has_a_winner()
    S = b.width * b.height
    legal = np.zeros(S, np.uint8)
    legal[b.availables] = 1
    b.get_move_target()
    tg = np.zeros(S, np.uint8)
    tg[b.move_target] = 1
    planes = None
    if b.width == b.height:
        planes = np.ascontiguousarray(b.current_state()).astype(np.float32)
    return int(end), int(winner), int(hw), legal, tg, planes, int(b.current_player)


def make_rules():
    rng = np.random.RandomState(0)
    shapes = [(6, 6, 4), (8, 8, 5), (9, 9, 5), (15, 15, 5), (3, 3, 3), (7, 7, 3), (8, 6, 4), (11, 11, 5)]
    kw = {}
    for (w, h, n) in shapes:
        S = w * h
        moves_all, offs, starts = [], [0], []
        ended, winner, haswin, legal, target, planes, cur = [], [], [], [], [], [], []
        for t in range(160):
            sp = t % 2
            b = G.Board(width=w, height=h, n_in_row=n)
            b.init_board(sp)
            k = int(rng.randint(0, S + 1))
            mv = []
            for _ in range(k):
                av = b.availables
                m = int(av[rng.randint(len(av))])
                # bias towards extending lines so that wins occur
                if mv and rng.rand() < 0.5:
                    cand = [x for x in (mv[-2] + d if len(mv) > 1 else mv[-1] + d
                                        for d in (1, -1, w, -w, w + 1, -w - 1, w - 1, -w + 1)) if x in av]
                    if cand:
                        m = int(cand[rng.randint(len(cand))])
                b.do_move(m)
                mv.append(m)
                if b.game_end()[0]:
                    break
            e, wi, hw, lg, tg, pl, cp = ref_board_outputs(b)
            moves_all += mv
            offs.append(len(moves_all))
            starts.append(sp)
            ended.append(e); winner.append(wi); haswin.append(hw); legal.append(lg); target.append(tg)
            cur.append(cp)
            if pl is not None:
                planes.append(pl)
        tag = "%dx%dx%d" % (w, h, n)
        kw[tag + "_moves"] = np.array(moves_all, np.int32)
        kw[tag + "_offsets"] = np.array(offs, np.int64)
        kw[tag + "_start"] = np.array(starts, np.int8)
        kw[tag + "_ended"] = np.array(ended, np.int8)
        kw[tag + "_winner"] = np.array(winner, np.int8)
        kw[tag + "_haswin"] = np.array(haswin, np.int8)
        kw[tag + "_legal"] = np.array(legal, np.uint8)
        kw[tag + "_target"] = np.array(target, np.uint8)
        kw[tag + "_cur"] = np.array(cur, np.int8)
        if planes:
            kw[tag + "_planes"] = np.array(planes, np.float32)
    save("rules.npz", **kw)


def replay_game_end(w, n, moves):
    """per-ply reference (ended, winner) for one recorded game; stops at first illegal move."""
    b = G.Board(width=w, height=w, n_in_row=n)
    b.init_board(0)
    res = []
    for m in moves:
        if m not in b.availables:
            break
        b.do_move(m)
        e, wi = b.game_end()
        res.append((int(e), int(wi)))
    return res


def make_corpus():
    games15 = []
    with open(os.path.join(rh.REF_DIR, "logs", "steps.log")) as f:
        for line in f:
            line = line.strip()
            if not line.startswith("{"):
                continue
            d = ast.literal_eval(line)
            games15.append([int(x) for x in d["steps"]])
            if len(games15) >= 120:
                break
    with open(os.path.join(rh.REF_DIR, "human_gomoku.txt")) as f:
        cnt = 0
        for line in f:
            line = line.strip()
            if not line.startswith("["):
                continue
            d = ast.literal_eval(line)
            games15.append([int(x) for x in d[1]])
            cnt += 1
            if cnt >= 80:
                break
    games9 = []
    with open(os.path.join(rh.REF_DIR, "run.log"), errors="ignore") as f:
        for line in f:
            m = re.search(r"\d-\[([0-9, ]+)\]", line)
            if m:
                mv = [int(x) for x in m.group(1).split(",")]
                if mv and max(mv) < 81:
                    games9.append(mv)
    kw = {}
    for tag, w, games in (("15", 15, games15), ("9", 9, games9[:120])):
        flat, offs, ended, winner = [], [0], [], []
        for mv in games:
            res = replay_game_end(w, 5, mv)
            mv = mv[:len(res)]
            flat += mv
            offs.append(len(flat))
            ended += [r[0] for r in res]
            winner += [r[1] for r in res]
        kw["g%s_moves" % tag] = np.array(flat, np.int32)
        kw["g%s_offsets" % tag] = np.array(offs, np.int64)
        kw["g%s_ended" % tag] = np.array(ended, np.int8)     # per ply (prefix) results
        kw["g%s_winner" % tag] = np.array(winner, np.int8)
    save("corpus.npz", **kw)


def make_net(weights):
    import torch
    kw = {}
    rng = np.random.RandomState(1)
    for tag, (w, n) in (("6_6_4", (6, 4)), ("8_8_5", (8, 5))):
        net = rh.make_ref_net(w, w, weights[tag])
        pls = []
        while len(pls) < 48:
            b, _ = rh.random_position(G.Board, w, w, n, rng)
            if b.game_end()[0]:
                continue
            pls.append(np.ascontiguousarray(b.current_state()).astype(np.float32))
        x = np.stack(pls)
        probs, vals = net.policy_value(x)
        kw[tag + "_planes"] = x
        kw[tag + "_probs"] = probs.astype(np.float32)
        kw[tag + "_values"] = vals.astype(np.float32).reshape(-1)
    save("net.npz", **kw)


def make_mcts(weights):
    kw = {}
    rng = np.random.RandomState(2)
    for tag, (w, n) in (("6_6_4", (6, 4)), ("8_8_5", (8, 5))):
        net = rh.make_ref_net(w, w, weights[tag])
        fn = rh.float_value_fn(net)
        moves_all, offs, vis1, vis2, mv2 = [], [0], [], [], []
        cnt = 0
        while cnt < 6:
            b, moves = rh.random_position(G.Board, w, w, n, rng, max_plies=12)
            if b.game_end()[0] or len(b.availables) < 3:
                continue
            m = AZ.MCTS(fn, c_puct=5, n_playout=400)
            acts, _ = m.get_move_probs(b, temp=1.0)
            v = np.full(w * w, -1, np.int32)
            for a in acts:
                v[a] = m._root._children[a]._n_visits
            # tree reuse: play the most visited move, search again (mcts_alphaZero.py:183-192)
            best = int(np.argmax(v))
            m.update_with_move(best)
            b.do_move(best)
            v2 = np.full(w * w, -1, np.int32)
            if not b.game_end()[0]:
                acts2, _ = m.get_move_probs(b, temp=1.0)
                for a in acts2:
                    v2[a] = m._root._children[a]._n_visits
            moves_all += moves
            offs.append(len(moves_all))
            vis1.append(v); vis2.append(v2); mv2.append(best)
            cnt += 1
            print("mcts", tag, cnt)
        kw[tag + "_moves"] = np.array(moves_all, np.int32)
        kw[tag + "_offsets"] = np.array(offs, np.int64)
        kw[tag + "_visits1"] = np.array(vis1, np.int32)
        kw[tag + "_visits2"] = np.array(vis2, np.int32)
        kw[tag + "_move"] = np.array(mv2, np.int32)
    save("mcts.npz", **kw)


def make_rollout():
    np.random.seed(0)
    random.seed(0)
    kw = {}
    positions = {
        "6x6x4_a": (6, 4, [14, 15, 20, 21]),
        "6x6x4_b": (6, 4, [14, 8, 15, 9, 16]),
        "8x8x5_a": (8, 5, [27, 28, 35, 36, 19]),
        "8x8x5_b": (8, 5, []),
    }
    for tag, (w, n, moves) in positions.items():
        b = G.Board(width=w, height=w, n_in_row=n)
        b.init_board(0)
        for m in moves:
            b.do_move(m)
        mc = PURE.MCTS(PURE.policy_value_fn, 5, 10)
        counts = np.zeros(3, np.int64)  # [-1, 0, +1]
        import copy
        with rh.quiet():
            for _ in range(3000):
                r = mc._evaluate_rollout(copy.deepcopy(b))
                counts[r + 1] += 1
        kw[tag + "_moves"] = np.array(moves, np.int32)
        kw[tag + "_counts"] = counts
        print("rollout", tag, counts)
    # pure MCTS move choice statistics from one position (first move distribution, 30 runs)
    w, n, moves = 6, 4, [14, 15, 20]
    first = np.zeros(36, np.int64)
    for r in range(30):
        b = G.Board(width=w, height=w, n_in_row=n)
        b.init_board(0)
        for m in moves:
            b.do_move(m)
        p = PURE.MCTSPlayer(c_puct=5, n_playout=300)
        with rh.quiet():
            mv = p.get_action(b)
        first[mv] += 1
    kw["pure_moves"] = np.array(moves, np.int32)
    kw["pure_first"] = first
    save("rollout.npz", **kw)


def make_equi():
    """get_equi_data (train.py:116-135) of the live reference on random samples."""
    kw = {}
    rng = np.random.RandomState(17)
    for w in (6, 8, 9, 15):
        n = 5
        states = (rng.rand(n, 4, w, w) < 0.4).astype(np.float64)
        pis = rng.dirichlet(np.ones(w * w), size=n)
        zs = rng.choice([-1.0, 0.0, 1.0], size=n)
        out = rh.equi_data([(states[i], pis[i], zs[i]) for i in range(n)], w, w)
        kw["w%d_states" % w], kw["w%d_pis" % w], kw["w%d_zs" % w] = states, pis, zs
        kw["w%d_out_states" % w] = np.array([o[0] for o in out])
        kw["w%d_out_pis" % w] = np.array([o[1] for o in out])
        kw["w%d_out_zs" % w] = np.array([o[2] for o in out])
    save("equi.npz", **kw)


def make_train(weights):
    """PolicyValueNet.train_step (policy_value_net_pytorch.py:117-148) of the live reference, CPU fp32:
    three Adam steps from the shipped weights on seeded batches of reference positions.  Stores the
    batches, (loss, entropy) per step and every 5th element of the final flat parameter vector."""
    kw = {}
    for tag, w, n_row in (("6_6_4", 6, 4), ("8_8_5", 8, 5)):
        rng = np.random.RandomState(23 + w)
        S = w * w
        lrs = [2e-3, 3e-3, 1.3e-3]
        batches = []
        for _ in lrs:
            st, pi, z = [], [], []
            while len(st) < 48:
                b, moves = rh.random_position(G.Board, w, w, n_row, rng)
                if b.game_end()[0]:
                    continue
                st.append(np.ascontiguousarray(b.current_state()))
                p = np.zeros(S)
                p[b.availables] = rng.dirichlet(0.3 * np.ones(len(b.availables)))
                pi.append(p)
                z.append(float(rng.choice([-1.0, 0.0, 1.0])))
            batches.append((np.array(st), np.array(pi), np.array(z)))
        out, final = rh.train_steps(w, w, weights[tag], batches, lrs)
        flat = np.concatenate([a.ravel() for a in final]).astype(np.float32)
        kw["t%s_states" % tag] = np.array([b[0] for b in batches]).astype(np.uint8)
        kw["t%s_pis" % tag] = np.array([b[1] for b in batches]).astype(np.float32)
        kw["t%s_zs" % tag] = np.array([b[2] for b in batches]).astype(np.float32)
        kw["t%s_lrs" % tag] = np.array(lrs, np.float64)
        kw["t%s_loss_entropy" % tag] = np.array(out, np.float64)
        kw["t%s_final_every5" % tag] = flat[::5].copy()
        kw["t%s_final_sums" % tag] = np.array([[a.astype(np.float64).sum(), np.abs(a.astype(np.float64)).sum()] for a in final])
        print(tag, out)
    save("train_step.npz", **kw)


def make_traindata():
    """Game.start_from_train_data (game.py:292-358) of the live reference on recorded games: the first
    three 15x15 games of human_gomoku.txt ([winner, moves] records) and two hand-made 6x6 games (one
    tie).  Stores the records and the reference's (state, pi, z) samples."""
    class _P(object):
        def reset_player(self):
            pass
    games15 = []
    with open(os.path.join(rh.REF_DIR, "human_gomoku.txt")) as f:
        for line in f:
            line = line.strip()
            if line.startswith("["):
                d = ast.literal_eval(line)
                games15.append([int(d[0]), [int(x) for x in d[1]]])
                if len(games15) >= 3:
                    break
    games6 = [[2, [14, 15, 20, 21, 8, 9, 27, 3]], [-1, [0, 35, 5, 30, 17]]]
    kw = {}
    for tag, w, n, games in (("15", 15, 5, games15), ("6", 6, 4, games6)):
        g = G.Game(G.Board(width=w, height=w, n_in_row=n))
        st, pi, z, offs, flat, winners = [], [], [], [0], [], []
        for rec in games:
            winner, data = g.start_from_train_data(_P(), rec)
            data = list(data)
            st += [np.ascontiguousarray(d[0]) for d in data]
            pi += [d[1] for d in data]
            z += [d[2] for d in data]
            flat += rec[1]
            offs.append(len(flat))
            winners.append(rec[0])
        kw["td%s_moves" % tag] = np.array(flat, np.int32)
        kw["td%s_offsets" % tag] = np.array(offs, np.int64)
        kw["td%s_winners" % tag] = np.array(winners, np.int32)
        kw["td%s_states" % tag] = np.array(st).astype(np.uint8)
        kw["td%s_pis" % tag] = np.array(pi, np.float64)
        kw["td%s_zs" % tag] = np.array(z, np.float64)
    save("traindata.npz", **kw)


if __name__ == "__main__":
    which = sys.argv[1:] or ["weights", "rules", "corpus", "net", "mcts", "rollout"]
    weights = make_weights() if {"weights", "net", "mcts", "train"} & set(which) else None
    if "rules" in which: make_rules()
    if "corpus" in which: make_corpus()
    if "net" in which: make_net(weights)
    if "mcts" in which: make_mcts(weights)
    if "rollout" in which: make_rollout()
    if "equi" in which: make_equi()
    if "train" in which: make_train(weights)
    if "traindata" in which: make_traindata()
