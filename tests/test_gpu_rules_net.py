"""GPU parity tests (through the C ABI): Board rules bit-exact, net within 1e-4 absolute.

Oracle = oracle/ (C restatement pinned against the reference) + committed golden vectors
generated from the reference itself (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from helpers import GOLDEN, load_golden_weights

pytestmark = pytest.mark.gpu

SHAPES = [(6, 6, 4), (8, 8, 5), (9, 9, 5), (15, 15, 5), (3, 3, 3), (7, 7, 3), (8, 6, 4), (11, 11, 5)]


def _engine(w, h, n, **kw):
    from alphazero_gomoku_b200 import Engine
    return Engine(width=w, height=h, n_in_row=n, **kw)


@pytest.mark.parametrize("w,h,n", SHAPES)
def test_rules_golden(w, h, n):
    z = np.load(os.path.join(GOLDEN, "rules.npz"))
    tag = "%dx%dx%d" % (w, h, n)
    moves, offs = z[tag + "_moves"], z[tag + "_offsets"]
    lists = [moves[offs[i]:offs[i + 1]] for i in range(len(offs) - 1)]
    e = _engine(w, h, n)
    want = ["status", "ended", "winner", "has_win", "legal", "target", "cur_player"]
    if w == h:
        want.append("planes")
    out = e.rules_eval(lists, z[tag + "_start"], want=want)
    assert (out["status"] == 0).all()
    assert np.array_equal(out["ended"], z[tag + "_ended"])
    assert np.array_equal(out["winner"], z[tag + "_winner"])
    assert np.array_equal(out["has_win"], z[tag + "_haswin"])
    assert np.array_equal(out["legal"], z[tag + "_legal"])
    assert np.array_equal(out["target"], z[tag + "_target"])
    assert np.array_equal(out["cur_player"], z[tag + "_cur"])
    if w == h:
        assert np.array_equal(out["planes"], z[tag + "_planes"])
    e.close()


@pytest.mark.parametrize("w,n", [(6, 4), (8, 5), (9, 5), (15, 5), (12, 6)])
def test_rules_vs_oracle_random(w, n):
    from oracle import orc
    rng = np.random.RandomState(w * 100 + n)
    S = w * w
    lists, exp = [], []
    for t in range(600):
        ob = orc.Board(w, w, n)
        ob.init_board(t % 2)
        k = int(rng.randint(0, S + 1))
        mv = []
        for _ in range(k):
            av = ob.availables
            m = av[rng.randint(len(av))]
            if mv and rng.rand() < 0.6:
                cand = [x for x in (mv[-1] + d for d in (2, -2, 2 * w, -2 * w, 2 * w + 2, -2 * w - 2,
                                                        2 * w - 2, -2 * w + 2, 1, w, w + 1, w - 1)) if x in av]
                if len(mv) > 1:
                    cand += [x for x in (mv[-2] + d for d in (1, -1, w, -w, w + 1, -w - 1, w - 1, 1 - w)) if x in av]
                if cand:
                    m = cand[rng.randint(len(cand))]
            ob.do_move(m)
            mv.append(m)
            if ob.game_end()[0]:
                break
        ob.get_move_target()
        lists.append(mv)
        exp.append((ob.game_end(), ob.has_a_winner()[0], ob.availables, ob.move_target,
                    ob.current_state().astype(np.float32), ob.current_player, t % 2))
    e = _engine(w, w, n)
    out = e.rules_eval(lists, [x[6] for x in exp])
    n_end = 0
    for i, (ge, hw, av, tg, st, cp, _) in enumerate(exp):
        assert (bool(out["ended"][i]), int(out["winner"][i])) == ge
        assert bool(out["has_win"][i]) == hw
        assert np.nonzero(out["legal"][i])[0].tolist() == av
        assert np.nonzero(out["target"][i])[0].tolist() == tg
        assert np.array_equal(out["planes"][i], st)
        assert int(out["cur_player"][i]) == cp
        n_end += ge[0]
    assert n_end > 20  # the sampler does produce finished games
    e.close()


@pytest.mark.parametrize("tag,w", [("15", 15), ("9", 9)])
def test_corpus_replay_every_ply(tag, w):
    """Every ply of the recorded games (logs/steps.log, human_gomoku.txt, run.log): the
    game_end of each prefix must equal the reference's."""
    z = np.load(os.path.join(GOLDEN, "corpus.npz"))
    moves, offs = z["g%s_moves" % tag], z["g%s_offsets" % tag]
    lists = []
    for i in range(len(offs) - 1):
        g = moves[offs[i]:offs[i + 1]]
        for p in range(1, len(g) + 1):
            lists.append(g[:p])
    e = _engine(w, w, 5)
    out = e.rules_eval(lists, want=("status", "ended", "winner"))
    assert (out["status"] == 0).all()
    assert np.array_equal(out["ended"], z["g%s_ended" % tag])
    assert np.array_equal(out["winner"], z["g%s_winner" % tag])
    assert out["ended"].sum() > 10
    e.close()


def test_illegal_moves_and_errors():
    from alphazero_gomoku_b200 import Engine
    e = _engine(6, 6, 4, n_games=3)
    out = e.rules_eval([[1, 2, 1], [40], [0, 1, 2]], want=("status",))
    assert out["status"].tolist() == [-2, -2, 0]
    e.boards_do_move([5, 6, 7])
    with pytest.raises(ValueError):
        e.boards_do_move([5, -1, -1])
    st = e.boards_status()
    assert st["n_stones"].tolist() == [1, 1, 1] and st["last_move"].tolist() == [5, 6, 7]
    with pytest.raises(ValueError):
        e.boards_set(0, [3, 3])
    e.close()
    with pytest.raises(Exception):
        Engine(width=4, height=4, n_in_row=5)  # game.py:61-62


def test_resident_boards_match_oracle():
    from oracle import orc
    w, n, G = 8, 5, 64
    rng = np.random.RandomState(5)
    e = _engine(w, w, n, n_games=G)
    obs = [orc.Board(w, w, n) for _ in range(G)]
    for b in obs:
        b.init_board(0)
    for ply in range(40):
        mv = np.full(G, -1, np.int32)
        for g, b in enumerate(obs):
            if b.game_end()[0]:
                continue
            av = b.availables
            mv[g] = av[rng.randint(len(av))]
            b.do_move(int(mv[g]))
        e.boards_do_move(mv)
        st = e.boards_status()
        planes = e.boards_planes()
        legal = e.boards_legal_mask()
        cells = e.boards_cells()
        for g, b in enumerate(obs):
            assert (bool(st["ended"][g]), int(st["winner"][g])) == b.game_end()
            assert int(st["cur_player"][g]) == b.current_player
            assert np.array_equal(planes[g], b.current_state().astype(np.float32))
            assert np.nonzero(legal[g])[0].tolist() == b.availables
            assert {m: int(cells[g][m]) for m in np.nonzero(cells[g])[0]} == b.states
    e.close()


# ---- net: 1e-4 absolute on probabilities and value (north_star tolerance) ----------------------
NET_TOL = 1e-4


@pytest.mark.parametrize("tag,w,n", [("6_6_4", 6, 4), ("8_8_5", 8, 5)])
def test_net_golden(tag, w, n):
    """Device forward vs the reference PyTorch fp32 Net outputs on shipped weights."""
    z = np.load(os.path.join(GOLDEN, "net.npz"))
    arrs = load_golden_weights("weights_" + tag)
    e = _engine(w, w, n)
    e.set_weights(arrs)
    probs, vals = e.net_forward(z[tag + "_planes"])
    dp = np.abs(probs - z[tag + "_probs"]).max()
    dv = np.abs(vals - z[tag + "_values"]).max()
    print("net golden", tag, "max|dp|", dp, "max|dv|", dv)
    assert dp < NET_TOL and dv < NET_TOL
    assert np.allclose(probs.sum(1), 1.0, atol=1e-5)
    got = e.get_weights()
    for a, b in zip(arrs, got):
        assert np.array_equal(a.reshape(-1), b.reshape(-1))
    e.close()


def _torch_forward(arrs, planes):
    """plain PyTorch fp32 CPU reference of the same op (policy_value_net_pytorch.py:43-57)."""
    import torch
    import torch.nn.functional as F
    p = [torch.from_numpy(np.ascontiguousarray(a)) for a in arrs]
    x = torch.from_numpy(planes)
    with torch.no_grad():
        x = F.relu(F.conv2d(x, p[0], p[1], padding=1))
        x = F.relu(F.conv2d(x, p[2], p[3], padding=1))
        x = F.relu(F.conv2d(x, p[4], p[5], padding=1))
        a = F.relu(F.conv2d(x, p[6], p[7])).flatten(1)
        a = F.log_softmax(F.linear(a, p[8], p[9]), dim=1)
        v = F.relu(F.conv2d(x, p[10], p[11])).flatten(1)
        v = torch.tanh(F.linear(F.relu(F.linear(v, p[12], p[13])), p[14], p[15]))
    return np.exp(a.numpy()), v.numpy().reshape(-1)


def _random_weights(rng, w, trunk=(32, 64, 128), scale=1.0):
    from alphazero_gomoku_b200.engine import weight_shapes
    arrs = []
    for s in weight_shapes(w, w, trunk):
        fan = int(np.prod(s[1:])) if len(s) > 1 else 1
        if len(s) > 1:
            arrs.append((rng.randn(*s) * scale * np.sqrt(2.0 / fan)).astype(np.float32))
        else:
            arrs.append((rng.randn(*s) * 0.1).astype(np.float32))
    return arrs


@pytest.mark.parametrize("w,trunk,B", [(6, (32, 64, 128), 37), (8, (32, 64, 128), 300), (9, (32, 64, 128), 65),
                                       (15, (32, 64, 128), 21), (15, (64, 128, 256), 9), (5, (32, 64, 128), 3)])
def test_net_vs_torch_and_oracle(w, trunk, B):
    from oracle import orc
    rng = np.random.RandomState(w + B)
    arrs = _random_weights(rng, w, trunk, scale=1.5)
    planes = (rng.rand(B, 4, w, w) < 0.35).astype(np.float32)
    planes[0] = 0.0
    e = _engine(w, w, min(w, 5), trunk=trunk)
    e.set_weights(arrs)
    probs, vals = e.net_forward(planes)
    tp, tv = _torch_forward(arrs, planes)
    olp, ov = orc.Net(w, w, arrs).forward(planes[:8])
    print("net", w, trunk, "vs torch", np.abs(probs - tp).max(), np.abs(vals - tv).max(),
          "vs oracle", np.abs(probs[:8] - np.exp(olp)).max(), np.abs(vals[:8] - ov).max())
    assert np.abs(probs - tp).max() < NET_TOL and np.abs(vals - tv).max() < NET_TOL
    assert np.abs(probs[:8] - np.exp(olp)).max() < NET_TOL and np.abs(vals[:8] - ov).max() < NET_TOL
    # non-binary inputs are legal for policy_value(state_batch) too
    x = rng.rand(5, 4, w, w).astype(np.float32)
    p2, v2 = e.net_forward(x)
    tp2, tv2 = _torch_forward(arrs, x)
    assert np.abs(p2 - tp2).max() < NET_TOL and np.abs(v2 - tv2).max() < NET_TOL
    e.close()


def test_net_theano_layout_and_device_pointers():
    import torch
    w = 6
    arrs = load_golden_weights("weights_6_6_4")
    theano = []
    for a in arrs:
        if a.ndim == 4:
            theano.append(np.ascontiguousarray(a[:, :, ::-1, ::-1]))
        elif a.ndim == 2:
            theano.append(np.ascontiguousarray(a.T))
        else:
            theano.append(a)
    from alphazero_gomoku_b200 import _lib
    e = _engine(w, w, 4)
    e.set_weights(theano, layout=_lib.LAYOUT_THEANO)
    z = np.load(os.path.join(GOLDEN, "net.npz"))
    probs, vals = e.net_forward(z["6_6_4_planes"])
    assert np.abs(probs - z["6_6_4_probs"]).max() < NET_TOL
    # torch CUDA tensors in, torch CUDA tensors out (zero copy)
    xp = torch.from_numpy(z["6_6_4_planes"]).cuda()
    pp, vv = e.net_forward(xp)
    e.sync()
    assert np.abs(pp.cpu().numpy() - z["6_6_4_probs"]).max() < NET_TOL
    assert np.abs(vv.cpu().numpy() - z["6_6_4_values"]).max() < NET_TOL
    e.close()


def test_net_requires_weights():
    from alphazero_gomoku_b200._lib import GzError
    e = _engine(6, 6, 4)
    with pytest.raises(GzError):
        e.net_forward(np.zeros((1, 4, 6, 6), np.float32))
    e.close()
