"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_build/liborc.so (the plain-C restatement of the reference's CPU
path, see orc_board.h).  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liborc.so")
MAXS = 225

EVAL_NET, EVAL_HASH, EVAL_PURE, EVAL_CALLBACK = 0, 1, 2, 3


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("orc_board.c", "orc_net.c", "orc_mcts.c", "orc_board.h")]
    if (not force and os.path.exists(_SO)
            and all(os.path.getmtime(_SO) >= os.path.getmtime(s) for s in srcs)):
        return _SO
    subprocess.check_call(["make", "-s", "-C", _HERE, "-B"])
    return _SO


class CBoard(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("n_in_row", C.c_int),
                ("states", C.c_int * MAXS), ("availables", C.c_int * MAXS),
                ("n_avail", C.c_int), ("n_moved", C.c_int), ("last_move", C.c_int),
                ("current_player", C.c_int), ("move_target", C.c_int * MAXS),
                ("n_target", C.c_int)]


class CNet(C.Structure):
    _fields_ = [("W", C.c_int), ("H", C.c_int), ("c1", C.c_int), ("c2", C.c_int),
                ("c3", C.c_int), ("p", C.POINTER(C.c_float) * 16)]


class CReplayStats(C.Structure):
    _fields_ = [("max_regret_rel", C.c_double), ("max_regret_margin", C.c_double), ("min_gap_rel", C.c_double),
                ("min_gap_margin", C.c_double), ("n_decisions", C.c_longlong), ("n_deviations", C.c_longlong),
                ("n_tie_order", C.c_longlong)]


EVAL_CB = C.CFUNCTYPE(None, C.POINTER(CBoard), C.POINTER(C.c_float), C.POINTER(C.c_float))

_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        assert L.orc_board_sizeof() == C.sizeof(CBoard)
        L.orc_mcts_new.restype = C.c_void_p
        L.orc_mcts_new.argtypes = [C.c_double, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_mcts_free.argtypes = [C.c_void_p]
        L.orc_mcts_set_stream.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.orc_mcts_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
        L.orc_mcts_replay.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_replay_stats_init.argtypes = [C.c_void_p]
        L.orc_mcts_search_vl.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_mcts_root_visits.argtypes = [C.c_void_p]
        L.orc_mcts_root_q.argtypes = [C.c_void_p]
        L.orc_mcts_root_q.restype = C.c_double
        L.orc_mcts_stat.argtypes = [C.c_void_p, C.c_int]
        L.orc_mcts_stat.restype = C.c_longlong
        L.orc_mcts_update_with_move.argtypes = [C.c_void_p, C.c_int]
        L.orc_pure_get_move.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_rollout_once.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_hash_eval.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_net_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_philox4x32.argtypes = [C.c_uint32] * 6 + [C.c_void_p]
        _lib = L
    return _lib


class Board:
    """Oracle board with the reference's attribute surface (game.py:11-172)."""

    def __init__(self, width=8, height=8, n_in_row=5):
        self.c = CBoard()
        self.width, self.height, self.n_in_row = int(width), int(height), int(n_in_row)
        self.players = [1, 2]

    def init_board(self, start_player=0):
        rc = lib().orc_board_init(C.byref(self.c), self.width, self.height, self.n_in_row,
                                  start_player)
        if rc == -1:
            raise Exception("board width and height can not be less than ", self.n_in_row)
        if rc != 0:
            raise ValueError("bad board parameters")

    def copy(self):
        b = Board(self.width, self.height, self.n_in_row)
        C.memmove(C.byref(b.c), C.byref(self.c), C.sizeof(CBoard))
        return b

    def do_move(self, move):
        if lib().orc_board_do_move(C.byref(self.c), int(move)) != 0:
            raise ValueError("list.remove(x): x not in list")

    def has_a_winner(self):
        w = C.c_int()
        r = lib().orc_board_has_a_winner(C.byref(self.c), C.byref(w))
        return bool(r), w.value

    def game_end(self):
        w = C.c_int()
        r = lib().orc_board_game_end(C.byref(self.c), C.byref(w))
        return bool(r), w.value

    def current_state(self):
        out = np.zeros((4, self.width, self.height), dtype=np.float64)
        if lib().orc_board_current_state(C.byref(self.c), out.ctypes.data) != 0:
            raise ValueError("current_state needs a square board")
        return out

    def get_move_target(self):
        lib().orc_board_get_move_target(C.byref(self.c))

    @property
    def move_target(self):
        return list(self.c.move_target[:self.c.n_target])

    @property
    def availables(self):
        return list(self.c.availables[:self.c.n_avail])

    @property
    def states(self):
        S = self.width * self.height
        return {m: self.c.states[m] for m in range(S) if self.c.states[m]}

    @property
    def last_move(self):
        return self.c.last_move

    @property
    def current_player(self):
        return self.c.current_player

    def get_current_player(self):
        return self.c.current_player


STATE_DICT_KEYS = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "conv3.weight",
                   "conv3.bias", "act_conv1.weight", "act_conv1.bias", "act_fc1.weight",
                   "act_fc1.bias", "val_conv1.weight", "val_conv1.bias", "val_fc1.weight",
                   "val_fc1.bias", "val_fc2.weight", "val_fc2.bias"]


class Net:
    """Oracle forward of reference Net (policy_value_net_pytorch.py:25-57).

    params: list of 16 float32 arrays in torch state_dict order and layout."""

    def __init__(self, width, height, params):
        self.width, self.height = int(width), int(height)
        self.params = [np.ascontiguousarray(np.asarray(p, dtype=np.float32)) for p in params]
        assert len(self.params) == 16
        self.c = CNet()
        self.c.W, self.c.H = self.width, self.height
        self.c.c1, self.c.c2, self.c.c3 = (self.params[0].shape[0], self.params[2].shape[0],
                                           self.params[4].shape[0])
        for i, p in enumerate(self.params):
            self.c.p[i] = p.ctypes.data_as(C.POINTER(C.c_float))

    def forward(self, planes):
        planes = np.ascontiguousarray(np.asarray(planes, dtype=np.float32))
        B = planes.shape[0]
        S = self.width * self.height
        logp = np.zeros((B, S), dtype=np.float32)
        val = np.zeros((B,), dtype=np.float32)
        rc = lib().orc_net_forward(C.byref(self.c), planes.ctypes.data, B, logp.ctypes.data,
                                   val.ctypes.data)
        assert rc == 0
        return logp, val

    def policy_value_fn(self, board):
        """policy_value_net_pytorch.py:96-115 (value as a Python float)."""
        st = np.ascontiguousarray(board.current_state().reshape(-1, 4, self.width, self.height))
        logp, val = self.forward(st.astype(np.float32))
        probs = np.exp(logp.flatten())
        legal = board.availables
        return zip(legal, probs[legal]), float(val[0])


class MCTS:
    """Oracle MCTS (mcts_alphaZero.py:95-192 / mcts_pure.py:124-243)."""

    def __init__(self, c_puct=5, n_playout=400, eval_kind=EVAL_HASH, net=None, callback=None,
                 salt=0):
        self._net = net
        self._cb = EVAL_CB(callback) if callback is not None else None
        self.h = lib().orc_mcts_new(float(c_puct), int(n_playout), int(eval_kind),
                                    C.addressof(net.c) if net is not None else None,
                                    C.cast(self._cb, C.c_void_p) if self._cb else None,
                                    C.c_uint64(salt))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mcts_free(self.h)
            self.h = None

    def set_stream(self, game_id, search_id):
        lib().orc_mcts_set_stream(self.h, game_id, search_id)

    def search(self, board, n_playout=-1):
        """Run playouts; returns (acts, visits, Q, P) of the root children."""
        acts = np.zeros(MAXS, dtype=np.int32)
        visits = np.zeros(MAXS, dtype=np.int32)
        q = np.zeros(MAXS, dtype=np.float64)
        p = np.zeros(MAXS, dtype=np.float64)
        k = lib().orc_mcts_search(self.h, C.addressof(board.c) if board is not None else None, n_playout, acts.ctypes.data,
                                  visits.ctypes.data, q.ctypes.data, p.ctypes.data)
        return acts[:k].copy(), visits[:k].copy(), q[:k].copy(), p[:k].copy()

    def get_move_probs(self, board, temp=1e-3):
        """mcts_alphaZero.py:154-181 including the float64 softmax of :179."""
        acts, visits, _, _ = self.search(board)
        x = 1.0 / temp * np.log(np.array(visits) + 1e-10)
        probs = np.exp(x - np.max(x))
        probs /= np.sum(probs)
        return tuple(int(a) for a in acts), probs

    def search_vl(self, board, n_playout, k):
        """n_playout playouts with up to k leaves in flight per wave (orc_mcts_search_vl: the build's own virtual-loss
        specification, not a reference behaviour).  Returns (acts, visits, Q, P, waves)."""
        waves = C.c_int()
        rc = lib().orc_mcts_search_vl(self.h, C.addressof(board.c), int(n_playout), int(k), C.byref(waves))
        assert rc == 0
        return self.root_children() + (waves.value,)

    def replay(self, board, tlen, trace):
        """Decision-level replay of a recorded search (orc_mcts_replay): tlen uint8 [n_playout] = depth of
        every playout, trace uint8 = the chosen moves back to back.  Returns the statistics dict; raises
        AssertionError when a recorded path is not a root-to-leaf path of the oracle's tree."""
        tlen = np.ascontiguousarray(tlen, dtype=np.uint8)
        trace = np.ascontiguousarray(trace, dtype=np.uint8)
        assert int(tlen.sum()) == trace.size, (int(tlen.sum()), trace.size)
        st = CReplayStats()
        lib().orc_replay_stats_init(C.byref(st))
        tr = trace if trace.size else np.zeros(1, np.uint8)
        rc = lib().orc_mcts_replay(self.h, C.addressof(board.c), tlen.ctypes.data, tr.ctypes.data,
                                   int(tlen.size), C.byref(st))
        if rc != 0:
            raise AssertionError("recorded path of playout %d is not a root-to-leaf path of the oracle tree "
                                 "(code %d: 1 = runs past a leaf, 2 = move is not a child, 3 = stops above the leaf)"
                                 % ((-rc) // 8 - 1, (-rc) % 8))
        return {k: getattr(st, k) for k, _ in CReplayStats._fields_}

    def root_children(self):
        """(acts, visits, Q, P) of the root children without running playouts."""
        return self.search(None, n_playout=0)

    def update_with_move(self, move):
        lib().orc_mcts_update_with_move(self.h, int(move))

    def pure_get_move(self, board):
        return lib().orc_pure_get_move(self.h, C.addressof(board.c))

    def root_visits(self):
        return lib().orc_mcts_root_visits(self.h)

    def root_q(self):
        return lib().orc_mcts_root_q(self.h)

    def stat(self, which):
        return lib().orc_mcts_stat(self.h, which)


def hash_eval(board, salt=0):
    probs = np.zeros(MAXS, dtype=np.float32)
    v = C.c_float()
    lib().orc_hash_eval(C.addressof(board.c), C.c_uint64(salt), probs.ctypes.data, C.byref(v))
    return probs[:board.c.n_avail].copy(), v.value


def rollout_once(board, seed, game_id, search_id, playout_idx):
    return lib().orc_rollout_once(C.addressof(board.c), C.c_uint64(seed), game_id, search_id,
                                  playout_idx)


def philox(key0, key1, c0, c1, c2, c3):
    out = (C.c_uint32 * 4)()
    lib().orc_philox4x32(key0, key1, c0, c1, c2, c3, out)
    return list(out)
