"""Engine: the batched, GPU-resident self-play engine (one per GPU), a thin ctypes wrapper.

Every method maps 1:1 onto a C-ABI entry point of include/gomoku_b200.h.  Array arguments
may be numpy arrays (host) or torch CUDA tensors (device, zero-copy).  The reference classes
(Board, MCTS, MCTSPlayer, PolicyValueNet, TrainPipeline) in this package are built on it.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import GzConfig, GzCounters, GzSelfplayStats, check

STATE_DICT_KEYS = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "conv3.weight",
                   "conv3.bias", "act_conv1.weight", "act_conv1.bias", "act_fc1.weight",
                   "act_fc1.bias", "val_conv1.weight", "val_conv1.bias", "val_fc1.weight",
                   "val_fc1.bias", "val_fc2.weight", "val_fc2.bias"]


def _ptr(x):
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"], "array must be C-contiguous"
        return x.ctypes.data
    if hasattr(x, "data_ptr"):  # torch tensor
        assert x.is_contiguous()
        return x.data_ptr()
    raise TypeError("expected numpy array or torch tensor, got %r" % type(x))


def weight_shapes(width, height, trunk=(32, 64, 128)):
    S = width * height
    c1, c2, c3 = trunk
    return [(c1, 4, 3, 3), (c1,), (c2, c1, 3, 3), (c2,), (c3, c2, 3, 3), (c3,), (4, c3, 1, 1), (4,),
            (S, 4 * S), (S,), (2, c3, 1, 1), (2,), (64, 2 * S), (64,), (1, 64), (1,)]


class Engine:
    def __init__(self, width=8, height=8, n_in_row=5, n_games=1, n_playout=400, c_puct=5.0,
                 nodes_per_game=0, dirichlet_alpha=0.1, dirichlet_eps=0.25, trunk=(32, 64, 128),
                 seed=0, game_id_base=0, device=0, net_impl=0, stream=None):
        self._h = None
        L = _lib.lib()
        cfg = GzConfig()
        cfg.width, cfg.height, cfg.n_in_row = int(width), int(height), int(n_in_row)
        cfg.n_games, cfg.n_playout, cfg.nodes_per_game = int(n_games), int(n_playout), int(nodes_per_game)
        self.nodes_per_game = int(nodes_per_game) if nodes_per_game > 0 else 8 * int(n_playout) + 8
        cfg.c_puct, cfg.dirichlet_alpha, cfg.dirichlet_eps = float(c_puct), float(dirichlet_alpha), float(dirichlet_eps)
        cfg.trunk[0], cfg.trunk[1], cfg.trunk[2] = [int(t) for t in trunk]
        cfg.seed, cfg.game_id_base = int(seed), int(game_id_base)
        cfg.device, cfg.net_impl = int(device), int(net_impl)
        cfg.stream = stream
        h = C.c_void_p()
        rc = L.gz_create(C.byref(cfg), C.byref(h))
        check(None, rc)
        self._h = h
        self.L = L
        self.width, self.height, self.n_in_row = cfg.width, cfg.height, cfg.n_in_row
        self.S = self.width * self.height
        self.G = cfg.n_games
        self.n_playout = cfg.n_playout
        self.trunk = tuple(int(t) for t in trunk)
        self.device = cfg.device
        self.game_id_base = int(game_id_base)

    def close(self):
        if getattr(self, "_h", None):
            self.L.gz_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        check(self._h, rc)

    def sync(self):
        self._ck(self.L.gz_sync(self._h))

    # ---- weights ---------------------------------------------------------------------------
    def set_weights(self, arrays, layout=_lib.LAYOUT_TORCH):
        """arrays: 16 float32 arrays in state_dict order (numpy or torch, host or device)."""
        keep = []
        ptrs = (C.c_void_p * 16)()
        shapes = weight_shapes(self.width, self.height, self.trunk)
        for i, a in enumerate(arrays):
            if isinstance(a, np.ndarray):
                a = np.ascontiguousarray(a, dtype=np.float32)
            else:
                a = a.detach().float().contiguous()
            n = int(np.prod(a.shape))
            if n != int(np.prod(shapes[i])):
                raise Exception("weight %d has %d elements, expected %s" % (i, n, shapes[i]))
            keep.append(a)
            ptrs[i] = _ptr(a)
        self._ck(self.L.gz_set_weights(self._h, ptrs, layout))

    def get_weights(self):
        shapes = weight_shapes(self.width, self.height, self.trunk)
        arrs = [np.zeros(s, dtype=np.float32) for s in shapes]
        ptrs = (C.c_void_p * 16)(*[a.ctypes.data for a in arrs])
        self._ck(self.L.gz_get_weights(self._h, ptrs))
        return arrs

    # ---- stateless rules -------------------------------------------------------------------
    def rules_eval(self, move_lists, start_players=None, want=("status", "ended", "winner", "has_win",
                                                               "legal", "planes", "target", "cur_player")):
        n = len(move_lists)
        offsets = np.zeros(n + 1, dtype=np.int64)
        offsets[1:] = np.cumsum([len(m) for m in move_lists])
        flat = np.zeros(max(int(offsets[-1]), 1), dtype=np.int32)
        if offsets[-1]:
            flat[:offsets[-1]] = np.concatenate([np.asarray(m, dtype=np.int32) for m in move_lists if len(m)])
        sp = None if start_players is None else np.ascontiguousarray(start_players, dtype=np.int8)
        S = self.S
        out = {}
        if "status" in want: out["status"] = np.zeros(n, np.int32)
        if "ended" in want: out["ended"] = np.zeros(n, np.int8)
        if "winner" in want: out["winner"] = np.zeros(n, np.int8)
        if "has_win" in want: out["has_win"] = np.zeros(n, np.int8)
        if "legal" in want: out["legal"] = np.zeros((n, S), np.uint8)
        if "planes" in want: out["planes"] = np.zeros((n, 4, self.width, self.height), np.float32)
        if "target" in want: out["target"] = np.zeros((n, S), np.uint8)
        if "cur_player" in want: out["cur_player"] = np.zeros(n, np.int8)
        g = lambda k: _ptr(out.get(k))
        self._ck(self.L.gz_rules_eval(self._h, _ptr(flat), _ptr(offsets), n, _ptr(sp), g("status"),
                                      g("ended"), g("winner"), g("has_win"), g("legal"), g("planes"),
                                      g("target"), g("cur_player")))
        return out

    # ---- resident boards -------------------------------------------------------------------
    def boards_reset(self, start_players=None):
        sp = None if start_players is None else np.ascontiguousarray(start_players, dtype=np.int8)
        self._ck(self.L.gz_boards_reset(self._h, _ptr(sp)))

    def boards_set(self, game, moves, start_player=0):
        mv = np.ascontiguousarray(moves, dtype=np.int32)
        self._ck(self.L.gz_boards_set(self._h, int(game), _ptr(mv) if len(mv) else None, len(mv),
                                      int(start_player)))

    def boards_do_move(self, moves, return_status=False):
        mv = moves if hasattr(moves, "data_ptr") else np.ascontiguousarray(moves, dtype=np.int32)
        st = np.zeros(self.G, np.int32)
        self._ck(self.L.gz_boards_do_move(self._h, _ptr(mv), _ptr(st)))
        if return_status:
            return st
        if (st != 0).any():
            raise ValueError("list.remove(x): x not in list (illegal move in game %d)"
                             % int(np.nonzero(st)[0][0]))

    def boards_status(self):
        G = self.G
        ended, winner, cur = np.zeros(G, np.int8), np.zeros(G, np.int8), np.zeros(G, np.int8)
        ns, last = np.zeros(G, np.int32), np.zeros(G, np.int32)
        self._ck(self.L.gz_boards_status(self._h, _ptr(ended), _ptr(winner), _ptr(cur), _ptr(ns), _ptr(last)))
        return {"ended": ended, "winner": winner, "cur_player": cur, "n_stones": ns, "last_move": last}

    def boards_legal_mask(self):
        out = np.zeros((self.G, self.S), np.uint8)
        self._ck(self.L.gz_boards_legal_mask(self._h, _ptr(out)))
        return out

    def boards_planes(self, out=None):
        if out is None:
            out = np.zeros((self.G, 4, self.width, self.height), np.float32)
        self._ck(self.L.gz_boards_planes(self._h, _ptr(out)))
        return out

    def boards_cells(self):
        out = np.zeros((self.G, self.S), np.int8)
        self._ck(self.L.gz_boards_cells(self._h, _ptr(out)))
        return out

    # ---- net -------------------------------------------------------------------------------
    def net_forward(self, planes, probs=None, values=None):
        """planes [B,4,W,H] float32 (numpy or torch cuda) -> (probs [B,S], values [B])."""
        if isinstance(planes, np.ndarray):
            planes = np.ascontiguousarray(planes, dtype=np.float32)
            B = planes.shape[0]
            probs = np.zeros((B, self.S), np.float32) if probs is None else probs
            values = np.zeros((B,), np.float32) if values is None else values
        else:
            import torch
            planes = planes.float().contiguous()
            B = planes.shape[0]
            probs = torch.empty((B, self.S), dtype=torch.float32, device=planes.device) if probs is None else probs
            values = torch.empty((B,), dtype=torch.float32, device=planes.device) if values is None else values
            torch.cuda.current_stream(planes.device).synchronize()
        self._ck(self.L.gz_net_forward(self._h, _ptr(planes), B, _ptr(probs), _ptr(values)))
        return probs, values

    def net_forward_noalloc(self, planes, probs, values):
        """gz_net_forward_noalloc: planes [B,4,W,H], probs [B,S], values [B] are caller-owned float32 CUDA tensors;
        nothing is allocated (static 8 MiB device scratch, B limited accordingly) and nothing is synchronised:
        the work is queued on the engine stream (sync() before reading)."""
        for t in (planes, probs, values):
            if not hasattr(t, "data_ptr"):
                raise TypeError("net_forward_noalloc takes torch CUDA tensors (device pointers)")
        self._ck(self.L.gz_net_forward_noalloc(self._h, _ptr(planes), int(planes.shape[0]), _ptr(probs), _ptr(values)))

    def debug_hash_evaluator(self, on=True, salt=0):
        """diagnostic: the oracle's integer-hash evaluator replaces the net inside the fused searches"""
        self._ck(self.L.gz_debug_hash_evaluator(self._h, int(bool(on)), int(salt)))

    def debug_net_layers(self, planes):
        """tensor-core path with per-layer activations: dict(act1 [B,S,c1], act2 [B,S,c2],
        act3 [B,S,c3], head [B,6,S], probs [B,S], values [B])."""
        planes = np.ascontiguousarray(planes, dtype=np.float32)
        B, S = planes.shape[0], self.S
        c1, c2, c3 = self.trunk
        out = {"act1": np.zeros((B, S, c1), np.float32), "act2": np.zeros((B, S, c2), np.float32),
               "act3": np.zeros((B, S, c3), np.float32), "head": np.zeros((B, 6, S), np.float32),
               "probs": np.zeros((B, S), np.float32), "values": np.zeros(B, np.float32)}
        self._ck(self.L.gz_debug_net_layers(self._h, _ptr(planes), B, _ptr(out["act1"]), _ptr(out["act2"]),
                                            _ptr(out["act3"]), _ptr(out["head"]), _ptr(out["probs"]),
                                            _ptr(out["values"])))
        return out

    # ---- search ----------------------------------------------------------------------------
    def search(self, n_playout=None, leaves_in_flight=1):
        """n_playout playouts in every active game.  leaves_in_flight = K > 1: up to K leaves of a game per wave
        (virtual loss, gz_search_vl) - faster for few games, not the reference's sequential search; returns the
        number of kernel chains launched."""
        n = int(self.n_playout if n_playout is None else n_playout)
        if int(leaves_in_flight) <= 1:
            self._ck(self.L.gz_search(self._h, n))
            return n
        waves = C.c_int32()
        self._ck(self.L.gz_search_vl(self._h, n, int(leaves_in_flight), C.byref(waves)))
        return int(waves.value)

    def search_select(self, want_cells=True):
        G, S = self.G, self.S
        st = np.zeros(G, np.int8)
        planes = np.zeros((G, 4, self.width, self.height), np.float32)
        cells = np.zeros((G, S), np.int8) if want_cells else None
        player = np.zeros(G, np.int8)
        self._ck(self.L.gz_search_select(self._h, _ptr(st), _ptr(planes), _ptr(cells), _ptr(player)))
        return st, planes, cells, player

    def search_expand_backup(self, priors, values):
        pr = priors if hasattr(priors, "data_ptr") else np.ascontiguousarray(priors, dtype=np.float32)
        va = values if hasattr(values, "data_ptr") else np.ascontiguousarray(values, dtype=np.float32)
        self._ck(self.L.gz_search_expand_backup(self._h, _ptr(pr), _ptr(va)))

    def search_paths(self):
        """(path_moves [G,S] with -1 padding, path_len [G]) of the last search_select descent."""
        mv = np.zeros((self.G, self.S), np.int32)
        ln = np.zeros(self.G, np.int32)
        self._ck(self.L.gz_search_paths(self._h, _ptr(mv), _ptr(ln)))
        return mv, ln

    def root_visits(self, want_q=False):
        G, S = self.G, self.S
        vis = np.zeros((G, S), np.int32)
        q = np.zeros((G, S), np.float64) if want_q else None
        rn = np.zeros(G, np.int32)
        self._ck(self.L.gz_root_visits(self._h, _ptr(vis), _ptr(q), _ptr(rn)))
        return (vis, q, rn) if want_q else (vis, rn)

    def update_with_move(self, moves):
        mv = np.ascontiguousarray(moves, dtype=np.int32)
        self._ck(self.L.gz_update_with_move(self._h, _ptr(mv)))

    def equi_data(self, states, pis, zs):
        """The 8 dihedral images of every sample, reference order (train.py:116-135); numpy float32 in
        and out, or torch cuda tensors in and out (no host round trip)."""
        if hasattr(states, "data_ptr"):
            import torch
            n = states.shape[0]
            st, pi, z = states.float().contiguous(), pis.float().contiguous(), zs.float().contiguous()
            torch.cuda.current_stream(st.device).synchronize()
            os_ = torch.empty((8 * n, 4, self.width, self.height), dtype=torch.float32, device=st.device)
            op = torch.empty((8 * n, self.S), dtype=torch.float32, device=st.device)
            oz = torch.empty((8 * n,), dtype=torch.float32, device=st.device)
        else:
            st = np.ascontiguousarray(states, dtype=np.float32)
            pi = np.ascontiguousarray(pis, dtype=np.float32)
            z = np.ascontiguousarray(zs, dtype=np.float32)
            n = st.shape[0]
            os_ = np.zeros((8 * n, 4, self.width, self.height), np.float32)
            op = np.zeros((8 * n, self.S), np.float32)
            oz = np.zeros((8 * n,), np.float32)
        self._ck(self.L.gz_equi_data(self._h, _ptr(st), _ptr(pi), _ptr(z), int(n), _ptr(os_), _ptr(op), _ptr(oz)))
        return os_, op, oz

    # ---- train step on the device -------------------------------------------------
    def _train_args(self, states, pis, zs):
        if hasattr(states, "data_ptr"):
            import torch
            st, pi, z = states.float().contiguous(), pis.float().contiguous(), zs.float().contiguous()
            torch.cuda.current_stream(st.device).synchronize()  # the engine runs on its own stream
        else:
            st = np.ascontiguousarray(states, dtype=np.float32)
            pi = np.ascontiguousarray(pis, dtype=np.float32)
            z = np.ascontiguousarray(zs, dtype=np.float32)
        return st, pi, z, int(st.shape[0])

    def train_step(self, states, pis, zs, lr):
        """PolicyValueNet.train_step (policy_value_net_pytorch.py:117-148) -> (loss, entropy)."""
        st, pi, z, n = self._train_args(states, pis, zs)
        loss, ent = C.c_float(0), C.c_float(0)
        self._ck(self.L.gz_train_step(self._h, _ptr(st), _ptr(pi), _ptr(z), n, float(lr), C.byref(loss), C.byref(ent)))
        return loss.value, ent.value

    def train_backward(self, states, pis, zs):
        """Forward + backward only; the flat gradient stays in the engine (see train_grads)."""
        st, pi, z, n = self._train_args(states, pis, zs)
        loss, ent = C.c_float(0), C.c_float(0)
        self._ck(self.L.gz_train_backward(self._h, _ptr(st), _ptr(pi), _ptr(z), n, C.byref(loss), C.byref(ent)))
        return loss.value, ent.value

    def train_param_count(self):
        n = C.c_int64(0)
        self._ck(self.L.gz_train_grads(self._h, None, 0, C.byref(n)))
        return int(n.value)

    def train_grads(self, out=None):
        """Flat gradient of the last train_backward in state_dict order / torch layout; `out` may be a
        torch cuda tensor (filled in place, e.g. for an NCCL all-reduce)."""
        n = self.train_param_count()
        if out is None:
            out = np.zeros(n, np.float32)
        self._ck(self.L.gz_train_grads(self._h, _ptr(out), n, None))
        return out

    def train_apply(self, lr, grads=None):
        """Adam step with `grads` (flat, host or device) or the engine's own gradient."""
        if grads is not None and hasattr(grads, "data_ptr"):
            import torch
            torch.cuda.current_stream(grads.device).synchronize()  # e.g. an NCCL all-reduce still in flight
        elif grads is not None:
            grads = np.ascontiguousarray(grads, dtype=np.float32)
        self._ck(self.L.gz_train_apply(self._h, _ptr(grads) if grads is not None else None, float(lr)))

    def train_forward(self, states):
        """policy_value with the weights as trained so far (fp32 path) -> (probs [n,S], values [n])."""
        if hasattr(states, "data_ptr"):
            import torch
            st = states.float().contiguous()
            torch.cuda.current_stream(st.device).synchronize()
        else:
            st = np.ascontiguousarray(states, dtype=np.float32)
        n = int(st.shape[0])
        probs = np.zeros((n, self.S), np.float32)
        values = np.zeros(n, np.float32)
        self._ck(self.L.gz_train_forward(self._h, _ptr(st), n, _ptr(probs), _ptr(values)))
        return probs, values

    def train_reset(self):
        self._ck(self.L.gz_train_reset(self._h))

    def set_active(self, mask=None):
        """Slots that take part in the next searches (None = all)."""
        if mask is None:
            self._ck(self.L.gz_set_active(self._h, None))
        else:
            m = np.ascontiguousarray(np.asarray(mask) != 0, dtype=np.int8)
            self._ck(self.L.gz_set_active(self._h, _ptr(m)))

    def root_select_move(self, temp=1e-3, selfplay=False):
        moves = np.zeros(self.G, np.int32)
        probs = np.zeros((self.G, self.S), np.float32)
        self._ck(self.L.gz_root_select_move(self._h, float(temp), int(bool(selfplay)), _ptr(moves), _ptr(probs)))
        return moves, probs

    # ---- self-play -------------------------------------------------------------------------
    def selfplay_begin(self, temp=1.0, recycle=False):
        """recycle: False = a slot idles after its game, True / 1 = next game id = id + G, 2 = next id = next start
        serial (ids order the games by when they began; TrainPipeline's continuous mode)."""
        self._ck(self.L.gz_selfplay_begin(self._h, float(temp), 2 if recycle == 2 else int(bool(recycle))))

    def selfplay_step(self, n_plies=1):
        self._ck(self.L.gz_selfplay_step(self._h, int(n_plies)))

    def selfplay_pending(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.L.gz_selfplay_pending(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def selfplay_drain(self, max_samples=None, bufs=None):
        """Returns dict(states [n,4,W,H], pis [n,S], zs [n], game_ids [n], plies [n])."""
        if max_samples is None:
            max_samples = self.selfplay_pending()[0]
        n_out = C.c_int64()
        if bufs is None:
            m = max(int(max_samples), 1)
            bufs = {"states": np.zeros((m, 4, self.width, self.height), np.float32),
                    "pis": np.zeros((m, self.S), np.float32), "zs": np.zeros(m, np.float32),
                    "game_ids": np.zeros(m, np.int64), "plies": np.zeros(m, np.int32)}
        self._ck(self.L.gz_selfplay_drain(self._h, int(max_samples), _ptr(bufs["states"]), _ptr(bufs["pis"]),
                                          _ptr(bufs["zs"]), _ptr(bufs["game_ids"]), _ptr(bufs["plies"]),
                                          C.byref(n_out)))
        n = n_out.value
        return {k: v[:n] for k, v in bufs.items()}

    def selfplay_drain_games(self, max_games=None):
        if max_games is None:
            max_games = 4 * self.G
        moves = np.zeros(max(max_games * self.S, 1), np.int32)
        offsets = np.zeros(max_games + 1, np.int64)
        winners = np.zeros(max(max_games, 1), np.int8)
        gids = np.zeros(max(max_games, 1), np.int64)
        n = C.c_int64()
        self._ck(self.L.gz_selfplay_drain_games(self._h, int(max_games), int(moves.size), _ptr(moves),
                                                _ptr(offsets), _ptr(winners), _ptr(gids), C.byref(n)))
        n = n.value
        return [(int(gids[i]), int(winners[i]), moves[offsets[i]:offsets[i + 1]].tolist()) for i in range(n)]

    def selfplay_game_ids(self):
        """ids of the games in progress, one per slot"""
        ids = np.zeros(self.G, np.int64)
        self._ck(self.L.gz_selfplay_state(self._h, _ptr(ids), None))
        return ids

    def selfplay_first_serial(self):
        """smallest game serial (id - game_id_base) in flight: right after selfplay_begin, the first of the run"""
        return int(self.selfplay_game_ids().min()) - self.game_id_base

    def selfplay_dropped(self):
        """samples lost because the device sample ring was full (0 unless the caller stopped draining)"""
        d = C.c_int64()
        self._ck(self.L.gz_selfplay_state(self._h, None, C.byref(d)))
        return int(d.value)

    def selfplay_last(self, moves=None, probs=None):
        moves = np.zeros(self.G, np.int32) if moves is None else moves
        probs = np.zeros((self.G, self.S), np.float32) if probs is None else probs
        self._ck(self.L.gz_selfplay_last(self._h, _ptr(moves), _ptr(probs)))
        return moves, probs

    def selfplay_stats(self, reset=False):
        s = GzSelfplayStats()
        self._ck(self.L.gz_selfplay_get_stats(self._h, C.byref(s), int(reset)))
        return {"n_games": s.n_games, "n_samples": s.n_samples, "n_playouts": s.n_playouts,
                "n_leaf_evals": s.n_leaf_evals, "n_plies": s.n_plies, "pool_overflows": s.pool_overflows,
                "n_wins": [s.n_wins[0], s.n_wins[1], s.n_wins[2]], "max_tree_nodes": s.max_tree_nodes}

    # ---- pure MCTS -------------------------------------------------------------------------
    def pure_search(self, n_playout, search_id=0):
        self._ck(self.L.gz_pure_search(self._h, int(n_playout), int(search_id)))

    def pure_best_move(self):
        moves = np.zeros(self.G, np.int32)
        self._ck(self.L.gz_pure_best_move(self._h, _ptr(moves)))
        return moves

    def rollouts(self, n_rollouts, search_id=0):
        res = np.zeros((self.G, n_rollouts), np.int8)
        tp = C.c_int64()
        self._ck(self.L.gz_rollouts(self._h, int(n_rollouts), int(search_id), _ptr(res), C.byref(tp)))
        return res, tp.value

    def counters(self, reset=False):
        c = GzCounters()
        self._ck(self.L.gz_get_counters(self._h, C.byref(c), int(reset)))
        return {k: getattr(c, k) for k, _ in GzCounters._fields_}

    # ---- per-kernel-class device timing ------------------------------------------------------
    def profile_enable(self, on=True, stride=1):
        """Per-kernel-class CUDA-event timing; stride N times every N-th launch of a class and
        scales (the event pairs themselves cost GPU idle time between dependent kernels)."""
        self._ck(self.L.gz_profile_enable(self._h, max(1, int(stride)) if on else 0))

    def profile_read(self, reset=True):
        """{class name: (device ms, timed launches)} since the last reset."""
        n = _lib.PROF_CLASSES
        ms = np.zeros(n, np.float64)
        cnt = np.zeros(n, np.int64)
        self._ck(self.L.gz_profile_read(self._h, _ptr(ms), _ptr(cnt), int(reset)))
        return {self.L.gz_profile_class_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}
