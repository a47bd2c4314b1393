"""MCTS / MCTSPlayer: drop-in for the reference's mcts_alphaZero.py (:95-265).

The tree lives in HBM inside an `Engine` (gz_tree.cu); this class keeps the reference's call
protocol.  Two evaluator modes, chosen from what `policy_value_fn` is:
  * a bound `PolicyValueNet.policy_value_fn` of this package -> the whole search runs on the
    device (gz_search: select -> tcgen05 net -> expand/backup per playout);
  * any other callable honouring the reference protocol (mcts_alphaZero.py:98-107) -> the
    search is split around it (gz_search_select / gz_search_expand_backup) and the callable is
    handed a leaf `Board` exactly as the reference's `_playout` does (:132).
"""
import logging
import random

import numpy as np

from .engine import Engine
from .game import Board

_logger = logging.getLogger(__name__)


def softmax(x):
    probs = np.exp(x - np.max(x))
    probs /= np.sum(probs)
    return probs


def _net_of(fn):
    """The PolicyValueNet behind a bound policy_value_fn, else None."""
    from .policy_value_net import PolicyValueNet
    owner = getattr(fn, "__self__", None)
    if isinstance(owner, PolicyValueNet) and getattr(fn, "__name__", "") == "policy_value_fn":
        return owner
    return None


def tree_pool_nodes(S, n_playout, budget_bytes=2 << 30):
    """Node blocks for a one-game tree that is re-rooted with subtree reuse (update_with_move(move)): a
    search adds at most n_playout nodes and a game has at most S searches, so S * n_playout can never
    overflow; that bound is taken while it fits `budget_bytes` of HBM (20 * Sp bytes per block), else the
    budget, never less than the engine default 8 * n_playout + 8 (measured high-water mark 6.9 * n_playout).
    A search that does run out raises (check_pool) instead of silently leaving leaves unexpanded."""
    sp = (S + 31) // 32 * 32
    return int(max(8 * n_playout + 8, min(S * n_playout + 8, budget_bytes // (20 * sp))))


def check_pool(owner, engine):
    """Raise if the search just run dropped an expansion because the tree pool was full (the reference
    never drops one: mcts_alphaZero.py:137 always expands a non-terminal leaf)."""
    n = engine.selfplay_stats()["pool_overflows"]
    if n > getattr(owner, "_overflows", 0):
        owner._overflows = n
        raise MemoryError("gomoku_b200: the device tree pool of this search is exhausted (%d expansions dropped); "
                          "create the player with a smaller n_playout or reset it with update_with_move(-1)" % n)


class MCTS(object):
    """mcts_alphaZero.py:95-192 over a one-slot device tree."""

    def __init__(self, policy_value_fn, c_puct=5, n_playout=10000, leaves_in_flight=1):
        """leaves_in_flight = K > 1 (not in the reference): up to K leaves per kernel chain with virtual loss
        (gz_search_vl) - about 3x lower latency for a single game at K = 8, but no longer the reference's sequential
        search (visit distributions ~0.04 total variation away at 400 playouts).  Default 1 = the reference."""
        self.leaves_in_flight = int(leaves_in_flight)
        self._policy = policy_value_fn
        self._c_puct = c_puct
        self._n_playout = int(n_playout)
        self._net = _net_of(policy_value_fn)
        self._engine = None
        self._geom = None
        self._weights_version = -1

    # ---- engine management -----------------------------------------------------------------
    def _ensure_engine(self, board):
        geom = (board.width, board.height, board.n_in_row)
        if self._engine is None or geom != self._geom:
            if self._engine is not None:
                self._engine.close()
            self._engine = Engine(width=geom[0], height=geom[1], n_in_row=geom[2], n_games=1,
                                  n_playout=self._n_playout, c_puct=float(self._c_puct),
                                  nodes_per_game=tree_pool_nodes(geom[0] * geom[1], self._n_playout))
            self._overflows = 0
            self._geom = geom
            self._weights_version = -1
        if self._net is not None and self._weights_version != self._net.version:
            self._engine.set_weights(self._net.get_policy_param())
            self._weights_version = self._net.version
        return self._engine

    def _sync_board(self, e, state):
        e.boards_set(0, list(state.states.keys()), getattr(state, "_start", 0))

    # ---- reference API -----------------------------------------------------------------------
    def get_move_probs(self, state, temp=1e-3):
        """:154-181: n_playout playouts from `state`, then visit counts -> softmax(log(v)/temp)."""
        e = self._ensure_engine(state)
        self._sync_board(e, state)
        if self._net is not None:
            e.search(self._n_playout, leaves_in_flight=self.leaves_in_flight)
        else:
            for _ in range(self._n_playout):
                self._playout_external(e, state)
        check_pool(self, e)
        vis, _ = e.root_visits()
        acts = np.nonzero(vis[0] >= 0)[0]
        visits = vis[0][acts].astype(np.float64)
        act_probs = softmax(1.0 / temp * np.log(visits + 1e-10))
        return tuple(int(a) for a in acts), act_probs

    def _playout_external(self, e, state):
        """One playout with a caller-supplied evaluator (:113-148)."""
        st, planes, cells, player = e.search_select(want_cells=True)
        S = state.width * state.height
        priors = np.zeros((1, S), np.float32)
        values = np.zeros(1, np.float32)
        if st[0] == 0:
            leaf = _board_from_cells(state, cells[0], int(player[0]), planes[0])
            action_probs, leaf_value = self._policy(leaf)
            for a, p in action_probs:
                priors[0, int(a)] = p
            values[0] = float(leaf_value)
        e.search_expand_backup(priors, values)

    def update_with_move(self, last_move):
        """:183-192: re-root at a child (subtree kept) or reset."""
        if self._engine is not None:
            self._engine.update_with_move([int(last_move)])

    def __str__(self):
        return "MCTS"


def _board_from_cells(template, cells, player, planes=None):
    """A Board holding the leaf position.  The move order inside each colour is not recoverable
    (nor meaningful) at a leaf, except for the last move, which plane 2 of the leaf state names."""
    W = template.width
    b = Board(width=W, height=template.height, n_in_row=template.n_in_row)
    ones = [int(m) for m in np.nonzero(cells == 1)[0]]
    twos = [int(m) for m in np.nonzero(cells == 2)[0]]
    start = 0 if len(ones) > len(twos) else 1 if len(twos) > len(ones) else player - 1
    first, second = (ones, twos) if start == 0 else (twos, ones)
    mover = second if len(first) == len(second) else first  # colour of the stone played last
    if planes is not None and mover:
        rs, c = np.nonzero(planes[2])
        if len(rs) == 1:
            last = int((W - 1 - rs[0]) * W + c[0])  # undo the row reversal of current_state
            if last in mover:
                mover.remove(last)
                mover.append(last)
    b.init_board(start)
    for i in range(len(first)):
        b.do_move(first[i])
        if i < len(second):
            b.do_move(second[i])
    return b


class MCTSPlayer(object):
    """AI player based on MCTS (mcts_alphaZero.py:198-265)"""

    def __init__(self, policy_value_function, c_puct=5, n_playout=2000, is_selfplay=0, leaves_in_flight=1):
        self.mcts = MCTS(policy_value_function, c_puct, n_playout, leaves_in_flight=leaves_in_flight)
        self._is_selfplay = is_selfplay

    def set_player_ind(self, p):
        self.player = p

    def reset_player(self):
        self.mcts.update_with_move(-1)

    def get_action(self, board, temp=1e-3, return_prob=0):
        sensible_moves = board.availables
        move_probs = np.zeros(board.width * board.height)
        if len(sensible_moves) > 0:
            acts, probs = self.mcts.get_move_probs(board, temp)
            move_probs[list(acts)] = probs
            if self._is_selfplay:
                # Dirichlet noise is mixed into the sampling distribution only (:236-240)
                move = np.random.choice(acts, p=0.75 * probs + 0.25 * np.random.dirichlet(0.1 * np.ones(len(probs))))
                self.mcts.update_with_move(move)
            else:
                move = np.random.choice(acts, p=probs)
                if len(board.states) == 0:
                    # the reference hard-codes the 15x15 centre (:251-253), illegal on smaller
                    # boards; reproduce on 15x15, keep the sampled move elsewhere (SURVEY A-13)
                    if board.width == 15 and board.height == 15:
                        move = random.choice([96, 97, 98, 111, 112, 113, 126, 127, 128])
                        print("move=", move)
                self.mcts.update_with_move(-1)
            move = int(move)
            if return_prob:
                return move, move_probs
            return move
        else:
            _logger.info("WARNING: the board is full")

    def __str__(self):
        return "MCTS {}".format(self.player)
