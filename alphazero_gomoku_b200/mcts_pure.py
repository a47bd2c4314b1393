"""Pure MCTS with random rollouts: drop-in for the reference's mcts_pure.py (:124-273).

The tree, the uniform priors over Board.move_target (mcts_pure.py:33-46) and the rollouts
(_evaluate_rollout :193-212, Philox-driven on the device) all run in gz_pure_search; the
reference draws from the unseeded global numpy RNG, so parity of this path is distributional.
"""
import logging

import numpy as np

from .engine import Engine
from .mcts_alphaZero import check_pool, tree_pool_nodes

_logger = logging.getLogger(__name__)


def rollout_policy_fn(board):
    """mcts_pure.py:15-30: uniform random scores over the candidate box."""
    board.get_move_target()
    return zip(board.move_target, np.random.rand(len(board.move_target)))


def policy_value_fn(board):
    """mcts_pure.py:33-46: uniform priors over the candidate box, value 0."""
    board.get_move_target()
    action_probs = np.ones(len(board.move_target)) / len(board.move_target)
    return zip(board.move_target, action_probs), 0


class MCTS(object):
    """mcts_pure.py:124-246 over a one-slot device tree."""

    def __init__(self, policy_value_fn=policy_value_fn, c_puct=5, n_playout=1000, seed=None):
        if policy_value_fn is not globals()["policy_value_fn"]:
            # the device search hard-wires mcts_pure.policy_value_fn (uniform priors over move_target)
            raise ValueError("mcts_pure.MCTS runs on the device with mcts_pure.policy_value_fn only")
        self._policy = policy_value_fn
        self._c_puct = c_puct
        self._n_playout = int(n_playout)
        self._engine = None
        self._geom = None
        self._search_id = 0
        self._seed = int(np.random.randint(0, 2 ** 31 - 1)) if seed is None else int(seed)

    def _ensure_engine(self, board):
        geom = (board.width, board.height, board.n_in_row)
        if self._engine is None or geom != self._geom:
            if self._engine is not None:
                self._engine.close()
            self._engine = Engine(width=geom[0], height=geom[1], n_in_row=geom[2], n_games=1,
                                  n_playout=self._n_playout, c_puct=float(self._c_puct),
                                  nodes_per_game=tree_pool_nodes(geom[0] * geom[1], self._n_playout),
                                  seed=self._seed)
            self._overflows = 0
            self._geom = geom
        return self._engine

    def get_move(self, board):
        """:214-231: n_playout playouts, then the most visited root child (first maximum)."""
        e = self._ensure_engine(board)
        e.boards_set(0, list(board.states.keys()), getattr(board, "_start", 0))
        self._search_id += 1
        e.pure_search(self._n_playout, self._search_id)
        move = int(e.pure_best_move()[0])
        check_pool(self, e)
        return move

    def root_visits(self):
        vis, _ = self._engine.root_visits()
        return vis[0]

    def update_with_move(self, last_move):
        if self._engine is not None:
            self._engine.update_with_move([int(last_move)])

    def __str__(self):
        return "MCTS"


class MCTSPlayer(object):
    """AI player based on MCTS (mcts_pure.py:249-273)"""

    def __init__(self, c_puct=5, n_playout=2000, seed=None):
        self.mcts = MCTS(policy_value_fn, c_puct, n_playout, seed=seed)

    def set_player_ind(self, p):
        self.player = p

    def reset_player(self):
        self.mcts.update_with_move(-1)

    def get_action(self, board):
        sensible_moves = board.availables
        if len(sensible_moves) > 0:
            move = self.mcts.get_move(board)
            self.mcts.update_with_move(-1)
            return move
        else:
            print("WARNING: the board is full")

    def __str__(self):
        return "MCTS_Pure {}".format(self.player)
