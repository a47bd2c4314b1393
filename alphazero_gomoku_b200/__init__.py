"""alphazero_gomoku_b200 — B200-native (sm_100a) self-play engine for AlphaZero Gomoku.

Drop-in for the hot path of AbcEric/AlphaZero_Gomoku: Board/Game (game.py), MCTS/MCTSPlayer
(mcts_alphaZero.py, mcts_pure.py), PolicyValueNet (policy_value_net.py) and
TrainPipeline.collect_selfplay_data (train.py), all thin wrappers over the C ABI declared in
include/gomoku_b200.h.  There is no CPU fallback: importing works anywhere the library has been
built, creating an engine needs a CUDA device.
"""
from .engine import Engine  # noqa: F401
from .game import Board, Game  # noqa: F401
from .mcts_alphaZero import MCTS, MCTSPlayer  # noqa: F401
from .mcts_pure import MCTSPlayer as MCTS_Pure  # noqa: F401
from .policy_value_net import PolicyValueNet  # noqa: F401
from .train import TrainPipeline  # noqa: F401

__all__ = ["Engine", "Board", "Game", "MCTS", "MCTSPlayer", "MCTS_Pure", "PolicyValueNet", "TrainPipeline"]
