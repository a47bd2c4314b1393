"""alphazero_gomoku_b200 — B200-native (sm_100a) self-play engine for AlphaZero Gomoku.

Drop-in for the hot path of AbcEric/AlphaZero_Gomoku: Board/Game, MCTS/MCTSPlayer (AlphaZero
and pure), PolicyValueNet and TrainPipeline.collect_selfplay_data, all thin wrappers over the
C ABI in include/gomoku_b200.h.  There is no CPU fallback.
"""
from .engine import Engine  # noqa: F401

__all__ = ["Engine"]
