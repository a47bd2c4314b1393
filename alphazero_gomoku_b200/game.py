"""Board / Game: drop-in for the reference's game.py (Board game.py:11-172, Game :175-416).

The Board keeps the reference's host-visible attributes (`states`, `availables`, `last_move`,
`current_player`, ...) because MCTS players, nets and GUIs read them directly, but every rule
*computation* (has_a_winner, game_end, current_state, get_move_target) is a call into the CUDA
library through the C ABI (gz_rules_eval): there is no CPU implementation of the rules in this
package.  The batched hot path never goes through this class; it exists so that code written
against the reference keeps working unchanged.
"""
from __future__ import print_function

import logging

import numpy as np

_logger = logging.getLogger(__name__)
_RULES = {}


def rules_engine(width, height, n_in_row):
    """Process-wide 1-slot engine used for single-board rule queries."""
    key = (int(width), int(height), int(n_in_row))
    if key not in _RULES:
        from .engine import Engine
        _RULES[key] = Engine(width=key[0], height=key[1], n_in_row=key[2], n_games=1, n_playout=1,
                             nodes_per_game=4)
    return _RULES[key]


class Board(object):
    """board for the game (game.py:11-172)"""

    def __init__(self, **kwargs):
        self.width = int(kwargs.get('width', 8))
        self.height = int(kwargs.get('height', 8))
        self.states = {}
        self.n_in_row = int(kwargs.get('n_in_row', 5))
        self.players = [1, 2]
        self.move_target = []
        self._start = 0

    # ---- helpers ---------------------------------------------------------------------------
    def _moves(self):
        return list(self.states.keys())  # dict insertion order == move order

    def _eval(self, want):
        if self.width < self.n_in_row or self.height < self.n_in_row:
            raise Exception('board width and height can not be less than ', self.n_in_row)
        e = rules_engine(self.width, self.height, self.n_in_row)
        return e.rules_eval([self._moves()], [self._start], want=want)

    # ---- reference API -----------------------------------------------------------------------
    def init_board(self, start_player=0):
        if self.width < self.n_in_row or self.height < self.n_in_row:
            raise Exception('board width and height can not be less than ', self.n_in_row)
        self.current_player = self.players[start_player]
        self._start = int(start_player)
        self.availables = list(range(self.width * self.height))
        self.states = {}
        self.last_move = -1

    def move_to_location(self, move):
        return [move // self.width, move % self.width]

    def location_to_move(self, location):
        if len(location) != 2:
            return -1
        move = location[0] * self.width + location[1]
        if move not in range(self.width * self.height):
            return -1
        return move

    def move_in_rect(self, move, rect):
        row, col = int(move / self.width), move % self.width
        return rect[1] <= row <= rect[0] and rect[2] <= col <= rect[3]

    def get_move_target(self):
        """game.py:35-58, computed on the device (bounding box of the stones grown by 2)."""
        t = self._eval(("target",))["target"][0]
        self.move_target = [int(m) for m in np.nonzero(t)[0]]

    def current_state(self):
        """game.py:92-116: float64 [4, W, H], rows reversed (the array the reference returns)."""
        if self.width != self.height:
            raise Exception("current_state needs a square board (game.py:97-110)")
        return self._eval(("planes",))["planes"][0].astype(np.float64)

    def do_move(self, move):
        self.availables.remove(move)  # ValueError on an occupied / unknown cell, as the reference
        self.states[move] = self.current_player
        self.current_player = self.players[0] if self.current_player == self.players[1] else self.players[1]
        self.last_move = move

    def has_a_winner(self):
        out = self._eval(("has_win", "winner", "ended"))
        if out["has_win"][0]:
            return True, int(out["winner"][0])
        return False, -1

    def game_end(self):
        out = self._eval(("ended", "winner"))
        if out["ended"][0]:
            return True, int(out["winner"][0])
        return False, -1

    def get_current_player(self):
        return self.current_player


class Game(object):
    """Match / self-play driver with the reference's entry points (game.py:175-416).  Everything
    that touches the rules goes through the Board above, i.e. through the device library."""

    def __init__(self, board, **kwargs):
        self.board = board

    def graphic(self, board, player1, player2):
        """Text rendering in the reference's format (game.py:183-209): X / O / - grid, top row first."""
        marks = {player1: 'X', player2: 'O'}
        out = ["Player %s %s" % (player1, "with X".rjust(3)), "Player %s %s" % (player2, "with O".rjust(3)), "", "\r\n"]
        for row in reversed(range(board.height)):
            cells = (marks.get(board.states.get(row * board.width + col, -1), '-') for col in range(board.width))
            out.append("%4d" % row + "".join("%2s " % c for c in cells))
        out.append("%4s" % ' ' + "".join("%2d " % col for col in range(board.width)) + "\r\n")
        print("\n".join(out))

    def _finished(self, show, *who):
        if show:
            self.graphic(self.board, *who)
        return self.board.game_end()

    def start_play(self, player1, player2, start_player=0, is_shown=1):
        """game.py:211-290: a match between two player objects; returns the winner (1, 2 or -1) and
        logs the move list in the reference's `{'steps': ...}` format."""
        if start_player not in (0, 1):
            raise Exception('start_player should be either 0 (player1 first) '
                            'or 1 (player2 first)')
        self.board.init_board(start_player)
        seat = dict(zip(self.board.players, (player1, player2)))
        for ind, pl in seat.items():
            pl.set_player_ind(ind)
        shown = (player1.player, player2.player)
        if is_shown:
            self.graphic(self.board, *shown)
        record = []
        ended, winner = False, -1
        while not ended:
            mover = seat[self.board.get_current_player()]
            record.append(mover.get_action(self.board))
            self.board.do_move(record[-1])
            ended, winner = self._finished(is_shown, *shown)
        # the log names the winner by seat order of the opening move (game.py:275-288)
        logged = winner if start_player == 0 else (2 if winner == 1 else 1)
        _logger.warning("{'steps':[%s],'winner':%d,'start_player':%d}"
                        % (",".join(map(str, record)), logged, start_player + 1))
        return winner

    def start_self_play(self, player, is_shown=0, temp=1e-3):
        """game.py:360-416: one self-play game of `player` against itself; returns
        (winner, zip(states, mcts_probs, winners_z)), z from each mover's point of view."""
        self.board.init_board()
        trace = []  # (state, pi, mover, move) per ply
        ended, winner = False, -1
        while not ended:
            move, pi = player.get_action(self.board, temp=temp, return_prob=1)
            trace.append((self.board.current_state(), pi, self.board.current_player, move))
            self.board.do_move(move)
            ended, winner = self._finished(is_shown, *self.board.players)
        movers = np.array([t[2] for t in trace])
        winners_z = np.zeros(len(trace))
        if winner != -1:
            winners_z = np.where(movers == winner, 1.0, -1.0)
        player.reset_player()
        if is_shown:
            print("Game end. Winner is player:", winner) if winner != -1 else print("Game end. Tie")
        _logger.warning("{'steps':[%s],'winner':%d}" % (",".join(str(t[3]) for t in trace), winner))
        return winner, zip([t[0] for t in trace], [t[1] for t in trace], winners_z)

    def start_from_train_data(self, player, train_data, is_shown=0):
        """game.py:292-358: samples from a recorded game `[winner, moves]` (player 1 moves first): the
        state before every move, pi = 0.99999 on the move played and 1e-6 elsewhere, z from each
        mover's point of view.  Returns (winner, zip(states, mcts_probs, winners_z)); an illegal move
        gives the reference's (1, None, None)."""
        winner, record = train_data[0], list(train_data[1])
        self.board.init_board()
        cells = self.board.width * self.board.height
        trace = []  # (state, pi, mover)
        for move in record:
            pi = np.full(cells, 0.000001)
            pi[move] = 0.99999
            trace.append((self.board.current_state(), pi, self.board.current_player))
            try:
                self.board.do_move(move)
            except Exception as e:
                print(str(e))
                return 1, None, None
            if is_shown:
                self.graphic(self.board, *self.board.players)
        if not trace:
            return None
        movers = np.array([t[2] for t in trace])
        winners_z = np.zeros(len(trace))
        if winner != -1:
            winners_z = np.where(movers == winner, 1.0, -1.0)
        player.reset_player()
        if is_shown:
            print("Game end. Winner is player:", winner) if winner != -1 else print("Game end. Tie")
        return winner, zip([t[0] for t in trace], [t[1] for t in trace], winners_z)
