"""Players with the reference's interface (alphazero/players.py:76-247): Random, Greedy, MCTS, AlphaZero.

get_move(board, temp) -> (move, action_probs | None, visit_counts | None, prior_probs | None); apply_move keeps an
MCTS player's root in sync with the board (tree reuse).  HumanPlayer (stdin UI) is out of scope.
For many games at once use alphazero_b200.arena.Arena.play_games_in_parallel / SelfPlayEngine, which run the same
search for all games in one device batch."""
from time import sleep

import numpy as np

from .base import Board, Player, PolicyValueNetwork
from .mcts import MCT
from .utils import fair_max


class RandomPlayer(Player):
    """Uniform random legal move (players.py:76-95)."""

    def __init__(self, lock_time: float = None, verbose: bool = False) -> None:
        super().__init__(verbose=verbose)
        self.lock_time = lock_time

    def clone(self) -> "RandomPlayer":
        return RandomPlayer(lock_time=self.lock_time, verbose=self.verbose)

    def get_move(self, board: Board, temp: float = None):
        if self.lock_time is not None:
            sleep(self.lock_time)
        return board.get_random_move(), None, None, None


class GreedyPlayer(Player):
    """One-ply look-ahead on board.get_score() with a random tie-break (players.py:98-123)."""

    def clone(self) -> "GreedyPlayer":
        return GreedyPlayer(verbose=self.verbose)

    def get_move(self, board: Board, temp: float = None):
        scores = {}
        for move in board.get_moves():
            nxt = board.clone()
            nxt.play_move(move)
            scores[move] = -nxt.get_score()     # get_score is from the next mover's viewpoint
        return fair_max(scores.items(), key=lambda x: x[1])[0], None, None, None


class MCTSPlayer(Player):
    """Rollout-mode MCTS player (players.py:126-195)."""

    def __init__(self, n_sim: int = None, compute_time: float = None, verbose: bool = False) -> None:
        super().__init__(verbose=verbose)
        self.n_sim = n_sim
        self.compute_time = compute_time
        if self.n_sim is None and self.compute_time is None:
            raise ValueError("MCTSPlayer needs to have either n_sim or compute_time specified.")
        if self.n_sim is not None and self.compute_time is not None:
            raise ValueError("MCTSPlayer can't have both n_sim and compute_time specified.")
        self.mct = MCT()

    def clone(self) -> "MCTSPlayer":
        return MCTSPlayer(n_sim=self.n_sim, compute_time=self.compute_time, verbose=self.verbose)

    def reset(self) -> None:
        old, self.mct = self.mct, MCT()
        self.mct._adopt(old)

    def apply_move(self, move, player: int = None) -> None:
        self.mct.change_root(move)

    def get_move(self, board: Board, temp: float = 0):
        if board.is_game_over():
            raise ValueError(f"{self}.get_move was called with a board in game over state...")
        self.mct.search(board=board, n_sim=self.n_sim, compute_time=self.compute_time)
        action_probs, visit_counts = self.mct.get_action_probs(board, temp)
        items = list(action_probs.items())
        if len(items) == 1:
            best_action = items[0][0]
        else:
            best_action = items[np.random.choice(len(items), p=[p for _, p in items])][0]
        return best_action, action_probs, visit_counts, self.mct.get_prior_probs()

    def get_stats_after_move(self) -> dict:
        n_rollouts, simulation_time = self.mct.get_stats()
        return {"n_rollouts": n_rollouts, "time": simulation_time}


class AlphaZeroPlayer(MCTSPlayer):
    """MCTS with network leaf evaluation and optional Dirichlet root noise (players.py:198-234)."""

    def __init__(self, n_sim: int = None, compute_time: float = None, nn: PolicyValueNetwork = None,
                 dirichlet_alpha: float = None, dirichlet_epsilon: float = None, verbose: bool = False,
                 leaves: int = None) -> None:
        """`leaves` is an extension (throughput mode, see mcts.MCT); None keeps the reference's search order."""
        super().__init__(n_sim=n_sim, compute_time=compute_time, verbose=verbose)
        self.leaves = leaves
        self.mct = MCT(eval_method="neural", nn=nn, leaves=leaves)
        self.mct.dirichlet_alpha = dirichlet_alpha
        self.mct.dirichlet_epsilon = dirichlet_epsilon

    def clone(self) -> "AlphaZeroPlayer":
        return AlphaZeroPlayer(n_sim=self.n_sim, compute_time=self.compute_time,
                               nn=self.mct.nn.clone() if self.mct.nn is not None else None,
                               dirichlet_alpha=self.mct.dirichlet_alpha, dirichlet_epsilon=self.mct.dirichlet_epsilon,
                               verbose=self.verbose, leaves=self.leaves)

    def reset(self) -> None:
        old = self.mct
        self.mct = MCT(eval_method="neural", nn=old.nn, dirichlet_alpha=old.dirichlet_alpha,
                       dirichlet_epsilon=old.dirichlet_epsilon, leaves=self.leaves)
        self.mct._adopt(old)


PLAYERS_SET = {"random", "greedy", "mcts", "alphazero"}
PLAYERS_REGISTER = {"random": RandomPlayer, "greedy": GreedyPlayer, "mcts": MCTSPlayer, "alphazero": AlphaZeroPlayer}
