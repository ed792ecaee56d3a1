"""Arena (alphazero/arena.py:12-277): games between two players, sequentially through the Player API or -- the
reference's "parallel" mode -- all games of a match as ONE device batch instead of a multiprocessing.Pool.

Batched mode supports the players whose move rule runs on the device: MCTSPlayer / AlphaZeroPlayer (TreeBatch search,
greedy on visit counts with random tie-break, tree reuse through both players' moves), RandomPlayer (uniform legal
move) and GreedyPlayer (az_greedy_move).  Other players (HumanPlayer) fall back to nothing: a ValueError says so."""
from collections import defaultdict
from time import time

import numpy as np
import torch

from . import boardops as bo
from .base import Board, Player
from .engine import TreeBatch
from .fused_net import make_evaluator
from .gamespec import spec_of
from .players import AlphaZeroPlayer, GreedyPlayer, MCTSPlayer, RandomPlayer


def _empty_stats() -> dict:
    return {"player1": [], "player2": [], "draw": 0, "player1_starts": defaultdict(int),
            "player2_starts": defaultdict(int)}


class _BatchedSide:
    """Device-side move rule of one arena player for G concurrent games."""

    def __init__(self, player: Player, G: int, spec, seed: int):
        self.player = player
        self.tree = None
        self.evaluator = None
        self.greedy = isinstance(player, GreedyPlayer)
        self.seed, self.n_calls = seed, 0
        self.gen = torch.Generator(device="cuda")
        self.gen.manual_seed(seed)
        if isinstance(player, AlphaZeroPlayer):
            nn = player.mct.nn
            self.n_sim = player.n_sim
            self.tree = TreeBatch(G, n=spec, mode="neural", n_sim_hint=self.n_sim, seed=seed, tie_random=True,
                                  dirichlet_alpha=player.mct.dirichlet_alpha,
                                  dirichlet_epsilon=player.mct.dirichlet_epsilon, enc_dtype=torch.bfloat16,
                                  leaves=getattr(player, "leaves", None))
            self.evaluator = make_evaluator(nn, self.tree.leaf_view)
        elif isinstance(player, MCTSPlayer):
            self.n_sim = player.n_sim
            self.tree = TreeBatch(G, n=spec, mode="rollout", n_sim_hint=self.n_sim, seed=seed, tie_random=True)
        elif not isinstance(player, (RandomPlayer, GreedyPlayer)):
            raise ValueError(f"{player} has no batched device implementation (supported: MCTSPlayer, "
                             "AlphaZeroPlayer, RandomPlayer, GreedyPlayer); use play_games instead")
        if self.tree is not None and self.n_sim is None:
            raise ValueError("batched arena needs n_sim players (compute_time is wall-clock bound per game)")

    def moves(self, own, opp, player, mask, spec) -> torch.Tensor:
        """Actions (uint8, 255 where mask is False) for the games in `mask`."""
        G = own.numel()
        if self.tree is not None:
            self.tree.active.copy_(mask.to(torch.uint8))
            self.tree.search(self.n_sim, self.evaluator, check=False)
            action, _ = self.tree.choose(None)        # temp = 0: MCTSPlayer.get_move default (players.py:161)
            return torch.where(mask, action, torch.full_like(action, 255))
        if self.greedy:        # one-ply look-ahead on get_score with a random tie-break (players.py:98-123)
            self.n_calls += 1
            return spec.greedy_moves(own, opp, mask=mask, seed=self.seed, ctr=self.n_calls, tie_random=True)
        bits, no_move_action = spec.legal_action_bits(spec.legal(own, opp))                         # [G, 64]
        has = bits.sum(dim=1) > 0
        bits[~has, 0] = 1.0                                # dummy support for rows without a move (Othello: pass)
        pick = torch.multinomial(bits, 1, generator=self.gen).squeeze(1).to(torch.uint8)
        action = torch.where(has, pick, torch.full_like(pick, no_move_action))
        return torch.where(mask, action, torch.full_like(action, 255))

    def observe(self, action, live) -> None:
        """apply_move for every live game (arena.py:98-99): both players follow both players' moves."""
        if self.tree is not None:
            self.tree.active.copy_(live.to(torch.uint8))
            self.tree.advance(action)


class Arena:
    def __init__(self, player1: Player, player2: Player, board: Board) -> None:
        self.board = board
        self.game = self.board.game
        self.player1 = player1
        self.player2 = player2

    @staticmethod
    def _check_inputs(n_rounds: int, n_process: int = None, start_player: int = None) -> None:
        if start_player is None and n_rounds % 2 != 0:
            raise ValueError("n_rounds must be an even number or the evaluation will be biased!")
        if n_process is not None:
            if n_process % 2 != 0:
                raise ValueError("n_process must be an even number when evaluating games in parallel or the evaluation will be biased!")
            if n_rounds % n_process != 0:
                raise ValueError("n_rounds must be divisible by the number of processes to play games in parallel!")
        if start_player is not None and start_player not in [1, 2]:
            raise ValueError("start_player must be either 1 or 2 or None (to alternate starts)")

    def play_game(self, player2_starts: bool = False, display: bool = False, save_frames: bool = False,
                  return_results: bool = False, show_indexes: bool = True, show_probs: bool = False,
                  verbose: bool = False) -> dict:
        """One game through the Player API (arena.py:36-117); display options are accepted but rendering is out of
        scope (no matplotlib backend in the engine)."""
        idx = 1 if player2_starts else 0
        self.board.reset()
        self.player1.reset()
        self.player2.reset()
        players = (self.player1, self.player2)
        while not self.board.is_game_over():
            move, _, _, _ = players[idx].get_move(self.board)
            self.board.play_move(move)
            if verbose:
                stats = players[idx].get_stats_after_move()
                print(f"{players[idx]} played {move} | score = {-self.board.get_score()} | " +
                      " | ".join(f"{k} = {v:.3f}" if isinstance(v, float) else f"{k} = {v}" for k, v in stats.items()))
            players[idx].apply_move(move, player=-self.board.player)
            players[1 - idx].apply_move(move, player=self.board.player)
            idx = 1 - idx
        score = abs(self.board.get_score())
        winner = self.board.get_winner()
        if winner == 0:
            result = {"winner": 0, "score": score}
        else:
            p1_wins = (winner == 1 and not player2_starts) or (winner == -1 and player2_starts)
            result = {"winner": 1 if p1_wins else 2, "score": score}
        if verbose:
            print("Draw!" if winner == 0 else f"Player {result['winner']} wins with score = {score}")
        return result if return_results else None

    def play_games(self, n_rounds: int, start_player: int = None, return_stats: bool = False, verbose: bool = False,
                   call_id: int = None) -> dict:
        """n_rounds sequential games (arena.py:119-185), same statistics schema."""
        self._check_inputs(n_rounds, start_player=start_player)
        stats = _empty_stats()
        start = time()
        for round_idx in range(n_rounds):
            p2s = {1: False, 2: True}.get(start_player, bool(round_idx % 2))
            res = self.play_game(player2_starts=p2s, return_results=True, verbose=verbose)
            starter = f"player{2 if p2s else 1}_starts"
            if res["winner"] == 0:
                stats["draw"] += 1
                stats[starter]["draw"] += 1
            else:
                stats[f"player{res['winner']}"].append(res["score"])
                stats[starter]["win" if res["winner"] == (2 if p2s else 1) else "loss"] += 1
        if verbose and call_id is None:
            print(f"Total time = {time() - start:.2f} seconds")
        return stats if return_stats else None

    def play_games_in_parallel(self, n_rounds: int, n_process: int = None, verbose: bool = False,
                               return_stats: bool = False, seed: int = None) -> dict:
        """arena.py:187-257 with the process pool replaced by one device batch of n_rounds games: the first half of
        the games is started by player 2, the second half by player 1 (the reference gives even-numbered workers
        start_player = 2, odd ones 1).  `n_process` only takes part in the reference's argument checks."""
        self._check_inputs(n_rounds, n_process if n_process is not None else 2)
        spec, G = spec_of(self.board), n_rounds      # NotImplementedError for games without device rules
        seed = int(np.random.randint(1 << 31)) if seed is None else seed
        sides = (_BatchedSide(self.player1, G, spec, seed), _BatchedSide(self.player2, G, spec, seed + 1))
        dev = torch.device("cuda")
        black, white = spec.start_position()
        own = torch.full((G,), bo.to_i64(black), dtype=torch.int64, device=dev)
        opp = torch.full((G,), bo.to_i64(white), dtype=torch.int64, device=dev)
        player = torch.ones(G, dtype=torch.int8, device=dev)
        p2_starts = torch.arange(G, device=dev) < G // 2
        # who moves: black (+1) is the starter
        live = torch.ones(G, dtype=torch.bool, device=dev)
        start = time()
        for ply in range(spec.max_plies + 2):
            starter_moves = player == 1
            p1_to_move = live & (starter_moves != p2_starts)
            p2_to_move = live & ~p1_to_move
            a1 = sides[0].moves(own, opp, player, p1_to_move, spec)
            a2 = sides[1].moves(own, opp, player, p2_to_move, spec)
            action = torch.where(p1_to_move, a1, a2)
            for s in sides:
                s.observe(action, live)
            own2, opp2, player2, st = spec.apply(own, opp, player, torch.where(live, action, torch.full_like(action, bo.PASS)))
            if bool((st[live] != 0).any().item()):
                raise RuntimeError("batched arena produced an illegal move")
            own = torch.where(live, own2, own); opp = torch.where(live, opp2, opp); player = torch.where(live, player2, player)
            term, _, _ = spec.terminal_score(own, opp, player)
            live = live & (term == 0)
            if not bool(live.any().item()):
                break
        for s in sides:
            if s.tree is not None:
                s.tree.check_status()            # overflow / depth / a move the tree's position does not have
        _, score, winner = spec.terminal_score(own, opp, player)
        score, winner, p2s = score.abs().cpu().numpy(), winner.cpu().numpy(), p2_starts.cpu().numpy()
        if verbose:
            print(f"Total time = {time() - start:.2f} seconds")
        stats = _empty_stats()
        for g in range(G):
            starter = f"player{2 if p2s[g] else 1}_starts"
            if winner[g] == 0:
                stats["draw"] += 1
                stats[starter]["draw"] += 1
                continue
            p1_wins = (winner[g] == 1 and not p2s[g]) or (winner[g] == -1 and p2s[g])
            who = 1 if p1_wins else 2
            stats[f"player{who}"].append(int(score[g]))
            stats[starter]["win" if who == (2 if p2s[g] else 1) else "loss"] += 1
        return stats if return_stats else None

    @staticmethod
    def print_stats_results(player1: Player, player2: Player, stats: dict):
        n_rounds = len(stats["player1"]) + len(stats["player2"]) + stats["draw"]
        print("\nRESULTS:")
        print(f"- {player1} wins = {len(stats['player1'])}/{n_rounds}")
        print(f"- {player2} wins = {len(stats['player2'])}/{n_rounds}")
        print(f"- Draws = {stats['draw']}/{n_rounds}")
        for who, p in (("player1", player1), ("player2", player2)):
            s = stats[f"{who}_starts"]
            print(f"- {p} results when starting: win={s['win']} | loose={s['loss']} | draw={s['draw']}")
