"""GPU: the device move rule of GreedyPlayer (az_greedy_move, players.py:98-123) against the reference fixture and the
oracle, its tie-break distribution, and the batched arena with a Greedy side."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc
from test_greedy_cpu import GAMES

pytestmark = pytest.mark.gpu


def _spec(key):
    from alphazero_b200.gamespec import Connect4Spec, OthelloSpec, TicTacToeSpec
    game, n, h = GAMES[key]
    return OthelloSpec(n) if game == 0 else (Connect4Spec(n, h) if game == 1 else TicTacToeSpec())


def _rel(black, white, player):
    from alphazero_b200 import boardops as bo
    own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
    return bo.bits_from_numpy(own), bo.bits_from_numpy(opp)


@pytest.mark.parametrize("key", sorted(GAMES))
def test_greedy_kernel_matches_reference_fixture(golden_dir, key):
    from alphazero_b200 import boardops as bo
    d = np.load(os.path.join(golden_dir, "greedy_moves.npz"))
    spec = _spec(key)
    own, opp = _rel(d[f"{key}_black"], d[f"{key}_white"], d[f"{key}_player"])
    action, ties = spec.greedy_moves(own, opp, tie_random=False, want_ties=True)
    assert (bo.bits_to_numpy(ties) == d[f"{key}_ties"]).all()
    assert (action.cpu().numpy() == d[f"{key}_best"].astype(np.uint8)).all()
    # random tie-break: always one of the reference's maximisers, and every maximiser gets drawn
    seen = np.zeros(len(d[f"{key}_ties"]), dtype=np.uint64)
    for ctr in range(24):
        a = spec.greedy_moves(own, opp, seed=5, ctr=ctr, tie_random=True).cpu().numpy()
        pick = np.where(a == 64, np.uint64(0), np.uint64(1) << np.minimum(a, 63).astype(np.uint64))
        assert ((pick & ~d[f"{key}_ties"]) == 0).all()
        seen |= pick
    few = np.array([bin(int(t)).count("1") for t in d[f"{key}_ties"]]) <= 3
    assert (seen[few] == d[f"{key}_ties"][few]).all()


def test_greedy_kernel_vs_oracle_on_random_walks():
    """200 k Othello 8x8 positions from random play: best action and tie set identical to the oracle; masked games
    are left alone."""
    from alphazero_b200 import boardops as bo
    from alphazero_b200.gamespec import OthelloSpec
    spec, n, G = OthelloSpec(8), 8, 4000
    b0, w0 = orc.start(n)
    black, white, player = np.full(G, b0, np.uint64), np.full(G, w0, np.uint64), np.ones(G, np.int8)
    rng = np.random.RandomState(3)
    checked = 0
    for ply in range(56):
        own, opp = _rel(black, white, player)
        mask = torch.from_numpy((rng.rand(G) < 0.9).astype(np.uint8)).cuda()
        action, ties = spec.greedy_moves(own, opp, mask=mask, tie_random=False, want_ties=True)
        ob, ot = orc.greedy_moves(0, black, white, player, n)
        m = mask.cpu().numpy().astype(bool)
        a = action.cpu().numpy()
        assert (a[m] == ob[m]).all() and (bo.bits_to_numpy(ties)[m] == ot[m]).all() and (a[~m] == 255).all()
        checked += int(m.sum())
        # advance by a random legal move (the oracle's playout step would do; use the legal masks)
        legal = orc.legal(black, white, player, n)
        act = np.full(G, 64, np.uint8)
        for g in range(G):
            bits = [i for i in range(64) if (int(legal[g]) >> i) & 1]
            if bits:
                act[g] = bits[rng.randint(len(bits))]
        black, white, player, _, st = orc.apply(black, white, player, act, n)
        assert (st == 0).all()
    assert checked > 190_000


def test_greedy_tie_break_is_uniform():
    """The four opening moves of Othello all flip one disc: the random tie-break must be uniform over them."""
    from alphazero_b200.gamespec import OthelloSpec
    spec, G = OthelloSpec(8), 40000
    b0, w0 = orc.start(8)
    own, opp = _rel(np.full(G, b0, np.uint64), np.full(G, w0, np.uint64), np.ones(G, np.int8))
    a = spec.greedy_moves(own, opp, seed=9, ctr=1, tie_random=True).cpu().numpy()
    vals, cnt = np.unique(a, return_counts=True)
    assert sorted(vals.tolist()) == [20, 29, 34, 43]
    assert np.abs(cnt / G - 0.25).max() < 0.012          # 5.5 sigma of a fair four-way draw


def test_batched_arena_greedy_vs_random_matches_the_report():
    """Othello 6x6 Greedy vs Random: the reference's report (Table 3) has 5 995 / 3 530 / 475 over 10 000 games; the
    batched arena must land in the binomial band of those rates (4 000 games, 4.5 sigma)."""
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloBoard
    from alphazero_b200.players import GreedyPlayer, RandomPlayer
    n_games = 4000
    stats = Arena(GreedyPlayer(), RandomPlayer(), OthelloBoard(n=6)).play_games_in_parallel(n_games, return_stats=True, seed=7)
    wins, losses, draws = len(stats["player1"]), len(stats["player2"]), stats["draw"]
    assert wins + losses + draws == n_games
    for got, p in ((wins, 0.5995), (losses, 0.3530), (draws, 0.0475)):
        sigma = (p * (1 - p) * (1 / n_games + 1 / 10000)) ** 0.5        # both are samples
        assert abs(got / n_games - p) < 4.5 * sigma, (wins, losses, draws)


def test_trainer_evaluate_uses_the_batched_arena_for_greedy():
    from alphazero_b200.games.othello import OthelloConfig
    from alphazero_b200.trainer import AlphaZeroTrainer
    cfg = OthelloConfig()
    cfg.board_size, cfg.simulations, cfg.eval_opponent, cfg.eval_episodes, cfg.do_eval = 6, 8, "greedy", 8, True
    tr = AlphaZeroTrainer(verbose=False)
    tr.setup(config=cfg)
    tr.evaluate(0)
    res = tr.eval_results["results"][0]
    total = sum(sum(res[k].values()) for k in ("player1_starts", "player2_starts"))
    assert total == 8 and "Greedy" in tr.eval_results["player2"]
