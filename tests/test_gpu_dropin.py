"""GPU: the reference-facing classes (OthelloBoard, MCT, players, Arena, trainer.self_play) on top of the kernels."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc
from hash_eval import bits_to_grid

pytestmark = pytest.mark.gpu


def _act_to_move(a, n):
    return (n, n) if a == 64 else (a // 8, a % 8)


@pytest.mark.parametrize("n", [4, 6, 8])
def test_othello_board_api_vs_reference_fixture(golden_dir, n):
    from alphazero_b200.games.othello import OthelloBoard
    d = np.load(os.path.join(golden_dir, f"othello_positions_n{n}.npz"))
    for i in range(0, len(d["black"]), max(1, len(d["black"]) // 150)):
        grid = bits_to_grid(int(d["black"][i]), int(d["white"][i]), n)
        b = OthelloBoard(n=n, grid=grid, player=int(d["player"][i]))
        assert (b.grid == grid).all() and b.get_board_shape() == (n, n) and b.get_action_size() == n * n + 1
        legal = int(d["legal"][i])
        want = {(a // 8, a % 8) for a in range(64) if (legal >> a) & 1} or {(n, n)}
        assert set(b.get_moves()) == want
        assert b.is_legal_move((n, n)) == (legal == 0)
        assert b.is_game_over() == bool(d["over"][i]) and b.get_score() == d["score"][i]
        assert b.get_random_move() in want
        if d["over"][i]:
            assert b.get_winner() == d["winner"][i]
            continue
        with pytest.raises(ValueError):
            b.get_winner()
        occupied = next((r, c) for r in range(n) for c in range(n) if grid[r, c] != 0)
        with pytest.raises(ValueError):
            b.play_move(occupied)
        c = b.clone()
        c.play_move(_act_to_move(int(d["action"][i]), n))
        assert (c.grid == bits_to_grid(int(d["black2"][i]), int(d["white2"][i]), n)).all()
        assert c.player == d["player2"][i] and b.player == d["player"][i]      # clone is independent
        # get_moves(player=...) overrides the side to move (othello.py:143)
        other = OthelloBoard(n=n, grid=grid, player=-int(d["player"][i]))
        assert set(b.get_moves(player=-b.player)) == set(other.get_moves())
    with pytest.raises(ValueError):
        OthelloBoard(n=5)
    assert str(OthelloBoard(n=8)) == "OthelloBoard8" and OthelloBoard(n=8).max_moves == 60


class _HashNet:
    """Network stand-in whose evaluator is the device hash evaluator (visit-count identity through the MCT API)."""

    def __init__(self, n, salt):
        self.n, self.salt = n, salt

    def make_batched_evaluator(self, tree):
        from alphazero_b200.evaluators import HashEvaluator
        return HashEvaluator(self.salt)


@pytest.mark.parametrize("n", [6, 8])
def test_mct_facade_identical_to_reference_fixture(golden_dir, n):
    """MCT.search / root.children / get_action_probs / change_root reproduce alphazero.mcts.MCT on the fixture."""
    from alphazero_b200.games.othello import OthelloBoard
    from alphazero_b200.mcts import MCT
    d = np.load(os.path.join(golden_dir, f"mct_visits_n{n}.npz"))
    salt = int(d["salt"][0])
    MCT.TIE_RANDOM = False
    try:
        rows = len(d["root"])
        i = 0
        checked = 0
        while i < rows and checked < 40:
            mct = MCT(eval_method="neural", nn=_HashNet(n, salt))
            j = i
            while j < rows and d["root"][j] == d["root"][i] and d["n_sim"][j] == d["n_sim"][i] and (j == i or d["step"][j] > d["step"][j - 1]):
                board = OthelloBoard(n=n, grid=bits_to_grid(int(d["black"][j]), int(d["white"][j]), n), player=int(d["player"][j]))
                mct.search(board, n_sim=int(d["n_sim"][j]))
                assert mct.n_rollouts == d["n_sim"][j] and mct.get_stats()[0] == d["n_sim"][j]
                root = mct.root
                assert root.N == d["rootN"][j]
                for a in range(65):
                    mv = _act_to_move(a, n)
                    if d["N"][j][a] < 0:
                        assert mv not in root.children
                    else:
                        ch = root.children[mv]
                        assert ch.N == d["N"][j][a] and ch.Q == d["Q"][j][a] and ch.P == d["P"][j][a]
                probs, visits = mct.get_action_probs(board, temp=1)
                tot = sum(max(int(v), 0) for v in d["N"][j])
                assert all(abs(probs[m] - visits[m] / tot) < 1e-12 for m in probs)
                best = _act_to_move(int(d["action"][j]), n)
                onehot, _ = mct.get_action_probs(board, temp=0)
                assert list(onehot.values()) == [1] and visits[list(onehot)[0]] == max(visits.values())
                assert set(mct.get_prior_probs()) == set(visits)
                mct.change_root(best)
                j += 1
                checked += 1
            i = j
    finally:
        MCT.TIE_RANDOM = True


def test_players_and_sequential_arena_small_board():
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloBoard, OthelloNet
    from alphazero_b200.players import AlphaZeroPlayer, GreedyPlayer, MCTSPlayer, RandomPlayer
    np.random.seed(0)
    torch.manual_seed(0)
    board = OthelloBoard(n=6)
    az = AlphaZeroPlayer(n_sim=12, nn=OthelloNet(n=6), dirichlet_alpha=0.03, dirichlet_epsilon=0.25)
    move, probs, visits, priors = az.get_move(board, temp=1)
    assert move in board.get_moves() and abs(sum(probs.values()) - 1) < 1e-9
    assert sum(visits.values()) == 12 and set(priors) == set(visits) and abs(sum(priors.values()) - 1) < 1e-5
    assert az.get_stats_after_move()["n_rollouts"] == 12
    clone = az.clone()
    assert clone.n_sim == 12 and clone.mct.dirichlet_alpha == 0.03 and clone.mct.nn is not az.mct.nn
    stats = Arena(MCTSPlayer(n_sim=8), RandomPlayer(), OthelloBoard(n=4)).play_games(4, return_stats=True)
    assert len(stats["player1"]) + len(stats["player2"]) + stats["draw"] == 4
    assert sum(stats["player1_starts"].values()) == 2 and sum(stats["player2_starts"].values()) == 2
    res = Arena(GreedyPlayer(), az, OthelloBoard(n=6)).play_game(player2_starts=True, return_results=True)
    assert res["winner"] in (0, 1, 2) and 0 <= res["score"] <= 36
    over = OthelloBoard(n=4, grid=np.ones((4, 4)))
    with pytest.raises(ValueError):
        MCTSPlayer(n_sim=2).get_move(over)


def test_batched_arena_win_rates():
    """Arena parallel mode as one device batch: rollout-MCTS(n_sim=30) crushes Random on 6x6 (report Table 3:
    MCTS(100) vs Random 99/100), Random vs Random is balanced, and the statistics schema adds up."""
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloBoard
    from alphazero_b200.players import MCTSPlayer, RandomPlayer
    G = 400
    stats = Arena(MCTSPlayer(n_sim=30), RandomPlayer(), OthelloBoard(n=6)).play_games_in_parallel(G, n_process=2, return_stats=True, seed=1)
    p1, p2 = len(stats["player1"]), len(stats["player2"])
    assert p1 + p2 + stats["draw"] == G
    assert p1 / G > 0.85, stats
    for who in ("player1_starts", "player2_starts"):
        assert sum(stats[who].values()) == G // 2
    assert stats["player1_starts"]["win"] + stats["player2_starts"]["loss"] == p1
    assert all(1 <= s <= 36 for s in stats["player1"] + stats["player2"])
    rr = Arena(RandomPlayer(), RandomPlayer(), OthelloBoard(n=6)).play_games_in_parallel(2000, return_stats=True, seed=2)
    w1, w2 = len(rr["player1"]), len(rr["player2"])
    assert abs(w1 - w2) / 2000 < 0.08 and w1 + w2 + rr["draw"] == 2000     # ~4 sigma band
    with pytest.raises(ValueError):
        Arena(RandomPlayer(), RandomPlayer(), OthelloBoard(n=6)).play_games_in_parallel(7)


def test_rollout_mode_matches_oracle_statistically():
    """Rollout mode cannot be stream-identical to the reference (NumPy global RNG, T7); the visit distribution over
    root moves after 150 simulations must agree with the oracle's UCT search over many independent trees."""
    from alphazero_b200 import boardops as bo
    from alphazero_b200.engine import TreeBatch
    n, n_sim, G = 6, 150, 1024
    rng = np.random.default_rng(3)
    b, w = orc.start(n)
    black, white, player = np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8)
    for _ in range(9):   # an asymmetric mid-game root
        m = int(orc.legal(black, white, player, n)[0]); bits = [i for i in range(64) if (m >> i) & 1] or [64]
        black, white, player, _, _ = orc.apply(black, white, player, np.array([bits[rng.integers(len(bits))]], np.uint8), n)
    own, opp = (black, white) if player[0] > 0 else (white, black)
    tree = TreeBatch(G, n=n, mode="rollout", tie_random=True, seed=5, n_sim_hint=n_sim)
    tree.set_roots(bo.bits_from_numpy(np.repeat(own, G)), bo.bits_from_numpy(np.repeat(opp, G)),
                   torch.full((G,), int(player[0]), dtype=torch.int8, device="cuda"))
    tree.search(n_sim)
    N, Q, _, rn = tree.root_stats()
    N = N.cpu().numpy().astype(np.float64)
    assert (rn.cpu().numpy() == n_sim).all() and (np.where(N >= 0, N, 0).sum(axis=1) == n_sim).all()
    dev_frac = np.where(N >= 0, N, 0).mean(axis=0) / n_sim
    runs = 300
    acc = np.zeros(65)
    for r in range(runs):
        o = orc.OracleMCT(n, "rollout", tie_random=True, seed=1000 + r)
        o.search(int(black[0]), int(white[0]), int(player[0]), n_sim)
        oN = o.stats()[0].astype(np.float64)
        acc += np.where(oN >= 0, oN, 0)
    orc_frac = acc / runs / n_sim
    assert set(np.nonzero(dev_frac)[0]) == set(np.nonzero(orc_frac)[0])
    assert np.abs(dev_frac - orc_frac).max() < 0.03, (dev_frac[dev_frac > 0], orc_frac[orc_frac > 0])


def test_trainer_self_play_samples_and_augmentation():
    from alphazero_b200.games.othello import OthelloConfig, OthelloNet
    from alphazero_b200.trainer import AlphaZeroTrainer
    torch.manual_seed(0)
    cfg = OthelloConfig(board_size=6, simulations=8, episodes=6, data_augmentation=True)
    tr = AlphaZeroTrainer()
    tr.setup("othello", cfg, OthelloNet(n=6))
    tr.self_play(0, seed=4)
    base = [s for s in tr.memory if s.transformation is None]
    aug = [s for s in tr.memory if s.transformation is not None]
    assert len(aug) == 7 * sum(1 for s in base if s.move_idx >= 2)          # x8 for move_idx >= 2 (trainer.py:275-284)
    assert {s.episode_idx for s in base} == set(range(6))
    for s in tr.memory:
        assert s.player == 1 and s.outcome in (-1, 0, 1) and s.state.shape == (6, 6) and abs(s.pi.sum() - 1) < 1e-5
    first = [s for s in base if s.move_idx == 0][0]
    assert first.state[2, 2] == 1 and first.state[2, 3] == -1
    kinds = {s.transformation for s in aug}
    assert "reflection_horizontal" in kinds and "reflection_horizontal+rotation_270" in kinds and "rotation_90" in kinds
    tr.update_network(OthelloNet(n=6))       # network swap refreshes the device snapshot
    tr.self_play(1, seed=5)
    assert len(tr.memory) > 0


def test_players_pickle_without_their_device_tree():
    """The reference ships boards / players into a multiprocessing.Pool (arena.py:211-234): GPU-backed players must
    pickle (the device tree stays behind, the copy starts from a fresh root) and a TreeBatch must refuse loudly."""
    import pickle
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.games.othello import OthelloBoard, OthelloNet
    from alphazero_b200.players import AlphaZeroPlayer, MCTSPlayer
    board = OthelloBoard(n=6)
    torch.manual_seed(0)
    for player in (MCTSPlayer(n_sim=20), AlphaZeroPlayer(n_sim=20, nn=OthelloNet(n=6))):
        move, _, counts, _ = player.get_move(board)
        assert player.mct._tree is not None and sum(counts.values()) == 20
        twin = pickle.loads(pickle.dumps(player))
        assert twin.mct._tree is None and twin.n_sim == 20
        move2, _, counts2, _ = twin.get_move(board)
        assert move2 in board.get_moves() and sum(counts2.values()) == 20
    b2 = pickle.loads(pickle.dumps(board))
    assert (b2.grid == board.grid).all() and b2.player == board.player
    with pytest.raises(TypeError, match="cannot be pickled"):
        pickle.dumps(TreeBatch(2, n=6, mode="rollout", capacity=64))


def test_mct_follows_in_place_weight_updates():
    """The reference MCT evaluates the live module (base.py:357-367): load_state_dict / optimizer steps on the same
    object must reach the evaluator's bf16 snapshot without going through the nn setter."""
    from alphazero_b200.games.othello import OthelloBoard, OthelloNet
    from alphazero_b200.players import AlphaZeroPlayer
    torch.manual_seed(1)
    net, other = OthelloNet(n=6), OthelloNet(n=6)
    with torch.no_grad():
        other.fc_probs.weight.mul_(6.0)
    board = OthelloBoard(n=6)
    player = AlphaZeroPlayer(n_sim=4, nn=net)
    _, _, _, prior_a = player.get_move(board)
    player.reset()
    net.load_state_dict(other.state_dict())               # same object, new weights
    _, _, _, prior_b = player.get_move(board)
    fresh = AlphaZeroPlayer(n_sim=4, nn=other)
    _, _, _, prior_c = fresh.get_move(board)
    assert prior_b == prior_c and prior_a != prior_b
