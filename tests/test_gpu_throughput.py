"""GPU: throughput mode (north_star (2)): up to L leaves of a game in flight per step under virtual
loss.  The node-for-node identity with the oracle's restatement of the same rules (and the L = 1 degeneration to the
reference) is in test_gpu_tree.py; here: how far the L > 1 search drifts from the reference's order, an arena match
between the two orders, and self-play through the engine."""
import numpy as np
import pytest
import torch

import oracle_lib as orc

pytestmark = pytest.mark.gpu


def _random_roots(n, G, seed):
    rng = np.random.default_rng(seed)
    b0, w0 = orc.start(n)
    black, white, player = np.full(G, b0, np.uint64), np.full(G, w0, np.uint64), np.ones(G, np.int8)
    depth = rng.integers(4, n * n - 20, size=G)
    for ply in range(int(depth.max())):
        legal = orc.legal(black, white, player, n)
        term, _, _ = orc.terminal(black, white, player, n)
        act = np.full(G, 64, np.uint8)
        for g in range(G):
            bits = [i for i in range(64) if (int(legal[g]) >> i) & 1]
            if bits:
                act[g] = bits[rng.integers(len(bits))]
        go = (ply < depth) & (term == 0)
        ob, ow, op, _, _ = orc.apply(black, white, player, act, n)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    term, _, _ = orc.terminal(black, white, player, n)
    keep = term == 0
    return black[keep], white[keep], player[keep]


@pytest.mark.parametrize("leaves,max_mean_tv,min_agree", [(2, 0.10, 0.75), (4, 0.15, 0.70)])
def test_visit_distribution_stays_close_to_the_reference_order(leaves, max_mean_tv, min_agree):
    """200 simulations from ~250 random mid-game roots under the hash evaluator (flat, noisy values: the hard case for
    virtual loss): the root visit distribution of the L-leaf search against the reference-order search.  Calibrated
    with the oracle on the CPU: mean total-variation distance 0.064 (L = 2) / 0.10 (L = 4), same best move 85 % / 80 %."""
    from alphazero_b200 import boardops as bo
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.evaluators import HashEvaluator
    n, n_sim = 8, 200
    black, white, player = _random_roots(n, 256, seed=1)
    G = len(black)
    own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
    fracs = {}
    for L in (None, leaves):
        tree = TreeBatch(G, n=n, mode="neural", tie_random=False, n_sim_hint=n_sim, leaves=L)
        tree.set_roots(bo.bits_from_numpy(own), bo.bits_from_numpy(opp), torch.from_numpy(player.astype(np.int8)).cuda())
        tree.search(n_sim, HashEvaluator(11, tree.leaf_view))
        N, _, _, rn = tree.root_stats()
        N = N.cpu().numpy().astype(np.float64)
        assert (rn.cpu().numpy() == n_sim).all()              # every game got exactly n_sim simulations
        assert (tree.sims_done.cpu().numpy() == n_sim).all()
        assert (np.where(N >= 0, N, 0).sum(axis=1) == n_sim).all()
        fracs[L] = np.where(N >= 0, N, 0) / n_sim
    tv = 0.5 * np.abs(fracs[None] - fracs[leaves]).sum(axis=1)
    agree = (fracs[None].argmax(axis=1) == fracs[leaves].argmax(axis=1)).mean()
    assert tv.mean() < max_mean_tv and agree > min_agree, (tv.mean(), agree)


def test_no_virtual_loss_is_left_behind():
    """After a search every node's in-flight counter is back to zero and W is the sum the running means imply."""
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.evaluators import HashEvaluator
    G, n_sim = 64, 100
    tree = TreeBatch(G, n=8, mode="neural", tie_random=False, n_sim_hint=n_sim, leaves=4, capacity=4096)
    tree.search(n_sim, HashEvaluator(3, tree.leaf_view))
    words = tree.nodes.view(torch.int32).view(G, 2, tree.capacity, 8)
    used = tree.n_alloc.cpu().numpy()
    half = tree.half.cpu().numpy()
    for g in range(0, G, 7):
        rec = words[g, int(half[g]), :int(used[g])].cpu()
        vloss = rec[:, 6].numpy()
        assert (vloss == 0).all()
        N = rec[:, 3].numpy()
        Q = rec[:, 0:2].contiguous().view(torch.float64).reshape(-1).numpy()
        W = rec[:, 7].contiguous().view(torch.float32).reshape(-1).numpy()
        assert np.allclose(W, N * Q, atol=1e-3 * np.maximum(1, N))
    assert (tree.leaf_kind == 0).all()


def test_arena_throughput_order_against_reference_order():
    """The same random-init network searched with 4 leaves in flight against the reference's order: over 600 games
    neither side may be ahead by more than a 4-sigma binomial band (a broken virtual loss loses badly)."""
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloBoard, OthelloNet
    from alphazero_b200.players import AlphaZeroPlayer
    torch.manual_seed(3)
    net = OthelloNet(n=6)
    G = 600
    stats = Arena(AlphaZeroPlayer(n_sim=32, nn=net, leaves=4), AlphaZeroPlayer(n_sim=32, nn=net),
                  OthelloBoard(n=6)).play_games_in_parallel(G, return_stats=True, seed=5)
    w1, w2 = len(stats["player1"]), len(stats["player2"])
    assert w1 + w2 + stats["draw"] == G
    assert abs(w1 - w2) < 4.0 * (w1 + w2) ** 0.5, (w1, w2, stats["draw"])


@pytest.mark.parametrize("game", ["othello6", "connect4"])
def test_self_play_with_leaves_in_flight(game):
    """SelfPlayEngine(leaves = 4) with the fused network: every move gets exactly n_sim simulations, the samples are
    legal positions with normalised policies, the evaluator saw G*4 leaf slots."""
    from alphazero_b200.gamespec import Connect4Spec, OthelloSpec
    from alphazero_b200.selfplay import SelfPlayEngine
    torch.manual_seed(0)
    if game == "connect4":
        from alphazero_b200.games.connect4 import Connect4Net
        spec, net = Connect4Spec(7, 6), Connect4Net(board_width=7, board_height=6)
    else:
        from alphazero_b200.games.othello import OthelloNet
        spec, net = OthelloSpec(6), OthelloNet(n=6)
    G, n_sim = 96, 30
    eng = SelfPlayEngine(net, G, n=spec, n_sim=n_sim, seed=4, leaves=4)
    assert eng.tree.leaf_view.G == 4 * G and eng.evaluator.tree.G == 4 * G
    batch = eng.play()
    assert eng.total_sims == eng.total_moves * n_sim and len(batch) == eng.total_moves
    pi = batch.pi.cpu().numpy()
    assert np.allclose(pi.sum(axis=1), 1.0, atol=1e-5)
    z = batch.z.cpu().numpy()
    assert set(np.unique(z)) <= {-1, 0, 1}
    # the same engine replays the whole-search graph from the second move on and stays consistent on a second batch
    batch2 = eng.play()
    assert eng.total_sims == eng.total_moves * n_sim and len(batch2) == eng.total_moves
