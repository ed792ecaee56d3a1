"""GPU: parity evidence at the sample sizes the north star asks for ("millions of random positions", "arena win-rate
parity must hold statistically", "policy/value within 1e-2 of the reference's fp32"):
  * the CUDA board kernels reproduce the SHA-256 digests the UNMODIFIED reference computed over ~1 M positions per game
    (legal sets, game over, score, winner, and -- through the walk -- every flip);
  * all 1 M playouts of BASELINE config 2 bit-exact against the oracle;
  * OthelloNet(n=8): every evaluator against the reference's own fp32 outputs on reference-held weights;
  * the batched arena against the reference Arena's win / loss / draw counts (two-sided test on proportions);
  * the Dirichlet root noise has the moments of Dir(alpha)."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc
from digest_walk import walk_digest

pytestmark = pytest.mark.gpu


def _dev_backend(spec):
    from alphazero_b200 import boardops as bo

    def rel(b, w, p):
        own, opp = np.where(p > 0, b, w), np.where(p > 0, w, b)
        return bo.bits_from_numpy(own), bo.bits_from_numpy(opp), torch.from_numpy(p.astype(np.int8)).cuda()

    def legal(b, w, p):
        own, opp, _ = rel(b, w, p)
        return bo.bits_to_numpy(spec.legal(own, opp))

    def apply(b, w, p, a):
        own, opp, pl = rel(b, w, p)
        o2, p2, pl2, st = spec.apply(own, opp, pl, torch.from_numpy(a).cuda())
        o2, p2, pl2 = bo.bits_to_numpy(o2), bo.bits_to_numpy(p2), pl2.cpu().numpy()
        return np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2), pl2

    def terminal(b, w, p):
        own, opp, pl = rel(b, w, p)
        t, s, wn = spec.terminal_score(own, opp, pl)
        return t.cpu().numpy(), s.cpu().numpy(), wn.cpu().numpy()

    return legal, apply, terminal


@pytest.mark.parametrize("key", ["othello8", "othello6", "connect4_7x6"])
def test_board_kernels_reproduce_the_reference_digests(golden_dir, key):
    from alphazero_b200.gamespec import Connect4Spec, OthelloSpec
    spec = {"othello8": OthelloSpec(8), "othello6": OthelloSpec(6), "connect4_7x6": Connect4Spec(7, 6)}[key]
    d = json.load(open(os.path.join(golden_dir, "reference_digests.json")))["games"][key]
    legal, apply, terminal = _dev_backend(spec)
    total = 0
    for ch in d["chunks"]:
        sha, n = walk_digest(ch["seed"], ch["games"], spec.start_position(), legal, apply, terminal)
        assert n == ch["positions"] and sha == ch["sha256"], (key, ch["seed"])
        total += n
    assert total >= 1_000_000


def test_all_one_million_playouts_match_the_oracle():
    """BASELINE config 2 (ii): 1 M random playouts from the start position, Philox4x32-10 -- final boards, winner and
    length of EVERY playout against the oracle (its C loop does the million in a few minutes on one core: the oracle
    side is cached in a digest per 100 k slice so that repeated runs only pay for the device side)."""
    from alphazero_b200 import boardops as bo
    n, seed, total = 8, 0x5EED, 1_000_000
    b, w = orc.start(n)
    own = bo.bits_from_numpy(np.array([b], np.uint64)); opp = bo.bits_from_numpy(np.array([w], np.uint64))
    fb, fw, win, plies = bo.playouts(own, opp, torch.ones(1, dtype=torch.int8, device="cuda"), n, total, seed)
    fb, fw, win, plies = bo.bits_to_numpy(fb), bo.bits_to_numpy(fw), win.cpu().numpy(), plies.cpu().numpy()
    step = 100_000
    for first in range(0, total, step):
        ofb, ofw, owin, opl = orc.playouts(np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8),
                                           n, step, seed, first_id=first)
        sl = slice(first, first + step)
        assert (fb[sl] == ofb).all() and (fw[sl] == ofw).all() and (win[sl] == owin).all() and (plies[sl] == opl).all()


def _golden_othello_net(golden_dir):
    from alphazero_b200.games.othello import OthelloNet
    d = np.load(os.path.join(golden_dir, "othello_net_n8.npz"))
    net = OthelloNet(n=8)
    net.load_state_dict({k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith("sd.")})
    net.eval()
    return net, d


@pytest.mark.parametrize("path", ["stack", "library", "torch"])
def test_othello_net_matches_reference_outputs(golden_dir, path):
    """The headline network on weights and inputs held by the reference: probs / value of every evaluator within 1e-2
    of the reference's own fp32 predict() (othello.py:366-382, base.py:350-355)."""
    from alphazero_b200 import boardops as bo
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.evaluators import NetEvaluator
    from alphazero_b200.fused_net import FusedNetEvaluator
    net, d = _golden_othello_net(golden_dir)
    player = d["player"].astype(np.int8)
    own = np.where(player > 0, d["black"], d["white"]); opp = np.where(player > 0, d["white"], d["black"])
    G = len(own)
    tree = TreeBatch(G, n=8, mode="neural", capacity=64, enc_dtype=torch.bfloat16)
    tree.leaf_own.copy_(bo.bits_from_numpy(own)); tree.leaf_opp.copy_(bo.bits_from_numpy(opp))
    assert (bo.encode(tree.leaf_own, tree.leaf_opp, 8).cpu().numpy() == d["input"]).all()
    if path == "torch":
        tree.enc.copy_(bo.encode(tree.leaf_own, tree.leaf_opp, 8, dtype=torch.bfloat16))
        ev = NetEvaluator(net, tree)
    else:
        ev = FusedNetEvaluator(net, tree, heads=path)
    ev(tree)
    torch.cuda.synchronize()
    assert np.abs(tree.probs.cpu().numpy() - d["probs"]).max() < 1e-2
    assert np.abs(tree.value.cpu().numpy() - d["value"].reshape(-1)).max() < 1e-2


def _two_sample_ok(k1, n1, k2, n2, z=4.0):
    p = (k1 + k2) / (n1 + n2)
    return abs(k1 / n1 - k2 / n2) <= z * (p * (1 - p) * (1 / n1 + 1 / n2)) ** 0.5 + 1e-12


@pytest.mark.parametrize("opponent", ["random", "mcts"])
def test_arena_win_rates_match_the_reference_arena(golden_dir, opponent):
    """AlphaZeroPlayer(n_sim = 25, seeded OthelloNet 6x6, bf16 engine) in the batched arena against the W / L / D counts
    of the reference's own Arena.play_games with the fp32 CPU player (tests/golden/make_golden_arena.py): two-sided
    4-sigma test on the win and the loss proportion."""
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloBoard, OthelloNet
    from alphazero_b200.players import AlphaZeroPlayer, MCTSPlayer, RandomPlayer
    fx = json.load(open(os.path.join(golden_dir, "arena_othello6.json")))
    torch.manual_seed(fx["net_seed"])
    net = OthelloNet(n=fx["board"])
    h = hashlib.sha256()
    for k, t in net.state_dict().items():
        h.update(k.encode()); h.update(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes())
    assert h.hexdigest() == fx["net_state_dict_sha256"], "the seeded network differs from the reference's"
    ref = fx["matches"][opponent]
    G = 2000
    other = RandomPlayer() if opponent == "random" else MCTSPlayer(n_sim=fx["n_sim"])
    stats = Arena(AlphaZeroPlayer(n_sim=fx["n_sim"], nn=net), other, OthelloBoard(n=fx["board"])).play_games_in_parallel(
        G, return_stats=True, seed=11)
    w1, w2 = len(stats["player1"]), len(stats["player2"])
    assert w1 + w2 + stats["draw"] == G
    assert _two_sample_ok(w1, G, ref["player1_wins"], ref["games"]), (w1, w2, stats["draw"], ref)
    assert _two_sample_ok(w2, G, ref["player2_wins"], ref["games"]), (w1, w2, stats["draw"], ref)


def test_dirichlet_root_noise_has_the_moments_of_dir_alpha():
    """mcts.py:234-240: P' = (1 - eps) P + eps Dir(alpha).  From 20 000 games at the start position (4 children):
    the recovered noise has mean 1/k and variance (1/k)(1 - 1/k)/(k alpha + 1), and sums to one per game."""
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.evaluators import HashEvaluator
    G, alpha, eps, k = 20000, 0.3, 0.25, 4
    base = TreeBatch(8, n=8, mode="neural", tie_random=False, capacity=64)
    base.search(2, HashEvaluator(3))
    P0 = base.root_stats()[2][0]
    tree = TreeBatch(G, n=8, mode="neural", tie_random=False, capacity=64, dirichlet_alpha=alpha, dirichlet_epsilon=eps, seed=9)
    tree.search(2, HashEvaluator(3))
    P = tree.root_stats()[2]
    legal = P0 > 0
    assert int(legal.sum()) == k
    noise = ((P[:, legal].double() - (1 - eps) * P0[legal].double()) / eps).cpu().numpy()
    assert np.allclose(noise.sum(axis=1), 1.0, atol=1e-5) and (noise > -1e-6).all()
    var = (1 / k) * (1 - 1 / k) / (k * alpha + 1)
    assert np.abs(noise.mean(axis=0) - 1 / k).max() < 4 * (var / G) ** 0.5 + 1e-4
    assert np.abs(noise.var(axis=0) - var).max() < 0.05 * var
    # the reference's default alpha = 0.03 (othello.py:26): almost all mass on one child
    t2 = TreeBatch(G, n=8, mode="neural", tie_random=False, capacity=64, dirichlet_alpha=0.03, dirichlet_epsilon=eps, seed=10)
    t2.search(2, HashEvaluator(3))
    n2 = ((t2.root_stats()[2][:, legal].double() - (1 - eps) * P0[legal].double()) / eps).cpu().numpy()
    var2 = (1 / k) * (1 - 1 / k) / (k * 0.03 + 1)
    assert np.abs(n2.mean(axis=0) - 1 / k).max() < 4 * (var2 / G) ** 0.5 + 1e-4
    assert np.abs(n2.var(axis=0) - var2).max() < 0.05 * var2
