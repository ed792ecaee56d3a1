"""GPU: batched MCTS tree kernels vs the reference MCT (golden fixtures) and vs the oracle MCT.

With the deterministic hash evaluator and the lowest-index tie-break on both sides, visit counts, float64 Q,
float32 priors, root N and tree reuse across moves must be IDENTICAL (BASELINE.json north_star)."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc

pytestmark = pytest.mark.gpu


def _rel(black, white, player):
    return np.where(player > 0, black, white), np.where(player > 0, white, black)


def _mk_tree(G, n, **kw):
    from alphazero_b200.engine import TreeBatch
    return TreeBatch(G, n=n, mode=kw.pop("mode", "neural"), tie_random=kw.pop("tie_random", False), **kw)


def _set_roots(tree, black, white, player):
    from alphazero_b200 import boardops as bo
    own, opp = _rel(black, white, player)
    tree.set_roots(bo.bits_from_numpy(own), bo.bits_from_numpy(opp), torch.from_numpy(player.astype(np.int8)).cuda())


@pytest.mark.parametrize("n,leaves", [(6, None), (8, None), (6, 1), (8, 1)])
def test_visit_counts_identical_to_reference_fixture(golden_dir, n, leaves):
    """leaves = 1: the throughput-mode kernel (k_tree_step_multi) with one leaf in flight must degenerate to the
    reference exactly (mcts.py:127-171, 197-223) -- same fixtures, same bits."""
    from alphazero_b200.evaluators import HashEvaluator
    d = np.load(os.path.join(golden_dir, f"mct_visits_n{n}.npz"))
    salt = int(d["salt"][0])
    for n_sim in sorted(set(d["n_sim"].tolist())):
        rows = np.where(d["n_sim"] == n_sim)[0]
        roots = sorted(set(d["root"][rows].tolist()))
        G = len(roots)
        tree = _mk_tree(G, n, n_sim_hint=n_sim, leaves=leaves)
        ev = HashEvaluator(salt)
        first = [rows[(d["root"][rows] == r) & (d["step"][rows] == 0)][0] for r in roots]
        _set_roots(tree, d["black"][first], d["white"][first], d["player"][first])
        max_step = int(d["step"][rows].max())
        for step in range(max_step + 1):
            idx = [rows[(d["root"][rows] == r) & (d["step"][rows] == step)] for r in roots]
            act = torch.tensor([1 if len(i) else 0 for i in idx], dtype=torch.uint8, device="cuda")
            tree.active.copy_(act)
            tree.search(n_sim, ev)
            N, Q, P, rn = [x.cpu().numpy() for x in tree.root_stats()]
            action, pi = tree.choose(None)
            action = action.cpu().numpy()
            adv = np.full(G, 255, np.uint8)
            for g, i in enumerate(idx):
                if not len(i):
                    continue
                j = int(i[0])
                assert (N[g] == d["N"][j]).all(), (n_sim, step, g, N[g][N[g] >= 0], d["N"][j][d["N"][j] >= 0])
                assert (Q[g] == d["Q"][j]).all()
                assert (P[g] == d["P"][j]).all()
                assert rn[g] == d["rootN"][j]
                assert action[g] == d["action"][j]
                adv[g] = d["action"][j]
            tree.advance(torch.from_numpy(adv).cuda())
            tree.check_status()


@pytest.mark.parametrize("n,n_sim,G,graph,leaves", [
    (8, 1, 256, False, None), (8, 2, 256, True, None), (8, 50, 512, True, None), (8, 50, 512, False, None),
    (8, 200, 256, True, None), (8, 400, 64, False, None), (6, 64, 256, True, None), (4, 40, 128, False, None),
    # throughput mode: L leaves in flight per game and step under virtual loss, against the oracle's sequential
    # restatement of the same rules (oracle/az_oracle.c: orc_search_multi)
    (8, 50, 256, True, 1), (8, 50, 256, True, 2), (8, 200, 128, True, 4), (8, 7, 128, False, 4), (6, 64, 128, False, 4),
    (4, 40, 64, False, 3), (8, 3, 64, True, 8)])
def test_visit_counts_identical_to_oracle(n, n_sim, G, graph, leaves):
    """>= 1000 roots overall: random mid-game roots, several consecutive searches with tree reuse, then an
    unexplored opponent move (fresh root) -- compared node statistics must be bit-identical.  graph=True: the
    evaluator is bound to the tree, so the searches from the second on replay the whole-search CUDA graph."""
    from alphazero_b200.evaluators import HashEvaluator
    rng = np.random.default_rng(n * 1000 + n_sim)
    salt = 4242 + n_sim
    b0, w0 = orc.start(n)
    black, white, player = np.full(G, b0, np.uint64), np.full(G, w0, np.uint64), np.ones(G, np.int8)
    depth = rng.integers(0, n * n - 8, size=G)
    depth[0] = 0
    for ply in range(int(depth.max())):
        legal = orc.legal(black, white, player, n)
        act = np.full(G, 255, np.uint8)
        for g in range(G):
            if ply < depth[g]:
                m = int(legal[g]); bits = [i for i in range(64) if (m >> i) & 1] or [64]
                act[g] = bits[rng.integers(len(bits))]
        term, _, _ = orc.terminal(black, white, player, n)
        go = (act != 255) & (term == 0)
        ob, ow, op, _, st = orc.apply(black, white, player, np.where(go, act, 64).astype(np.uint8), n)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    term, _, _ = orc.terminal(black, white, player, n)
    tree = _mk_tree(G, n, n_sim_hint=n_sim, leaves=leaves)
    ev = HashEvaluator(salt, tree.leaf_view if graph else None)
    _set_roots(tree, black, white, player)
    tree.active.copy_(torch.from_numpy((term == 0).astype(np.uint8)).cuda())
    oracles = [orc.OracleMCT(n, "neural", tie_random=False, salt=salt, leaves=leaves) for _ in range(G)]
    alive = term == 0
    for move in range(4):
        tree.search(n_sim, ev)
        N, Q, P, rn = [x.cpu().numpy() for x in tree.root_stats()]
        action, pi = tree.choose(None)
        action, pi = action.cpu().numpy(), pi.cpu().numpy()
        adv = np.full(G, 255, np.uint8)
        for g in range(G):
            if not alive[g]:
                continue
            o = oracles[g]
            o.search(black[g], white[g], player[g], n_sim)
            oN, oQ, oP, orn, _, _ = o.stats()
            assert (N[g] == oN).all(), (move, g, N[g][N[g] >= 0], oN[oN >= 0])
            assert (Q[g] == oQ).all() and (P[g] == oP.astype(np.float32)).all() and rn[g] == orn
            best = o.best_action()
            assert action[g] == best
            a_idx = n * n if best == 64 else (best // 8) * n + best % 8
            assert pi[g, a_idx] == 1.0 and pi[g].sum() == 1.0
            if move == 2:   # play a legal move the search did not favour: often an unexplored child / fresh root
                m = int(orc.legal(black[g:g + 1], white[g:g + 1], player[g:g + 1], n)[0])
                bits = [i for i in range(64) if (m >> i) & 1] or [64]
                best = bits[-1]
            adv[g] = best
            o.change_root(best)
        tree.advance(torch.from_numpy(adv).cuda())
        tree.check_status()
        go = adv != 255
        ob, ow, op, _, st = orc.apply(black, white, player, np.where(go, adv, 64).astype(np.uint8), n)
        assert (st[go] == 0).all()
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
        # root boards on the device follow
        from alphazero_b200 import boardops as bo
        own, opp = _rel(black, white, player)
        assert (bo.bits_to_numpy(tree.root_own)[alive] == own[alive]).all()
        assert (bo.bits_to_numpy(tree.root_opp)[alive] == opp[alive]).all()
        term, _, _ = orc.terminal(black, white, player, n)
        alive = alive & (term == 0)
        tree.active.copy_(torch.from_numpy(alive.astype(np.uint8)).cuda())
    assert any(e[0] is not None for e in tree._search_graphs.values()) == graph
    if (leaves or 1) == 1:
        assert ev.n_calls == 4 * (n_sim + 1)
    else:       # fewer, larger leaf batches
        assert ev.n_calls < 4 * (n_sim + 1) or n_sim <= 3


def test_advance_illegal_move_and_overflow_are_loud():
    from alphazero_b200.evaluators import HashEvaluator
    tree = _mk_tree(4, 8, n_sim_hint=10)
    tree.search(10, HashEvaluator(1))
    tree.advance(torch.tensor([0, 255, 255, 255], dtype=torch.uint8))  # (0,0) is not legal at the start
    with pytest.raises(ValueError):
        tree.check_status()
    small = _mk_tree(2, 8, capacity=64)
    with pytest.raises(RuntimeError):
        small.search(100, HashEvaluator(1))


def test_temperature_sampling_distribution():
    """temp = 1: the chosen move is distributed as N / sum(N) (players.py:185-189), pi equals that distribution."""
    from alphazero_b200.evaluators import HashEvaluator
    G = 4096
    tree = _mk_tree(G, 8, n_sim_hint=60, tie_random=True, seed=11)
    tree.search(60, HashEvaluator(77))
    N, _, _, _ = tree.root_stats()
    N = N.cpu().numpy()
    action, pi = tree.choose(torch.ones(G))
    action, pi = action.cpu().numpy(), pi.cpu().numpy()
    # all games share the root position and evaluator, but random tie-breaks make trees differ: check per game
    # that pi = N/sum(N) and overall that the empirical frequency of the chosen action matches the mean pi
    freq = np.zeros(65)
    mean_pi = np.zeros(65)
    for g in range(G):
        cnt = np.where(N[g] >= 0, N[g], 0).astype(np.float64)
        want = cnt / cnt.sum()
        assert np.allclose(pi[g, :64], want[:64].astype(np.float32), atol=1e-6)
        assert N[g][action[g]] > 0
        freq[action[g]] += 1
        mean_pi += want
    freq /= G; mean_pi /= G
    assert np.abs(freq - mean_pi).max() < 0.03


def test_search_graph_follows_dirichlet_settings_and_evaluator_swaps():
    """The whole-search CUDA graph bakes az_tree in by value: changing the Dirichlet settings or swapping the evaluator
    between searches must re-capture, never replay a stale graph (the reference reads mct.dirichlet_* and mct.nn live,
    mcts.py:78-83, 234-240)."""
    from alphazero_b200.evaluators import HashEvaluator
    G, n_sim = 64, 40

    def root_priors(tree, ev):
        tree.reset_roots()
        tree.search(n_sim, ev)
        return tree.root_stats()[2].clone()

    tree = _mk_tree(G, 8, n_sim_hint=n_sim, seed=3)
    ev = HashEvaluator(5, tree)
    plain = root_priors(tree, ev)
    assert torch.equal(root_priors(tree, ev), plain)            # second use: captured and replayed
    assert any(e[0] is not None for e in tree._search_graphs.values())
    tree.set_dirichlet(0.3, 0.25)
    noisy = root_priors(tree, ev)
    legal = plain > 0
    assert (noisy[legal] != plain[legal]).any()                  # the noise is ON in the third search
    assert torch.allclose(noisy.sum(1), torch.ones(G, device="cuda"), atol=1e-5)
    root_priors(tree, ev)                                        # captured with the noise on
    tree.set_dirichlet(None, None)
    assert torch.equal(root_priors(tree, ev), plain)            # and OFF again
    # a new evaluator object (different hash salt): its graph must not be confused with the dropped one's
    fresh = _mk_tree(G, 8, n_sim_hint=n_sim, seed=3)
    want = root_priors(fresh, HashEvaluator(6, fresh))
    del ev
    for _ in range(3):
        ev2 = HashEvaluator(6, tree)
        got = root_priors(tree, ev2)
        assert torch.equal(root_priors(tree, ev2), got)
        assert torch.equal(got, want)
        del ev2
