"""GPU: TicTacToe (the reference's third game) -- board kernels on ALL 5 478 reachable positions, device
perft, playouts, the tree kernels instantiated for TicTacToe, the fused TicTacToeNet kernel and the drop-in classes,
against the reference's fixtures (tests/golden/tictactoe_*.npz, mct_visits_ttt.npz) and the oracle."""
import json
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc

pytestmark = pytest.mark.gpu


def _dev(a):
    from alphazero_b200 import boardops
    return boardops.bits_from_numpy(a)


def _np(t):
    from alphazero_b200 import boardops
    return boardops.bits_to_numpy(t)


def _rel(black, white, player):
    return np.where(player > 0, black, white), np.where(player > 0, white, black)


def _spec():
    from alphazero_b200.gamespec import TicTacToeSpec
    return TicTacToeSpec()


def test_board_kernels_on_every_reachable_position(golden_dir):
    spec = _spec()
    d = np.load(os.path.join(golden_dir, "tictactoe_positions.npz"))
    player = d["player"].astype(np.int8)
    own, opp = _rel(d["black"], d["white"], player)
    t_own, t_opp, t_pl = _dev(own), _dev(opp), torch.from_numpy(player).cuda()
    assert (_np(spec.legal(t_own, t_opp)) == d["legal"]).all()
    term, score, winner = spec.terminal_score(t_own, t_opp, t_pl)
    assert (term.cpu().numpy() == d["over"]).all() and (score.cpu().numpy() == d["score"]).all()
    over = d["over"] == 1
    assert (winner.cpu().numpy()[over] == d["winner"][over]).all() and (winner.cpu().numpy()[~over] == 0).all()
    live = ~over
    act = np.where(live, d["action"], 0).astype(np.uint8)
    o2, p2, pl2, st = spec.apply(t_own, t_opp, t_pl, torch.from_numpy(act).cuda())
    o2, p2, pl2, st = _np(o2), _np(p2), pl2.cpu().numpy(), st.cpu().numpy()
    assert (st[live] == 0).all()
    nb, nw = np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2)
    assert (nb[live] == d["black2"][live]).all() and (nw[live] == d["white2"][live]).all()
    assert (pl2[live] == d["player2"][live]).all()
    for bad in (3, 11, 24, 63, 64, 255):            # not a cell: refused, board unchanged
        o3, p3, _, st3 = spec.apply(t_own, t_opp, t_pl, torch.full((len(own),), bad, dtype=torch.uint8, device="cuda"))
        assert (st3.cpu().numpy() == 1).all() and (_np(o3) == own).all() and (_np(p3) == opp).all()
    has = d["illegal"] != 255                       # occupied cell
    ill = np.where(has, d["illegal"], 0).astype(np.uint8)
    _, _, _, st4 = spec.apply(t_own, t_opp, t_pl, torch.from_numpy(ill).cuda())
    assert (st4.cpu().numpy()[has] == 1).all()
    enc = spec.encode(t_own, t_opp).cpu().numpy()
    for i in range(0, len(own), 97):
        g = np.array([[((int(own[i]) >> (r * 8 + c)) & 1) - ((int(opp[i]) >> (r * 8 + c)) & 1) for c in range(3)] for r in range(3)])
        assert (enc[i] == g).all()


def test_device_perft_by_frontier_expansion(golden_dir):
    """perft(1..9) with the legal / apply / terminal kernels driving a breadth-first frontier: 255 168 games."""
    spec = _spec()
    want = json.load(open(os.path.join(golden_dir, "tictactoe_misc.json")))["perft_1_to_9"]
    own = torch.zeros(1, dtype=torch.int64, device="cuda")
    opp = torch.zeros(1, dtype=torch.int64, device="cuda")
    player = torch.ones(1, dtype=torch.int8, device="cuda")
    finished = 0
    shifts = torch.arange(64, device="cuda")
    for depth in range(1, 10):
        term, _, _ = spec.terminal_score(own, opp, player)
        live = term == 0
        finished += int((~live).sum().item())
        own, opp, player = own[live], opp[live], player[live]
        legal = spec.legal(own, opp)
        bits = ((legal.unsqueeze(1) >> shifts) & 1).bool()
        idx, act = torch.nonzero(bits, as_tuple=True)
        o2, p2, pl2, st = spec.apply(own[idx], opp[idx], player[idx], act.to(torch.uint8))
        assert int(st.sum().item()) == 0
        own, opp, player = o2, p2, pl2
        assert finished + own.numel() == want[depth - 1], depth
    assert want[-1] == 255168


def test_playouts_vs_oracle():
    spec = _spec()
    z = np.zeros(1, np.uint64)
    total, seed = 200_000, 0x7717
    fb, fw, win, pl = spec.playouts(_dev(z), _dev(z), torch.ones(1, dtype=torch.int8, device="cuda"), total, seed)
    ofb, ofw, owin, opl = orc.ttt_playouts(z, z, np.ones(1, np.int8), 20_000, seed)
    assert (_np(fb)[:20_000] == ofb).all() and (_np(fw)[:20_000] == ofw).all()
    assert (win.cpu().numpy()[:20_000] == owin).all() and (pl.cpu().numpy()[:20_000] == opl).all()
    end_player = torch.where(pl % 2 == 0, 1, -1).to(torch.int8)
    own, opp = torch.where(end_player > 0, fb, fw), torch.where(end_player > 0, fw, fb)
    term, _, w2 = spec.terminal_score(own, opp, end_player)
    assert bool((term == 1).all()) and bool((w2 == win).all())
    frac = [(win == v).float().mean().item() for v in (1, -1, 0)]
    # exact values for uniformly random play: 131184 / 77904 / 46080 out of 255168 games weighted by probability
    assert abs(frac[0] - 0.585) < 0.01 and abs(frac[1] - 0.288) < 0.01 and abs(frac[2] - 0.127) < 0.01


def _mk_tree(G, **kw):
    from alphazero_b200.engine import TreeBatch
    return TreeBatch(G, n=_spec(), mode=kw.pop("mode", "neural"), tie_random=kw.pop("tie_random", False), **kw)


def _set_roots(tree, black, white, player):
    own, opp = _rel(black, white, player)
    tree.set_roots(_dev(own), _dev(opp), torch.from_numpy(player.astype(np.int8)).cuda())


def test_visit_counts_identical_to_reference_fixture(golden_dir):
    from alphazero_b200.evaluators import HashEvaluator
    d = np.load(os.path.join(golden_dir, "mct_visits_ttt.npz"))
    salt = int(d["salt"][0])
    for n_sim in sorted(set(d["n_sim"].tolist())):
        rows = np.where(d["n_sim"] == n_sim)[0]
        roots = sorted(set(d["root"][rows].tolist()))
        G = len(roots)
        tree = _mk_tree(G, n_sim_hint=n_sim)
        ev = HashEvaluator(salt)
        first = [rows[(d["root"][rows] == r) & (d["step"][rows] == 0)][0] for r in roots]
        _set_roots(tree, d["black"][first], d["white"][first], d["player"][first])
        for step in range(int(d["step"][rows].max()) + 1):
            idx = [rows[(d["root"][rows] == r) & (d["step"][rows] == step)] for r in roots]
            tree.active.copy_(torch.tensor([1 if len(i) else 0 for i in idx], dtype=torch.uint8, device="cuda"))
            tree.search(n_sim, ev)
            N, Q, P, rn = [x.cpu().numpy() for x in tree.root_stats()]
            action, pi = tree.choose(None)
            action, pi = action.cpu().numpy(), pi.cpu().numpy()
            assert pi.shape == (G, 9)
            adv = np.full(G, 255, np.uint8)
            for g, i in enumerate(idx):
                if not len(i):
                    continue
                j = int(i[0])
                assert (N[g] == d["N"][j]).all(), (n_sim, step, g, N[g][N[g] >= 0], d["N"][j][d["N"][j] >= 0])
                assert (Q[g] == d["Q"][j]).all() and (P[g] == d["P"][j]).all() and rn[g] == d["rootN"][j]
                assert action[g] == d["action"][j]
                a = int(action[g])
                assert pi[g, 3 * (a // 8) + a % 8] == 1.0 and pi[g].sum() == 1.0     # network order 3r + c
                adv[g] = d["action"][j]
            tree.advance(torch.from_numpy(adv).cuda())
            tree.check_status()


@pytest.mark.parametrize("n_sim,G", [(1, 128), (30, 512), (200, 256), (600, 64)])
def test_visit_counts_identical_to_oracle(n_sim, G):
    """Random roots, consecutive searches with tree reuse down to the end of the game (the whole game tree fits in a
    600-simulation search: terminal leaves are re-visited), bit-identical stats."""
    from alphazero_b200.evaluators import HashEvaluator
    rng = np.random.default_rng(n_sim)
    salt = 4242 + n_sim
    black, white, player = np.zeros(G, np.uint64), np.zeros(G, np.uint64), np.ones(G, np.int8)
    depth = rng.integers(0, 5, size=G)
    depth[0] = 0
    for ply in range(int(depth.max())):
        legal = orc.ttt_legal(black, white, player)
        term, _, _ = orc.ttt_terminal(black, white, player)
        go = (ply < depth) & (term == 0)
        act = np.zeros(G, np.uint8)
        for g in np.nonzero(go)[0]:
            ok = [a for a in range(64) if (int(legal[g]) >> a) & 1]
            act[g] = ok[rng.integers(len(ok))]
        ob, ow, op, st = orc.ttt_apply(black, white, player, act)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    term, _, _ = orc.ttt_terminal(black, white, player)
    tree = _mk_tree(G, n_sim_hint=n_sim)
    ev = HashEvaluator(salt)
    _set_roots(tree, black, white, player)
    alive = term == 0
    tree.active.copy_(torch.from_numpy(alive.astype(np.uint8)).cuda())
    oracles = [orc.OracleMCT(3, "neural", tie_random=False, salt=salt, game="tictactoe") for _ in range(G)]
    for move in range(6):
        if not alive.any():
            break
        tree.search(n_sim, ev)
        N, Q, P, rn = [x.cpu().numpy() for x in tree.root_stats()]
        action, _ = tree.choose(None)
        action = action.cpu().numpy()
        adv = np.full(G, 255, np.uint8)
        legal = orc.ttt_legal(black, white, player)
        for g in np.nonzero(alive)[0]:
            o = oracles[g]
            o.search(black[g], white[g], player[g], n_sim)
            oN, oQ, oP, orn, _, _ = o.stats()
            assert (N[g] == oN).all(), (move, g, N[g][N[g] >= 0], oN[oN >= 0])
            assert (Q[g] == oQ).all() and (P[g] == oP.astype(np.float32)).all() and rn[g] == orn
            best = o.best_action()
            assert action[g] == best
            if move == 1:
                best = [a for a in range(64) if (int(legal[g]) >> a) & 1][-1]
            adv[g] = best
            o.change_root(best)
        tree.advance(torch.from_numpy(adv).cuda())
        tree.check_status()
        go = adv != 255
        ob, ow, op, st = orc.ttt_apply(black, white, player, np.where(go, adv, 0).astype(np.uint8))
        assert (st[go] == 0).all()
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
        term, _, _ = orc.ttt_terminal(black, white, player)
        alive = alive & (term == 0)
        tree.active.copy_(torch.from_numpy(alive.astype(np.uint8)).cuda())


def test_rollout_mode_takes_the_win_and_blocks():
    G = 64
    b, w_, pl = np.zeros(1, np.uint64), np.zeros(1, np.uint64), np.ones(1, np.int8)
    for a in (0, 8, 1, 9):              # X: (0,0), (0,1); O: (1,0), (1,1); X to move wins at (0,2) = action 2
        b, w_, pl, st = orc.ttt_apply(b, w_, pl, np.array([a], np.uint8))
    tree = _mk_tree(G, mode="rollout", tie_random=True, n_sim_hint=200, seed=5)
    _set_roots(tree, np.repeat(b, G), np.repeat(w_, G), np.repeat(pl, G))
    tree.search(200)
    action, _ = tree.choose(None)
    assert (action.cpu().numpy() == 2).mean() > 0.9
    b2, w2, pl2, st = orc.ttt_apply(b, w_, pl, np.array([18], np.uint8))     # X plays (2,2) instead; O must win at (1,2) = 10
    _set_roots(tree, np.repeat(b2, G), np.repeat(w2, G), np.repeat(pl2, G))
    tree.search(200)
    action, _ = tree.choose(None)
    assert (action.cpu().numpy() == 10).mean() > 0.9


def _golden_net(golden_dir):
    from alphazero_b200.games.tictactoe import TicTacToeNet
    d = np.load(os.path.join(golden_dir, "tictactoe_net.npz"))
    net = TicTacToeNet(device="cpu")
    net.load_state_dict({k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith("sd.")})
    return net.eval(), d


def test_fused_net_matches_reference_outputs(golden_dir):
    """az_tictactoe_net on every live position vs the reference's fp32 predict (fp32 kernel: 1e-5, far inside the
    1e-2 of the north star)."""
    from alphazero_b200.fused_net import TicTacToeNetEvaluator, make_evaluator
    net, d = _golden_net(golden_dir)
    pos = np.load(os.path.join(golden_dir, "tictactoe_positions.npz"))
    idx = d["index"]
    player = pos["player"][idx].astype(np.int8)
    own, opp = _rel(pos["black"][idx], pos["white"][idx], player)
    G = len(idx)
    tree = _mk_tree(G)
    ev = make_evaluator(net, tree)
    assert isinstance(ev, TicTacToeNetEvaluator)
    tree.leaf_own.copy_(_dev(own)); tree.leaf_opp.copy_(_dev(opp))
    ev()
    assert np.abs(tree.probs.cpu().numpy() - d["probs"]).max() < 1e-5
    assert np.abs(tree.value.cpu().numpy() - d["value"].reshape(-1)).max() < 1e-5
    # the spec's encoding is the reference's network input
    enc = _spec().encode(_dev(own), _dev(opp)).cpu().numpy()
    assert (enc == d["input"]).all()


def test_board_and_players_drop_in(golden_dir):
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.tictactoe import TicTacToeBoard, TicTacToeNet
    from alphazero_b200.players import AlphaZeroPlayer, GreedyPlayer, MCTSPlayer, RandomPlayer
    d = np.load(os.path.join(golden_dir, "tictactoe_positions.npz"))
    for i in range(0, len(d["black"]), 61):
        grid = np.array([[((int(d["black"][i]) >> (r * 8 + c)) & 1) - ((int(d["white"][i]) >> (r * 8 + c)) & 1)
                          for c in range(3)] for r in range(3)], dtype=float)
        b = TicTacToeBoard(grid=grid, player=int(d["player"][i]))
        legal = int(d["legal"][i])
        assert b.get_moves() == [(a // 8, a % 8) for a in range(64) if (legal >> a) & 1]       # row-major order
        assert b.is_game_over() == bool(d["over"][i])
        assert b.get_score() == (float("inf") if d["score"][i] == 127 else 0)
        if d["over"][i]:
            assert b.get_winner() == d["winner"][i]
        else:
            with pytest.raises(ValueError):
                b.get_winner()
            c = b.clone()
            a = int(d["action"][i])
            c.play_move((a // 8, a % 8))
            assert c.player == d["player2"][i] and (b.grid == grid).all()
            with pytest.raises(ValueError):
                c.play_move((a // 8, a % 8))
        assert not b.is_legal_move((3, 0)) and str(b) == "TicTacToeBoard"
    np.random.seed(1)
    torch.manual_seed(1)
    # greedy takes an immediate win; MCTS (rollout) never loses to random often; AlphaZero player plays legal games
    stats = Arena(MCTSPlayer(n_sim=60), RandomPlayer(), TicTacToeBoard()).play_games_in_parallel(200, n_process=2, return_stats=True)
    assert len(stats["player1"]) > 150 and len(stats["player2"]) < 20
    stats = Arena(AlphaZeroPlayer(n_sim=20, nn=TicTacToeNet()), GreedyPlayer(), TicTacToeBoard()).play_games(4, return_stats=True)
    assert len(stats["player1"]) + len(stats["player2"]) + stats["draw"] == 4
