"""GPU: Connect4 (BASELINE.json config 5) -- board kernels, playouts, perft and the tree kernels instantiated for
Connect4, all through the C ABI, bit-exact against the reference fixtures (tests/golden/connect4_*.npz,
mct_visits_c4_*.npz) and the oracle's cell-grid restatement (oracle/az_oracle.c)."""
import json
import os

import numpy as np
import pytest
import torch

import oracle_lib as orc

pytestmark = pytest.mark.gpu

SIZES = [(7, 6), (4, 4), (8, 8), (5, 7)]


def _dev(a):
    from alphazero_b200 import boardops
    return boardops.bits_from_numpy(a)


def _np(t):
    from alphazero_b200 import boardops
    return boardops.bits_to_numpy(t)


def _rel(black, white, player):
    return np.where(player > 0, black, white), np.where(player > 0, white, black)


def _spec(w, h):
    from alphazero_b200.gamespec import Connect4Spec
    return Connect4Spec(w, h)


@pytest.mark.parametrize("w,h", SIZES)
def test_board_kernels_vs_reference_fixtures(golden_dir, w, h):
    spec = _spec(w, h)
    d = np.load(os.path.join(golden_dir, f"connect4_positions_{w}x{h}.npz"))
    player = d["player"].astype(np.int8)
    own, opp = _rel(d["black"], d["white"], player)
    t_own, t_opp, t_pl = _dev(own), _dev(opp), torch.from_numpy(player).cuda()
    assert (_np(spec.legal(t_own, t_opp)) == d["legal"]).all()
    term, score, winner = spec.terminal_score(t_own, t_opp, t_pl)
    assert (term.cpu().numpy() == d["over"]).all()
    assert (score.cpu().numpy() == d["score"]).all()
    over = d["over"] == 1
    assert (winner.cpu().numpy()[over] == d["winner"][over]).all()
    live = ~over
    act = np.where(live, d["action"], 0).astype(np.uint8)
    o2, p2, pl2, st = spec.apply(t_own, t_opp, t_pl, torch.from_numpy(act).cuda())
    o2, p2, pl2, st = _np(o2), _np(p2), pl2.cpu().numpy(), st.cpu().numpy()
    assert (st[live] == 0).all()
    nb, nw = np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2)
    assert (nb[live] == d["black2"][live]).all() and (nw[live] == d["white2"][live]).all()
    assert (pl2[live] == d["player2"][live]).all()
    # out-of-range column: refused, board unchanged (the reference raises ValueError)
    o3, p3, pl3, st3 = spec.apply(t_own, t_opp, t_pl, torch.full((len(own),), w, dtype=torch.uint8, device="cuda"))
    assert (st3.cpu().numpy() == 1).all() and (_np(o3) == own).all() and (_np(p3) == opp).all()
    assert (pl3.cpu().numpy() == player).all()
    # encode = player * grid, row-major (height, width)
    enc = spec.encode(t_own, t_opp).cpu().numpy()
    assert enc.shape == (len(own), h, w)
    for g in range(0, len(own), max(1, len(own) // 60)):
        for r in range(h):
            for c in range(w):
                want = ((int(own[g]) >> (r * 8 + c)) & 1) - ((int(opp[g]) >> (r * 8 + c)) & 1)
                assert enc[g, r, c] == want
    assert (spec.encode(t_own, t_opp, dtype=torch.bfloat16).float().cpu().numpy() == enc).all()


@pytest.mark.parametrize("w,h", [(7, 6), (8, 8), (4, 5)])
def test_board_kernels_vs_oracle_random_walk(w, h):
    """Every position of 6000 random games: legal columns, terminal flag, winner, score and the drop compared with
    the oracle, plus refusal of full / out-of-range columns."""
    spec = _spec(w, h)
    rng = np.random.default_rng(100 * w + h)
    G = 6000
    black, white, player = np.zeros(G, np.uint64), np.zeros(G, np.uint64), np.ones(G, np.int8)
    n_pos = 0
    for ply in range(w * h + 1):
        own, opp = _rel(black, white, player)
        t_own, t_opp, t_pl = _dev(own), _dev(opp), torch.from_numpy(player).cuda()
        legal = orc.c4_legal(black, white, player, w, h)
        assert (_np(spec.legal(t_own, t_opp)) == legal).all()
        term, score, winner = orc.c4_terminal(black, white, player, w, h)
        d_term, d_score, d_win = spec.terminal_score(t_own, t_opp, t_pl)
        assert (d_term.cpu().numpy() == term).all() and (d_score.cpu().numpy() == score).all()
        assert (d_win.cpu().numpy()[term == 1] == winner[term == 1]).all()
        live = term == 0
        n_pos += int(live.sum())
        # mostly a random playable column, sometimes an arbitrary (possibly full / out-of-range) one
        cols = rng.integers(0, w + 2, size=G)
        pick = rng.random(G) < 0.9
        for g in np.nonzero(pick & live)[0]:
            ok = [c for c in range(w) if (int(legal[g]) >> c) & 1]
            cols[g] = ok[rng.integers(len(ok))]
        act = cols.astype(np.uint8)
        ob, ow, op, st = orc.c4_apply(black, white, player, act, w, h)
        o2, p2, pl2, st2 = spec.apply(t_own, t_opp, t_pl, torch.from_numpy(act).cuda())
        o2, p2, pl2, st2 = _np(o2), _np(p2), pl2.cpu().numpy(), st2.cpu().numpy()
        assert (st2 == st).all()
        nb, nw = np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2)
        assert (nb == ob).all() and (nw == ow).all() and (pl2 == op).all()
        go = live & (st == 0)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    assert n_pos > G * 10


def test_perft_vs_reference_and_oracle(golden_dir):
    t = json.load(open(os.path.join(golden_dir, "perft_connect4.json")))
    for key in ("7x6", "4x4", "5x4"):
        w, h = (int(x) for x in key.split("x"))
        spec = _spec(w, h)
        assert [spec.perft(d) for d in range(1, len(t[key]) + 1)] == t[key]
    spec = _spec(7, 6)
    assert spec.perft(7) == 7 ** 7 - 7
    assert spec.perft(8) == orc.c4_perft(7, 6, 8)
    # BFS/DFS split must not matter
    assert spec.perft(6, dfs_depth=0) == spec.perft(6, dfs_depth=6) == 7 ** 6
    # a mid-game root
    b, w_ = np.zeros(1, np.uint64), np.zeros(1, np.uint64)
    pl = np.ones(1, np.int8)
    for a in (3, 3, 2, 4, 1):
        b, w_, pl, st = orc.c4_apply(b, w_, pl, np.array([a], np.uint8), 7, 6)
    own, opp = _rel(b, w_, pl)
    assert spec.perft(5, own=int(own[0]), opp=int(opp[0])) == orc.c4_perft(7, 6, 5, int(b[0]), int(w_[0]), int(pl[0]))


def test_playouts_vs_oracle_and_properties():
    """20k playouts bit-exact with the oracle (same Philox stream); 1M playouts through size-independent
    properties: every final position is terminal, the winner matches it, plies = discs on the board."""
    spec = _spec(7, 6)
    rng = np.random.default_rng(5)
    R = 200
    black, white, player = np.zeros(R, np.uint64), np.zeros(R, np.uint64), np.ones(R, np.int8)
    for ply in range(12):       # roots after 0..11 random plies
        legal = orc.c4_legal(black, white, player, 7, 6)
        term, _, _ = orc.c4_terminal(black, white, player, 7, 6)
        go = (term == 0) & (np.arange(R) % 12 > ply)
        act = np.zeros(R, np.uint8)
        for g in np.nonzero(go)[0]:
            ok = [c for c in range(7) if (int(legal[g]) >> c) & 1]
            act[g] = ok[rng.integers(len(ok))]
        ob, ow, op, st = orc.c4_apply(black, white, player, act, 7, 6)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    own, opp = _rel(black, white, player)
    fb, fw, win, pl = spec.playouts(_dev(own), _dev(opp), torch.from_numpy(player).cuda(), 100, seed=0xC4, first_id=17)
    ofb, ofw, owin, opl = orc.c4_playouts(black, white, player, 7, 6, 100, 0xC4, 17)
    assert (_np(fb) == ofb).all() and (_np(fw) == ofw).all()
    assert (win.cpu().numpy() == owin).all() and (pl.cpu().numpy() == opl).all()
    # 1M playouts from the empty board
    z = torch.zeros(1, dtype=torch.int64, device="cuda")
    T = 1_000_000
    fb, fw, win, pl = spec.playouts(z, z, torch.ones(1, dtype=torch.int8, device="cuda"), T, seed=0x5EED)
    discs = torch.tensor([bin(x).count("1") for x in range(256)], dtype=torch.int32, device="cuda")

    def popc(t):
        b = t.view(torch.uint8).reshape(-1, 8).to(torch.int64)
        return discs[b].sum(dim=1)

    nb, nw = popc(fb), popc(fw)
    assert bool(((nb + nw) == pl.to(torch.int32)).all())
    assert bool(((nb - nw) == (pl % 2).to(torch.int32)).all())           # black moves first
    end_player = torch.where(pl % 2 == 0, 1, -1).to(torch.int8)
    own_t = torch.where(end_player > 0, fb, fw); opp_t = torch.where(end_player > 0, fw, fb)
    term, _, w2 = spec.terminal_score(own_t, opp_t, end_player)
    assert bool((term == 1).all()) and bool((w2 == win).all())
    # the winner is the side that moved last unless the board is full without a line
    last = -end_player
    assert bool(((win == last) | ((win == 0) & (pl == 42))).all())
    frac = [(win == v).float().mean().item() for v in (1, -1, 0)]
    assert 0.52 < frac[0] < 0.60 and 0.40 < frac[1] < 0.47 and frac[2] < 0.01   # first-player edge of random play


# ------------------------------------------------------------------------------------------------ tree kernels

def _mk_tree(G, w, h, **kw):
    from alphazero_b200.engine import TreeBatch
    return TreeBatch(G, n=_spec(w, h), mode=kw.pop("mode", "neural"), tie_random=kw.pop("tie_random", False), **kw)


def _set_roots(tree, black, white, player):
    own, opp = _rel(black, white, player)
    tree.set_roots(_dev(own), _dev(opp), torch.from_numpy(player.astype(np.int8)).cuda())


@pytest.mark.parametrize("w,h", [(7, 6), (5, 4)])
def test_visit_counts_identical_to_reference_fixture(golden_dir, w, h):
    from alphazero_b200.evaluators import HashEvaluator
    d = np.load(os.path.join(golden_dir, f"mct_visits_c4_{w}x{h}.npz"))
    salt = int(d["salt"][0])
    for n_sim in sorted(set(d["n_sim"].tolist())):
        rows = np.where(d["n_sim"] == n_sim)[0]
        roots = sorted(set(d["root"][rows].tolist()))
        G = len(roots)
        tree = _mk_tree(G, w, h, n_sim_hint=n_sim)
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
            assert pi.shape == (G, w)
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
                assert pi[g, action[g]] == 1.0 and pi[g].sum() == 1.0
                adv[g] = d["action"][j]
            tree.advance(torch.from_numpy(adv).cuda())
            tree.check_status()


@pytest.mark.parametrize("w,h,n_sim,G", [(7, 6, 1, 256), (7, 6, 50, 512), (7, 6, 200, 256), (7, 6, 400, 64), (8, 8, 64, 128),
                                         (4, 4, 40, 128)])
def test_visit_counts_identical_to_oracle(w, h, n_sim, G):
    """Random mid-game roots, 4 consecutive searches with tree reuse, one unexplored re-root: bit-identical stats."""
    from alphazero_b200.evaluators import HashEvaluator
    rng = np.random.default_rng(w * 1000 + h * 100 + n_sim)
    salt = 777 + n_sim
    black, white, player = np.zeros(G, np.uint64), np.zeros(G, np.uint64), np.ones(G, np.int8)
    depth = rng.integers(0, w * h - 6, size=G)
    depth[0] = 0
    for ply in range(int(depth.max())):
        legal = orc.c4_legal(black, white, player, w, h)
        term, _, _ = orc.c4_terminal(black, white, player, w, h)
        go = (ply < depth) & (term == 0)
        act = np.zeros(G, np.uint8)
        for g in np.nonzero(go)[0]:
            ok = [c for c in range(w) if (int(legal[g]) >> c) & 1]
            act[g] = ok[rng.integers(len(ok))]
        ob, ow, op, st = orc.c4_apply(black, white, player, act, w, h)
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
    term, _, _ = orc.c4_terminal(black, white, player, w, h)
    tree = _mk_tree(G, w, h, n_sim_hint=n_sim)
    ev = HashEvaluator(salt)
    _set_roots(tree, black, white, player)
    alive = term == 0
    tree.active.copy_(torch.from_numpy(alive.astype(np.uint8)).cuda())
    oracles = [orc.OracleMCT((w, h), "neural", tie_random=False, salt=salt, game="connect4") for _ in range(G)]
    for move in range(4):
        tree.search(n_sim, ev)
        N, Q, P, rn = [x.cpu().numpy() for x in tree.root_stats()]
        action, pi = tree.choose(None)
        action = action.cpu().numpy()
        adv = np.full(G, 255, np.uint8)
        legal = orc.c4_legal(black, white, player, w, h)
        for g in np.nonzero(alive)[0]:
            o = oracles[g]
            o.search(black[g], white[g], player[g], n_sim)
            oN, oQ, oP, orn, _, _ = o.stats()
            assert (N[g] == oN).all(), (move, g, N[g][N[g] >= 0], oN[oN >= 0])
            assert (Q[g] == oQ).all() and (P[g] == oP.astype(np.float32)).all() and rn[g] == orn
            best = o.best_action()
            assert action[g] == best
            if move == 2:       # a legal column the search did not favour: often an unexplored child / fresh root
                best = [c for c in range(w) if (int(legal[g]) >> c) & 1][-1]
            adv[g] = best
            o.change_root(best)
        tree.advance(torch.from_numpy(adv).cuda())
        tree.check_status()
        go = adv != 255
        ob, ow, op, st = orc.c4_apply(black, white, player, np.where(go, adv, 0).astype(np.uint8), w, h)
        assert (st[go] == 0).all()
        black, white, player = np.where(go, ob, black), np.where(go, ow, white), np.where(go, op, player)
        own, opp = _rel(black, white, player)
        assert (_np(tree.root_own)[alive] == own[alive]).all() and (_np(tree.root_opp)[alive] == opp[alive]).all()
        term, _, _ = orc.c4_terminal(black, white, player, w, h)
        alive = alive & (term == 0)
        tree.active.copy_(torch.from_numpy(alive.astype(np.uint8)).cuda())


def test_rollout_mode_prefers_the_winning_column():
    """UCT + random playouts (MCTSPlayer): with three in a row on the bottom line the search must find the win, and
    the opponent's search must find the block -- distribution-level check of rollout mode on Connect4."""
    G = 64
    b, w_, pl = np.zeros(1, np.uint64), np.zeros(1, np.uint64), np.ones(1, np.int8)
    for a in (0, 0, 1, 1, 2):           # black: cols 0,1,2 on the bottom row; white to move must block col 3
        b, w_, pl, st = orc.c4_apply(b, w_, pl, np.array([a], np.uint8), 7, 6)
    tree = _mk_tree(G, 7, 6, mode="rollout", tie_random=True, n_sim_hint=300, seed=3)
    _set_roots(tree, np.repeat(b, G), np.repeat(w_, G), np.repeat(pl, G))
    tree.search(300)
    action, _ = tree.choose(None)
    assert (action.cpu().numpy() == 3).mean() > 0.9
    b2, w2, pl2, st = orc.c4_apply(b, w_, pl, np.array([6], np.uint8), 7, 6)     # white wastes the move
    _set_roots(tree, np.repeat(b2, G), np.repeat(w2, G), np.repeat(pl2, G))
    tree.search(300)
    action, _ = tree.choose(None)
    assert (action.cpu().numpy() == 3).mean() > 0.9


# ------------------------------------------------------------------------------------------------ network

def _golden_net(golden_dir):
    from alphazero_b200.games.connect4 import Connect4Net
    d = np.load(os.path.join(golden_dir, "connect4_net_7x6.npz"))
    net = Connect4Net(board_width=7, board_height=6, device="cpu")
    sd = {k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith("sd.")}
    net.load_state_dict(sd)            # the reference's parameter names load as they are
    return net.eval(), d


def test_connect4_net_matches_reference_outputs(golden_dir):
    """fp32 on the device from the kernel-made encoding == the reference's fp32 predict (same weights), and the
    bf16 CUDA-graph evaluator used by the search stays within 1e-2 of it (north_star tolerance)."""
    from alphazero_b200.evaluators import NetEvaluator
    net, d = _golden_net(golden_dir)
    spec = _spec(7, 6)
    player = d["player"].astype(np.int8)
    own, opp = _rel(d["black"], d["white"], player)
    t_own, t_opp = _dev(own), _dev(opp)
    enc = spec.encode(t_own, t_opp)
    assert (enc.cpu().numpy().reshape(len(own), -1) == d["input"].reshape(len(own), -1)).all()
    net32 = net.to("cuda"); net32.device = torch.device("cuda")
    probs, v = net32.predict(enc)
    assert np.abs(probs.cpu().numpy() - d["probs"]).max() < 1e-5      # tolerance: fp32 conv algorithm differences
    assert np.abs(v.cpu().numpy() - d["value"]).max() < 1e-5
    G = len(own)
    tree = _mk_tree(G, 7, 6, enc_dtype=torch.bfloat16)
    tree.enc.copy_(enc.to(torch.bfloat16))
    ev = NetEvaluator(net, tree, dtype=torch.bfloat16)
    tree.enc.copy_(enc.to(torch.bfloat16))
    ev(tree)
    assert np.abs(tree.probs.cpu().numpy() - d["probs"]).max() < 1e-2  # tolerance stated by BASELINE.json north_star
    assert np.abs(tree.value.cpu().numpy() - d["value"].reshape(-1)).max() < 1e-2


# ------------------------------------------------------------------------------------------------ drop-in layer

def _bits_to_grid(black, white, w, h):
    g = np.zeros((h, w))
    for r in range(h):
        for c in range(w):
            if (black >> (r * 8 + c)) & 1:
                g[r, c] = 1
            elif (white >> (r * 8 + c)) & 1:
                g[r, c] = -1
    return g


@pytest.mark.parametrize("w,h", [(7, 6), (5, 7)])
def test_connect4_board_api_vs_reference_fixture(golden_dir, w, h):
    from alphazero_b200.games.connect4 import Connect4Board
    d = np.load(os.path.join(golden_dir, f"connect4_positions_{w}x{h}.npz"))
    for i in range(0, len(d["black"]), max(1, len(d["black"]) // 120)):
        grid = _bits_to_grid(int(d["black"][i]), int(d["white"][i]), w, h)
        b = Connect4Board(width=w, height=h, grid=grid, player=int(d["player"][i]))
        assert (b.grid == grid).all() and b.get_board_shape() == (h, w) and b.get_action_size() == w
        assert (b.free_rows == (grid == 0).sum(axis=0)).all()
        legal = int(d["legal"][i])
        want = [c for c in range(w) if (legal >> c) & 1]
        moves = b.get_moves()
        assert [int(m) for m in moves] == want and all(isinstance(m, np.int64) for m in moves)   # ascending columns
        assert b.is_game_over() == bool(d["over"][i]) and b.get_score() == d["score"][i]
        with pytest.raises(ValueError):
            b.is_legal_move(w)
        with pytest.raises(ValueError):
            b.play_move(-1)
        for c in range(w):
            assert b.is_legal_move(c) == (c in want)
            if c not in want:
                with pytest.raises(ValueError):
                    b.clone().play_move(c)
        if d["over"][i]:
            assert b.get_winner() == d["winner"][i]
            continue
        with pytest.raises(ValueError):
            b.get_winner()
        assert int(b.get_random_move()) in want
        c2 = b.clone()
        c2.play_move(int(d["action"][i]))
        assert (c2.grid == _bits_to_grid(int(d["black2"][i]), int(d["white2"][i]), w, h)).all()
        assert c2.player == d["player2"][i] and b.player == d["player"][i]
    with pytest.raises(ValueError):
        Connect4Board(width=3, height=6)
    b = Connect4Board(width=7, height=6)
    assert str(b) == "Connect4Board7x6" and b.max_moves == 42 and b.pass_move is None and b.game == "connect4"


class _HashNet:
    def __init__(self, salt):
        self.salt = salt

    def make_batched_evaluator(self, tree):
        from alphazero_b200.evaluators import HashEvaluator
        return HashEvaluator(self.salt)


def test_mct_facade_identical_to_reference_fixture(golden_dir):
    from alphazero_b200.games.connect4 import Connect4Board
    from alphazero_b200.mcts import MCT
    w, h = 7, 6
    d = np.load(os.path.join(golden_dir, f"mct_visits_c4_{w}x{h}.npz"))
    salt = int(d["salt"][0])
    MCT.TIE_RANDOM = False
    try:
        rows, i, checked = len(d["root"]), 0, 0
        while i < rows and checked < 30:
            mct = MCT(eval_method="neural", nn=_HashNet(salt))
            j = i
            while j < rows and d["root"][j] == d["root"][i] and d["n_sim"][j] == d["n_sim"][i] and (j == i or d["step"][j] > d["step"][j - 1]):
                board = Connect4Board(width=w, height=h, grid=_bits_to_grid(int(d["black"][j]), int(d["white"][j]), w, h),
                                      player=int(d["player"][j]))
                mct.search(board, n_sim=int(d["n_sim"][j]))
                root = mct.root
                assert root.N == d["rootN"][j]
                for a in range(w):
                    if d["N"][j][a] < 0:
                        assert a not in root.children
                    else:
                        ch = root.children[a]
                        assert ch.N == d["N"][j][a] and ch.Q == d["Q"][j][a] and ch.P == d["P"][j][a]
                onehot, visits = mct.get_action_probs(board, temp=0)
                assert list(onehot.values()) == [1] and visits[list(onehot)[0]] == max(visits.values())
                mct.change_root(np.int64(d["action"][j]))
                j += 1
                checked += 1
            i = j
    finally:
        MCT.TIE_RANDOM = True


def test_players_arena_and_batched_arena_connect4():
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.connect4 import Connect4Board, Connect4Net
    from alphazero_b200.players import AlphaZeroPlayer, GreedyPlayer, MCTSPlayer, RandomPlayer
    np.random.seed(0)
    torch.manual_seed(0)
    board = Connect4Board(width=7, height=6)
    az = AlphaZeroPlayer(n_sim=16, nn=Connect4Net(board_width=7, board_height=6), dirichlet_alpha=0.03, dirichlet_epsilon=0.25)
    move, probs, visits, priors = az.get_move(board, temp=1)
    assert 0 <= move < 7 and abs(sum(probs.values()) - 1) < 1e-9 and sum(visits.values()) == 16
    assert abs(sum(priors.values()) - 1) < 1e-5
    res = Arena(GreedyPlayer(), az, Connect4Board(width=7, height=6)).play_game(player2_starts=True, return_results=True)
    assert res["winner"] in (0, 1, 2)
    stats = Arena(MCTSPlayer(n_sim=8), RandomPlayer(), Connect4Board(width=5, height=4)).play_games(4, return_stats=True)
    assert len(stats["player1"]) + len(stats["player2"]) + stats["draw"] == 4
    # batched "parallel" mode: rollout MCTS(50) beats Random on 7x6 nearly always; Random vs Random shows the
    # first-mover edge (~55.6 % of decided games for the starter), symmetric between the two players
    G = 400
    st = Arena(MCTSPlayer(n_sim=50), RandomPlayer(), Connect4Board(width=7, height=6)).play_games_in_parallel(
        G, n_process=2, return_stats=True, seed=1)
    p1, p2 = len(st["player1"]), len(st["player2"])
    assert p1 + p2 + st["draw"] == G and p1 / G > 0.9, st
    rr = Arena(RandomPlayer(), RandomPlayer(), Connect4Board(width=7, height=6)).play_games_in_parallel(
        4000, return_stats=True, seed=2)
    w1, w2 = len(rr["player1"]), len(rr["player2"])
    assert w1 + w2 + rr["draw"] == 4000 and abs(w1 - w2) / 4000 < 0.07
    starter_wins = rr["player1_starts"]["win"] + rr["player2_starts"]["win"]
    assert 0.52 < starter_wins / 4000 < 0.60


def test_self_play_engine_and_trainer_connect4():
    """BASELINE config 5 in small: Connect4 self-play through the same tree kernels; samples are consistent with
    replaying the game with the oracle, augmentation is the left-right mirror only (registers.py:42-45)."""
    from alphazero_b200.games.connect4 import Connect4Config, Connect4Net
    from alphazero_b200.gamespec import Connect4Spec
    from alphazero_b200.selfplay import SelfPlayEngine
    from alphazero_b200.trainer import AlphaZeroTrainer
    torch.manual_seed(0)
    net = Connect4Net(board_width=7, board_height=6)
    G, n_sim = 64, 24
    eng = SelfPlayEngine(net, G, n=Connect4Spec(7, 6), n_sim=n_sim, seed=9)
    batch = eng.play().cpu()
    assert eng.total_sims == eng.total_moves * n_sim and batch.pi.shape[1] == 7
    states = batch.states()
    assert states.shape[1:] == (6, 7)
    ep, mv, z = batch.episode_idx.numpy(), batch.move_idx.numpy(), batch.z.numpy()
    pi = batch.pi.numpy()
    assert np.allclose(pi.sum(axis=1), 1, atol=1e-5)
    own, opp = _np(batch.own), _np(batch.opp)
    for g in range(G):
        idx = np.nonzero(ep == g)[0]
        assert (mv[idx] == np.arange(len(idx))).all() and 7 <= len(idx) <= 42
        # consecutive states differ by exactly one dropped disc of the mover, seen from the other side
        for a, b in zip(idx[:-1], idx[1:]):
            assert own[b] == opp[a] and bin(int(opp[b]) ^ int(own[a])).count("1") == 1
            col = (int(opp[b]) ^ int(own[a])).bit_length() - 1
            assert pi[a, col % 8] > 0                       # the played column had visits
        # outcome: z of the last sample = +1 if its mover completed a line, 0 if the board filled up
        last = idx[-1]
        player_last = 1 if len(idx) % 2 == 1 else -1
        assert z[last] in (0, 1) and (z[idx] * np.where(np.arange(len(idx)) % 2 == (len(idx) - 1) % 2, 1, -1) == z[last]).all()
        assert player_last in (1, -1)
    cfg = Connect4Config(simulations=8, episodes=6, data_augmentation=True)
    tr = AlphaZeroTrainer()
    tr.setup("connect4", cfg, net)
    tr.self_play(0, seed=4)
    base = [s for s in tr.memory if s.transformation is None]
    aug = [s for s in tr.memory if s.transformation is not None]
    assert len(aug) == sum(1 for s in base if s.move_idx >= 2)
    assert {s.transformation for s in aug} == {"reflection_horizontal"}
    for s in tr.memory:
        assert s.state.shape == (6, 7) and s.pi.shape == (7,) and s.outcome in (-1, 0, 1)
    s0 = next(s for s in base if s.move_idx >= 2)
    tw = next(s for s in aug if s.episode_idx == s0.episode_idx and s.move_idx == s0.move_idx)
    assert (tw.state == s0.state[:, ::-1]).all() and (tw.pi == s0.pi[::-1]).all()


def test_fused_connect4_net_matches_reference_outputs(golden_dir):
    """The hand-written forward (az_connect4_image_bits -> tcgen05 trunk on the 7x6 image -> fc GEMMs -> fused heads)
    on the reference's own weights: conv features against the fp32 module, probs / value within 1e-2 of the
    reference's fp32 predict (BASELINE.json north_star tolerance), with and without the CUDA graph."""
    import torch.nn.functional as F
    from alphazero_b200.fused_net import FusedNetEvaluator
    net, d = _golden_net(golden_dir)
    assert FusedNetEvaluator.supports(net)
    player = d["player"].astype(np.int8)
    own, opp = _rel(d["black"], d["white"], player)
    G = len(own)
    tree = _mk_tree(G, 7, 6)
    tree.leaf_own.copy_(_dev(own)); tree.leaf_opp.copy_(_dev(opp))
    for graph in (False, True):
        ev = FusedNetEvaluator(net, tree, use_cuda_graph=graph)
        tree.probs.zero_(); tree.value.zero_()
        ev(tree)
        torch.cuda.synchronize()
        assert np.abs(tree.probs.cpu().numpy() - d["probs"]).max() < 1e-2
        assert np.abs(tree.value.cpu().numpy() - d["value"].reshape(-1)).max() < 1e-2
        assert torch.allclose(tree.probs.sum(dim=1).cpu(), torch.ones(G), atol=1e-4)
    x = torch.from_numpy(d["input"]).reshape(-1, 1, 7, 6)      # the (width, height) view of Connect4Net.forward
    with torch.no_grad():
        a = F.relu(net.bn1(net.conv1(x))); a = F.relu(net.bn2(net.conv2(a)))
        a = F.relu(net.bn3(net.conv3(a))); a4 = F.relu(net.bn4(net.conv4(a)))     # [G, 32, 3, 2]
    feat = ev.features_rowmajor().float().cpu().reshape(G, 6, 32).permute(0, 2, 1).reshape(G, 32, 3, 2)
    assert (feat - a4).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02
    # a batch larger than one round of CTAs, with a partially filled last group
    G2 = 4099
    idx = np.arange(G2) % G
    big = _mk_tree(G2, 7, 6, capacity=64)
    big.leaf_own.copy_(_dev(own[idx])); big.leaf_opp.copy_(_dev(opp[idx]))
    ev2 = FusedNetEvaluator(net, big)
    ev2(big)
    torch.cuda.synchronize()
    p = big.probs.cpu().numpy()
    assert np.abs(p[:G] - d["probs"]).max() < 1e-2 and (p[G:2 * G] == p[:G]).all()
    # that batch has more groups of 18 boards than the GPU has SMs: its trunk is k_net_trunk_rows2<7, 6> (two CTAs per SM);
    # its features against the fp32 module and against the first kernel (trunk_variant = 1) on the same batch
    from alphazero_b200 import _lib
    assert _lib.lib().az_net_trunk_boards_per_group(7, G2) == 18
    f2 = ev2.features_rowmajor().float().cpu().reshape(G2, 6, 32).permute(0, 2, 1).reshape(G2, 32, 3, 2)
    assert (f2 - a4[idx]).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02
    ev1 = FusedNetEvaluator(net, big, use_cuda_graph=False, trunk_variant=1)
    ev1.trunk()
    torch.cuda.synchronize()
    f1 = ev1.features_rowmajor().float().cpu().reshape(G2, 6, 32).permute(0, 2, 1).reshape(G2, 32, 3, 2)
    assert (f2 - f1).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02


def test_self_play_connect4_fused_path_is_used():
    from alphazero_b200.fused_net import FusedNetEvaluator
    from alphazero_b200.games.connect4 import Connect4Net
    from alphazero_b200.gamespec import Connect4Spec
    from alphazero_b200.selfplay import SelfPlayEngine
    torch.manual_seed(1)
    eng = SelfPlayEngine(Connect4Net(board_width=7, board_height=6), 32, n=Connect4Spec(7, 6), n_sim=16, seed=3)
    assert isinstance(eng.evaluator, FusedNetEvaluator)
    batch = eng.play()
    assert eng.total_sims == eng.total_moves * 16 and len(batch) == eng.total_moves
    other = SelfPlayEngine(Connect4Net(board_width=5, board_height=6), 8, n=Connect4Spec(5, 6), n_sim=8, seed=3)
    assert not isinstance(other.evaluator, FusedNetEvaluator)      # no fused kernel for this size: module's own graph
    other.play()
