"""GPU: Othello board kernels (through the C ABI) bit-exact against the reference fixtures and the oracle."""
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


def _rel(black, white, player):
    return np.where(player > 0, black, white), np.where(player > 0, white, black)


@pytest.mark.parametrize("n", [4, 6, 8])
def test_kernels_vs_reference_fixtures(golden_dir, n):
    from alphazero_b200 import boardops as bo
    d = np.load(os.path.join(golden_dir, f"othello_positions_n{n}.npz"))
    player = d["player"].astype(np.int8)
    own, opp = _rel(d["black"], d["white"], player)
    t_own, t_opp, t_pl = _dev(own), _dev(opp), torch.from_numpy(player).cuda()
    assert (bo.bits_to_numpy(bo.legal(t_own, t_opp, n)) == d["legal"]).all()
    term, score, winner = bo.terminal_score(t_own, t_opp, t_pl, n)
    assert (term.cpu().numpy() == d["over"]).all()
    assert (score.cpu().numpy() == d["score"]).all()
    over = d["over"] == 1
    assert (winner.cpu().numpy()[over] == d["winner"][over]).all()
    live = ~over
    act = np.where(live, d["action"], 64).astype(np.uint8)
    o2, p2, pl2, st = bo.apply(t_own, t_opp, t_pl, torch.from_numpy(act).cuda(), n)
    o2, p2, pl2, st = bo.bits_to_numpy(o2), bo.bits_to_numpy(p2), pl2.cpu().numpy(), st.cpu().numpy()
    assert (st[live] == 0).all()
    nb, nw = np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2)
    assert (nb[live] == d["black2"][live]).all() and (nw[live] == d["white2"][live]).all()
    assert (pl2[live] == d["player2"][live]).all()
    # encode = player * grid
    enc = bo.encode(t_own, t_opp, n).cpu().numpy()
    for g in range(0, len(own), max(1, len(own) // 50)):
        for r in range(n):
            for c in range(n):
                want = ((int(own[g]) >> (r * 8 + c)) & 1) - ((int(opp[g]) >> (r * 8 + c)) & 1)
                assert enc[g, r, c] == want
    encb = bo.encode(t_own, t_opp, n, dtype=torch.bfloat16).float().cpu().numpy()
    assert (encb == enc).all()


@pytest.mark.parametrize("n", [6, 8])
def test_kernels_vs_oracle_random_walk(n):
    """~100k positions per size reached by uniform random play; every kernel compared with the ray-walking oracle,
    including the refusal of illegal moves (status 1, board unchanged)."""
    from alphazero_b200 import boardops as bo
    rng = np.random.default_rng(100 + n)
    G = 3000
    b0, w0 = orc.start(n)
    black, white, player = np.full(G, b0, np.uint64), np.full(G, w0, np.uint64), np.ones(G, np.int8)
    for ply in range(n * n + 2):
        own, opp = _rel(black, white, player)
        t_own, t_opp, t_pl = _dev(own), _dev(opp), torch.from_numpy(player).cuda()
        legal = orc.legal(black, white, player, n)
        assert (bo.bits_to_numpy(bo.legal(t_own, t_opp, n)) == legal).all()
        term, score, winner = orc.terminal(black, white, player, n)
        d_term, d_score, d_win = bo.terminal_score(t_own, t_opp, t_pl, n)
        assert (d_term.cpu().numpy() == term).all() and (d_score.cpu().numpy() == score).all()
        assert (d_win.cpu().numpy() == winner).all()
        # a random action per board: mostly legal, sometimes arbitrary (illegal) cells
        act = np.full(G, 64, np.uint8)
        for g in range(G):
            m = int(legal[g])
            if m and rng.random() < 0.9:
                bits = [i for i in range(64) if (m >> i) & 1]
                act[g] = bits[rng.integers(len(bits))]
            elif rng.random() < 0.5:
                act[g] = rng.integers(0, 65)
        ob, ow, op, fl, st = orc.apply(black, white, player, act, n)
        t_act = torch.from_numpy(act).cuda()
        d_fl = bo.bits_to_numpy(bo.flips(t_own, t_opp, t_act, n))
        ok = st == 0
        assert (d_fl[ok] == fl[ok]).all()
        o2, p2, pl2, d_st = bo.apply(t_own, t_opp, t_pl, t_act, n)
        o2, p2, pl2 = bo.bits_to_numpy(o2), bo.bits_to_numpy(p2), pl2.cpu().numpy()
        assert (d_st.cpu().numpy() == st).all()
        nb, nw = np.where(pl2 > 0, o2, p2), np.where(pl2 > 0, p2, o2)
        assert (nb == ob).all() and (nw == ow).all() and (pl2 == op).all()
        black, white, player = ob, ow, op


def test_perft_device(golden_dir):
    from alphazero_b200 import boardops as bo
    t = json.load(open(os.path.join(golden_dir, "perft.json")))
    assert [bo.perft(8, d) for d in range(1, 10)] == t["known_8x8_to_depth_9"]
    assert [bo.perft(6, d) for d in range(1, 8)] == t["known_6x6_to_depth_7"]
    assert [bo.perft(4, d) for d in range(1, 9)] == t["4"]
    # different BFS/DFS splits give the same count
    assert bo.perft(8, 8, dfs_depth=1) == bo.perft(8, 8, dfs_depth=5) == 390216
    # perft from a mid-game position with passes (oracle as the checker)
    rng = np.random.default_rng(5)
    for n in (6, 8):
        b, w = orc.start(n)
        black, white, player = np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8)
        for _ in range(n * n - 14):
            legal = int(orc.legal(black, white, player, n)[0])
            bits = [i for i in range(64) if (legal >> i) & 1] or [64]
            black, white, player, _, _ = orc.apply(black, white, player, np.array([bits[rng.integers(len(bits))]], np.uint8), n)
        own, opp = _rel(black, white, player)
        for d in (1, 3, 5, 6):
            want = orc.perft(n, d, int(black[0]), int(white[0]), int(player[0]))
            assert bo.perft(n, d, int(own[0]), int(opp[0])) == want


def test_playouts_vs_oracle_and_properties():
    """1M playouts from the start position (BASELINE config 2): a 20k-playout slice is compared bit-exactly with
    the oracle (same Philox stream), the full million through size-independent properties."""
    from alphazero_b200 import boardops as bo
    n, seed = 8, 0x5EED
    b, w = orc.start(n)
    own = _dev(np.array([b], np.uint64)); opp = _dev(np.array([w], np.uint64))
    pl = torch.ones(1, dtype=torch.int8, device="cuda")
    total = 1_000_000
    fb, fw, win, plies = bo.playouts(own, opp, pl, n, total, seed)
    fbn, fwn, winn, pliesn = bo.bits_to_numpy(fb), bo.bits_to_numpy(fw), win.cpu().numpy(), plies.cpu().numpy()
    # exact slice vs oracle (ids 0..9999 and 990000..999999)
    for first in (0, 990_000):
        ofb, ofw, owin, opl = orc.playouts(np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8),
                                           n, 10_000, seed, first_id=first)
        sl = slice(first, first + 10_000)
        assert (fbn[sl] == ofb).all() and (fwn[sl] == ofw).all()
        assert (winn[sl] == owin).all() and (pliesn[sl] == opl).all()
    # properties on the full million: final positions are finished games, winner = sign of disc difference,
    # no overlap, plausible lengths, both colours win sometimes
    assert (fbn & fwn == 0).all()
    term, score, _ = bo.terminal_score(fb, fw, torch.ones(total, dtype=torch.int8, device="cuda"), n)
    assert bool((term == 1).all())
    assert bool((torch.sign(score.int()) == win.int()).all())
    assert pliesn.min() >= 9 and pliesn.max() <= 80
    frac_black = float((winn == 1).mean())
    assert 0.40 < frac_black < 0.52 and float((winn == 0).mean()) < 0.08
    # the stream is a pure function of (seed, id): a second launch with an offset reproduces the tail
    fb2, fw2, _, _ = bo.playouts(own, opp, pl, n, 1000, seed, first_id=total - 1000)
    assert (bo.bits_to_numpy(fb2) == fbn[-1000:]).all() and (bo.bits_to_numpy(fw2) == fwn[-1000:]).all()


def test_edge_cases():
    from alphazero_b200 import boardops as bo
    empty = torch.empty(0, dtype=torch.int64, device="cuda")
    assert bo.legal(empty, empty, 8).numel() == 0
    with pytest.raises(ValueError):
        bo.legal(torch.zeros(1, dtype=torch.int64, device="cuda"), torch.zeros(1, dtype=torch.int64, device="cuda"), 5)
    with pytest.raises(ValueError):
        bo.legal(torch.zeros(1, dtype=torch.int64), torch.zeros(1, dtype=torch.int64), 8)
    # full board / one-colour board are terminal
    full = _dev(np.array([0xFFFFFFFFFFFFFFFF, 0xFFFF], np.uint64)); none = _dev(np.array([0, 0], np.uint64))
    term, score, winner = bo.terminal_score(full, none, torch.tensor([1, -1], dtype=torch.int8, device="cuda"), 8)
    assert term.tolist() == [1, 1] and score.tolist() == [64, 16] and winner.tolist() == [1, -1]
