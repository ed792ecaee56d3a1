"""CPU: Connect4.  Pins the oracle's cell-grid restatement (oracle/az_oracle.c, connect4 section) to the fixtures
generated from the unmodified Python reference (tests/golden/make_golden_connect4.py), then checks the HOST side of
the device bitboard header (connect4_bits.cuh / rules.cuh) against the oracle on random walks."""
import ctypes
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIZES = [(7, 6), (4, 4), (8, 8), (5, 7)]


@pytest.mark.parametrize("w,h", SIZES)
def test_oracle_vs_reference_positions(golden_dir, w, h):
    d = np.load(os.path.join(golden_dir, f"connect4_positions_{w}x{h}.npz"))
    black, white = d["black"], d["white"]
    player = d["player"].astype(np.int8)
    assert (orc.c4_legal(black, white, player, w, h) == d["legal"]).all()
    term, score, winner = orc.c4_terminal(black, white, player, w, h)
    assert (term == d["over"]).all()
    assert (score == d["score"]).all()
    over = d["over"] == 1
    assert (winner[over] == d["winner"][over]).all()
    live = ~over
    act = d["action"][live].astype(np.uint8)
    ob, ow, op, st = orc.c4_apply(black[live], white[live], player[live], act, w, h)
    assert (st == 0).all()
    assert (ob == d["black2"][live]).all() and (ow == d["white2"][live]).all() and (op == d["player2"][live]).all()
    # out-of-range column and full column are refused (the reference raises ValueError, connect4.py:145-175)
    _, _, _, st2 = orc.c4_apply(black, white, player, np.full(len(black), w, np.uint8), w, h)
    assert (st2 == 1).all()
    full = np.array([[c for c in range(w) if not (int(l) >> c) & 1] for l in d["legal"]], dtype=object)
    idx = [i for i, f in enumerate(full) if len(f)]
    if idx:
        cols = np.array([full[i][0] for i in idx], np.uint8)
        _, _, _, st3 = orc.c4_apply(black[idx], white[idx], player[idx], cols, w, h)
        assert (st3 == 1).all()


def test_oracle_perft_tables(golden_dir):
    t = json.load(open(os.path.join(golden_dir, "perft_connect4.json")))
    for key in ("7x6", "4x4", "5x4"):
        w, h = (int(x) for x in key.split("x"))
        assert [orc.c4_perft(w, h, d) for d in range(1, len(t[key]) + 1)] == t[key]
    # known: 7x6 perft(7) = 7^7 minus the 7 sequences that put 7 discs in one column
    assert orc.c4_perft(7, 6, 7) == 7 ** 7 - 7


@pytest.mark.parametrize("w,h", [(7, 6), (5, 4)])
def test_oracle_mct_visit_counts_vs_reference(golden_dir, w, h):
    """Identical N, float64 Q, float32 P, root N and evaluation counts as alphazero.mcts.MCT on Connect4."""
    d = np.load(os.path.join(golden_dir, f"mct_visits_c4_{w}x{h}.npz"))
    salt = int(d["salt"][0])
    i, rows = 0, len(d["root"])
    while i < rows:
        t = orc.OracleMCT((w, h), "neural", tie_random=False, salt=salt, game="connect4")
        j = i
        while j < rows and d["root"][j] == d["root"][i] and d["n_sim"][j] == d["n_sim"][i] and (j == i or d["step"][j] > d["step"][j - 1]):
            t.search(d["black"][j], d["white"][j], d["player"][j], d["n_sim"][j])
            N, Q, P, rn, ne, nr = t.stats()
            assert (N == d["N"][j]).all()
            assert (Q == d["Q"][j]).all()
            assert (P.astype(np.float32) == d["P"][j]).all()
            assert rn == d["rootN"][j] and ne == d["evals"][j] and nr == d["n_sim"][j]
            assert t.best_action() == d["action"][j]
            t.change_root(int(d["action"][j]))
            j += 1
        i = j


def test_oracle_playouts_are_complete_games():
    z = np.zeros(1, np.uint64)
    fb, fw, win, pl = orc.c4_playouts(z, z, np.ones(1, np.int8), 7, 6, 300, 0xC4)
    # the player to move at the end alternates with the plies played
    player_end = np.where(pl % 2 == 0, 1, -1).astype(np.int8)
    t, s, wn = orc.c4_terminal(fb, fw, player_end, 7, 6)
    assert (t == 1).all() and (wn == win).all()
    assert pl.min() >= 7 and pl.max() <= 42
    assert {-1, 1} <= set(win.tolist())


@pytest.fixture(scope="module")
def hb():
    so = os.path.join(ROOT, "tests", "_host_bits.so")
    subprocess.run(["g++", "-O2", "-x", "c++", "-shared", "-fPIC", "-o", so,
                    os.path.join(ROOT, "tests", "host_bits_harness.cpp")], check=True)
    return ctypes.CDLL(so)


@pytest.mark.parametrize("w,h", SIZES + [(8, 4), (4, 8)])
def test_bitboard_logic_vs_oracle(hb, w, h):
    P = orc.P
    rng = np.random.default_rng(w * 10 + h)
    G = 3000
    black, white = np.zeros(G, np.uint64), np.zeros(G, np.uint64)
    player = np.ones(G, np.int8)
    n_term = 0
    for ply in range(w * h + 1):
        own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
        legal = orc.c4_legal(black, white, player, w, h)
        l2 = np.zeros(G, np.uint64)
        hb.hb_c4_legal(P(own), P(opp), w, ctypes.c_int64(G), P(l2))
        assert (legal == l2).all()
        term, _, winner = orc.c4_terminal(black, white, player, w, h)
        t2, w2 = np.zeros(G, np.uint8), np.zeros(G, np.int8)
        hb.hb_c4_terminal(P(own), P(opp), P(player), w, h, ctypes.c_int64(G), P(t2), P(w2))
        assert (term == t2).all()
        assert (winner[term == 1] == w2[term == 1]).all()
        live = term == 0
        n_term += int((~live).sum())
        act = np.zeros(G, np.uint8)
        for g in np.nonzero(live)[0]:
            cols = [c for c in range(w) if (int(legal[g]) >> c) & 1]
            act[g] = cols[rng.integers(len(cols))]
        ob, ow, op, st = orc.c4_apply(black, white, player, act, w, h)
        assert (st[live] == 0).all()
        o2, p2 = own.copy(), opp.copy()
        idx = np.nonzero(live)[0]
        oo, pp = np.ascontiguousarray(o2[idx]), np.ascontiguousarray(p2[idx])
        aa = np.ascontiguousarray(act[idx])
        hb.hb_c4_apply(P(oo), P(pp), P(aa), w, h, ctypes.c_int64(len(idx)))
        nb, nw = np.where(op[idx] > 0, oo, pp), np.where(op[idx] > 0, pp, oo)
        assert (nb == ob[idx]).all() and (nw == ow[idx]).all()
        black, white, player = np.where(live, ob, black), np.where(live, ow, white), np.where(live, op, player)
    term, _, _ = orc.c4_terminal(black, white, player, w, h)
    assert (term == 1).all() and n_term > G
