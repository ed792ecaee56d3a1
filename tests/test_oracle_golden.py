"""CPU: pins the oracle (oracle/az_oracle.c) to the fixtures generated from the unmodified Python reference
(tests/golden/make_golden.py) and to the known-answer tables."""
import ctypes
import json
import os

import numpy as np
import pytest

import oracle_lib as orc
from hash_eval import hash_eval


def _abs_to_rel(black, white, player):
    own = np.where(player > 0, black, white)
    opp = np.where(player > 0, white, black)
    return own, opp


@pytest.mark.parametrize("n", [4, 6, 8])
def test_board_vs_reference_positions(golden_dir, n):
    d = np.load(os.path.join(golden_dir, f"othello_positions_n{n}.npz"))
    black, white = d["black"], d["white"]
    player = d["player"].astype(np.int8)
    assert (orc.legal(black, white, player, n) == d["legal"]).all()
    term, score, winner = orc.terminal(black, white, player, n)
    assert (term == d["over"]).all()
    assert (score == d["score"]).all()
    over = d["over"] == 1
    assert (winner[over] == d["winner"][over]).all()
    live = ~over
    act = d["action"][live].astype(np.uint8)
    ob, ow, op, fl, st = orc.apply(black[live], white[live], player[live], act, n)
    assert (st == 0).all()
    assert (ob == d["black2"][live]).all() and (ow == d["white2"][live]).all()
    assert (op == d["player2"][live]).all()
    # illegal moves are refused: occupied cell, and pass while a placement exists
    occ = np.array([int(b | w).bit_length() - 1 for b, w in zip(black[live], white[live])], dtype=np.uint8)
    _, _, _, _, st2 = orc.apply(black[live], white[live], player[live], occ, n)
    assert (st2 == 1).all()
    has_move = d["legal"][live] != 0
    _, _, _, _, st3 = orc.apply(black[live][has_move], white[live][has_move], player[live][has_move],
                               np.full(has_move.sum(), 64, np.uint8), n)
    assert (st3 == 1).all()


def test_perft_tables(golden_dir):
    t = json.load(open(os.path.join(golden_dir, "perft.json")))
    for n in (4, 6, 8):
        ref = t[str(n)]
        assert [orc.perft(n, d) for d in range(1, len(ref) + 1)] == ref
    assert [orc.perft(8, d) for d in range(1, 9)] == t["known_8x8_to_depth_9"][:8]
    assert [orc.perft(6, d) for d in range(1, 8)] == t["known_6x6_to_depth_7"]


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    kat = [
        ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
        ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
        ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
         [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
    ]
    for c, k, want in kat:
        cc, kk, oo = (ctypes.c_uint32 * 4)(*c), (ctypes.c_uint32 * 2)(*k), (ctypes.c_uint32 * 4)()
        orc.lib().orc_philox4x32_10(cc, kk, oo)
        assert list(oo) == want


def test_hash_eval_matches_python():
    rng = np.random.default_rng(3)
    for _ in range(50):
        own = int(rng.integers(0, 2**63)); opp = int(rng.integers(0, 2**63)) & ~own
        for A in (37, 65):
            p, v = hash_eval(own, opp, A, 99)
            p2, v2 = orc.hash_eval(own, opp, A, 99)
            assert (p == p2).all() and v == v2


@pytest.mark.parametrize("n", [6, 8])
def test_mct_visit_counts_vs_reference(golden_dir, n):
    """Identical N, float64 Q, float32 P, root N and evaluation counts as alphazero.mcts.MCT (deterministic
    evaluator + lowest-index tie-break on both sides), including tree reuse across moves."""
    d = np.load(os.path.join(golden_dir, f"mct_visits_n{n}.npz"))
    salt = int(d["salt"][0])
    i, rows = 0, len(d["root"])
    while i < rows:
        t = orc.OracleMCT(n, "neural", tie_random=False, salt=salt)
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


def test_playouts_are_complete_games():
    b, w = orc.start(8)
    fb, fw, win, pl = orc.playouts(np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8), 8, 200, 0x5EED)
    t, s, wn = orc.terminal(fb, fw, np.ones(200, np.int8), 8)
    assert (t == 1).all()
    assert (np.sign(s) == win).all()
    assert pl.min() >= 9 and pl.max() <= 70
    assert len(set(zip(fb.tolist(), fw.tolist()))) > 150
