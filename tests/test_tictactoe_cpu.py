"""CPU: TicTacToe.  Pins the oracle's restatement (oracle/az_oracle.c, tictactoe section) to the fixtures generated
from the unmodified Python reference (tests/golden/make_golden_tictactoe.py): all 5 478 reachable positions, perft,
MCT visit counts."""
import json
import os

import numpy as np

import oracle_lib as orc


def test_oracle_vs_reference_all_positions(golden_dir):
    d = np.load(os.path.join(golden_dir, "tictactoe_positions.npz"))
    assert len(d["black"]) == 5478
    black, white, player = d["black"], d["white"], d["player"].astype(np.int8)
    assert (orc.ttt_legal(black, white, player) == d["legal"]).all()
    term, score, winner = orc.ttt_terminal(black, white, player)
    assert (term == d["over"]).all() and (score == d["score"]).all()
    over = d["over"] == 1
    assert (winner[over] == d["winner"][over]).all()
    live = ~over
    ob, ow, op, st = orc.ttt_apply(black[live], white[live], player[live], d["action"][live].astype(np.uint8))
    assert (st == 0).all()
    assert (ob == d["black2"][live]).all() and (ow == d["white2"][live]).all() and (op == d["player2"][live]).all()
    has = d["illegal"] != 255            # occupied cell -> ValueError in the reference (tictactoe.py:159-160)
    _, _, _, st2 = orc.ttt_apply(black[has], white[has], player[has], d["illegal"][has].astype(np.uint8))
    assert (st2 == 1).all()
    for bad in (3, 11, 24, 63, 64, 255):  # cells outside the 3x3 board
        _, _, _, st3 = orc.ttt_apply(black, white, player, np.full(len(black), bad, np.uint8))
        assert (st3 == 1).all()


def test_oracle_perft(golden_dir):
    t = json.load(open(os.path.join(golden_dir, "tictactoe_misc.json")))
    assert [orc.ttt_perft(d) for d in range(1, 10)] == t["perft_1_to_9"]
    assert t["perft_1_to_9"][-1] == 255168        # the well-known number of TicTacToe games


def test_oracle_mct_visit_counts_vs_reference(golden_dir):
    d = np.load(os.path.join(golden_dir, "mct_visits_ttt.npz"))
    salt = int(d["salt"][0])
    i, rows = 0, len(d["root"])
    while i < rows:
        t = orc.OracleMCT(3, "neural", tie_random=False, salt=salt, game="tictactoe")
        j = i
        while j < rows and d["root"][j] == d["root"][i] and d["n_sim"][j] == d["n_sim"][i] and (j == i or d["step"][j] > d["step"][j - 1]):
            t.search(d["black"][j], d["white"][j], d["player"][j], d["n_sim"][j])
            N, Q, P, rn, ne, nr = t.stats()
            assert (N == d["N"][j]).all() and (Q == d["Q"][j]).all() and (P.astype(np.float32) == d["P"][j]).all()
            assert rn == d["rootN"][j] and ne == d["evals"][j] and nr == d["n_sim"][j]
            assert t.best_action() == d["action"][j]
            t.change_root(int(d["action"][j]))
            j += 1
        i = j


def test_oracle_playouts_are_complete_games():
    z = np.zeros(1, np.uint64)
    fb, fw, win, pl = orc.ttt_playouts(z, z, np.ones(1, np.int8), 500, 0x77)
    player_end = np.where(pl % 2 == 0, 1, -1).astype(np.int8)
    t, _, wn = orc.ttt_terminal(fb, fw, player_end)
    assert (t == 1).all() and (wn == win).all() and pl.min() >= 5 and pl.max() <= 9
    assert {-1, 0, 1} == set(win.tolist())
