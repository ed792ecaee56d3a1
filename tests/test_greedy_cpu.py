"""CPU: the oracle's GreedyPlayer restatement (oracle/az_oracle.c: orc_greedy_moves) against the fixture recorded from
the unmodified reference GreedyPlayer.get_move (tests/golden/make_golden_greedy.py)."""
import os

import numpy as np
import pytest

import oracle_lib as orc

GAMES = {"othello8": (0, 8, 8), "othello6": (0, 6, 6), "connect4_7x6": (1, 7, 6), "tictactoe": (2, 3, 3)}


@pytest.mark.parametrize("key", sorted(GAMES))
def test_oracle_greedy_matches_reference_fixture(golden_dir, key):
    d = np.load(os.path.join(golden_dir, "greedy_moves.npz"))
    game, n, h = GAMES[key]
    best, ties = orc.greedy_moves(game, d[f"{key}_black"], d[f"{key}_white"], d[f"{key}_player"].astype(np.int8), n, h)
    assert (ties == d[f"{key}_ties"]).all()
    assert (best == d[f"{key}_best"].astype(np.uint8)).all()
    assert len(best) >= 600
