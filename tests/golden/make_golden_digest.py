"""SHA-256 digests of what the UNMODIFIED reference boards compute on a million positions per game.

    python tests/golden/make_golden_digest.py [positions_per_game]      (build container only; ~10 min on 8 cores)

Storing a million positions is pointless; their digest pins the oracle just as hard.  A chunk = `games` random games
from the start position, every position of every game (the terminal one included), game after game.  The walk is
reproducible by anyone who can list legal moves: game j of a chunk starts its own 64-bit LCG at
lcg(chunk seed + j * 0x9E3779B97F4A7C15) and at every position plays the k-th legal action IN ASCENDING ACTION ORDER,
k = (next LCG state >> 33) % (number of legal moves) -- never the order of the reference's hash-ordered get_moves() (T1).
Per position the digest absorbs, as little-endian integers:
    black u64, white u64 (8-wide canvas bitboards), player i8, legal-action set u64 (Othello: 0 = pass only),
    is_game_over u8, get_score i16 (TicTacToe's inf as 127), get_winner i8 (0 when the game is live).
The oracle regenerates the same walk with its own move generator (tests/test_oracle_digest.py): a single differing
legal set, score, terminal flag or flip changes the digest.
Output (committed): tests/golden/reference_digests.json.
"""
import hashlib
import json
import os
import struct
import sys
import time
from multiprocessing import Pool

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

GAMES = {"othello8": ("othello", 8, 8), "othello6": ("othello", 6, 6), "connect4_7x6": ("connect4", 7, 6)}
MASK = (1 << 64) - 1


def lcg(state):
    return (state * 6364136223846793005 + 1442695040888963407) & MASK


def pack(black, white, player, legal, over, score, winner):
    return struct.pack("<QQbQBhb", black, white, player, legal, over, score, winner)


def chunk(args):
    key, seed, games = args
    from ref_import import import_reference
    from hash_eval import grid_to_bits
    import_reference()
    from alphazero.games.connect4 import Connect4Board
    from alphazero.games.othello import OthelloBoard
    game, n, h = GAMES[key]
    h256 = hashlib.sha256()
    count = 0
    for j in range(games):
        state = lcg((seed + j * 0x9E3779B97F4A7C15) & MASK)
        b = OthelloBoard(n=n) if game == "othello" else Connect4Board(width=n, height=h)
        while True:
            black, white = grid_to_bits(b.grid)
            moves = b.get_moves()
            if game == "othello":
                acts = sorted(64 if tuple(m) == tuple(b.pass_move) else m[0] * 8 + m[1] for m in moves)
                back = {64 if tuple(m) == tuple(b.pass_move) else m[0] * 8 + m[1]: m for m in moves}
            else:
                acts = sorted(int(m) for m in moves)
                back = {int(m): m for m in moves}
            legal = 0
            for a in acts:
                if a < 64:
                    legal |= 1 << a
            over = bool(b.is_game_over())
            h256.update(pack(black, white, int(b.player), legal, int(over), int(b.get_score()),
                             int(b.get_winner()) if over else 0))
            count += 1
            if over:
                break
            state = lcg(state)
            b.play_move(back[acts[(state >> 33) % len(acts)]])
    return key, seed, games, count, h256.hexdigest()


if __name__ == "__main__":
    target = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    per_game = {"othello8": 61.0, "othello6": 33.0, "connect4_7x6": 22.0}      # mean positions per random game
    jobs = []
    for key in GAMES:
        n_chunks = 64
        games = int(target / per_game[key] / n_chunks) + 1
        jobs += [(key, 0x9E3779B97F4A7C15 * (i + 1) & MASK, games) for i in range(n_chunks)]
    t0 = time.time()
    with Pool(os.cpu_count()) as pool:
        res = pool.map(chunk, jobs, chunksize=1)
    out = {"format": "see tests/golden/make_golden_digest.py", "games": {}}
    for key in GAMES:
        rows = [r for r in res if r[0] == key]
        out["games"][key] = {"chunks": [{"seed": r[1], "games": r[2], "positions": r[3], "sha256": r[4]} for r in rows],
                             "positions": sum(r[3] for r in rows)}
        print(key, out["games"][key]["positions"], "positions")
    json.dump(out, open(os.path.join(HERE, "reference_digests.json"), "w"), indent=1)
    print(f"done in {time.time() - t0:.0f}s")
