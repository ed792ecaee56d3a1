"""Generates the golden fixtures of tests/golden/ by IMPORTING THE UNMODIFIED REFERENCE (/root/reference).

Run in the build container only (the reference does not exist on the GPU box):
    python tests/golden/make_golden.py
Outputs (committed):
    othello_positions_n{4,6,8}.npz  random positions reached by random play of the reference OthelloBoard with,
                                    for each: legal mask (get_moves as a set), is_game_over, get_score,
                                    get_winner (when over), one played move and the resulting grid.
    perft.json                      perft(1..d) of the reference OthelloBoard (pass = ply, finished game = 1 leaf).
    mct_visits_n{6,8}.npz           alphazero.mcts.MCT("neural") visit counts / Q / P of the root's children under the
                                    deterministic hash evaluator (tests/hash_eval.py) and a deterministic tie-break
                                    (fair_max -> lowest action index), with tree reuse across moves.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ref_import import import_reference  # noqa: E402
from hash_eval import grid_to_bits, hash_eval  # noqa: E402

import_reference()
import alphazero.mcts as ref_mcts  # noqa: E402
from alphazero.games.othello import OthelloBoard, OthelloNet  # noqa: E402


def move_to_action(move, n):
    return 64 if move == (n, n) else move[0] * 8 + move[1]


def positions(n, count, seed):
    rng = np.random.RandomState(seed)
    np.random.seed(seed)
    rec = {k: [] for k in ("black", "white", "player", "legal", "over", "score", "winner", "action",
                           "black2", "white2", "player2")}
    while len(rec["black"]) < count:
        b = OthelloBoard(n=n)
        while True:
            black, white = grid_to_bits(b.grid)
            moves = b.get_moves()
            legal = 0
            for m in moves:
                if m != b.pass_move:
                    legal |= 1 << (m[0] * 8 + m[1])
            over = b.is_game_over()
            keep = over or rng.rand() < 0.35
            if keep:
                rec["black"].append(black); rec["white"].append(white); rec["player"].append(b.player)
                rec["legal"].append(legal); rec["over"].append(int(over)); rec["score"].append(int(b.get_score()))
                rec["winner"].append(int(b.get_winner()) if over else 0)
            if over:
                if keep:
                    rec["action"].append(255); rec["black2"].append(black); rec["white2"].append(white)
                    rec["player2"].append(b.player)
                break
            mv = moves[rng.randint(len(moves))]
            b.play_move(mv)
            if keep:
                b2, w2 = grid_to_bits(b.grid)
                rec["action"].append(move_to_action(mv, n)); rec["black2"].append(b2); rec["white2"].append(w2)
                rec["player2"].append(b.player)
    out = {k: np.array(v[:count], dtype=np.uint64 if k in ("black", "white", "legal", "black2", "white2") else np.int16)
           for k, v in rec.items()}
    np.savez_compressed(os.path.join(HERE, f"othello_positions_n{n}.npz"), **out)
    print(f"positions n={n}: {count}")


def perft(b, d):
    if d == 0 or b.is_game_over():
        return 1
    t = 0
    for m in b.get_moves():
        c = b.clone(); c.play_move(m); t += perft(c, d - 1)
    return t


class HashNet:
    """Fake PolicyValueNetwork for the reference MCT: evaluate() per alphazero/base.py:357-367 semantics."""

    def __init__(self, n, salt):
        self.n = n
        self.salt = salt
        self.n_evals = 0

    def evaluate(self, board):
        black, white = grid_to_bits(board.grid)
        own, opp = (black, white) if board.player == 1 else (white, black)
        probs, v_net = hash_eval(own, opp, self.n * self.n + 1, self.salt)
        self.n_evals += 1
        return probs, board.player * float(np.float32(v_net))

    def get_normalized_probs(self, probs, legal_moves):
        return OthelloNet.get_normalized_probs(self, probs, legal_moves)


def det_fair_max(elements, key=lambda x: x):
    """Deterministic stand-in for alphazero.utils.fair_max: ties -> lowest action index (row-major, pass last)."""
    elements = list(elements)
    best = max(key(x) for x in elements)
    ties = [x for x in elements if key(x) == best]
    return min(ties, key=lambda x: x[0][0] * 1000 + x[0][1])


def mct_fixture(n, n_roots, n_moves, sims_list, seed):
    ref_mcts.fair_max = det_fair_max
    rng = np.random.RandomState(seed)
    A = 65
    rows = []
    for root_idx in range(n_roots):
        b = OthelloBoard(n=n)
        for _ in range(0 if root_idx == 0 else rng.randint(0, n * n - 10)):
            if b.is_game_over():
                break
            mv = b.get_moves(); b.play_move(mv[rng.randint(len(mv))])
        if b.is_game_over():
            continue
        for n_sim in sims_list:
            board = b.clone()
            net = HashNet(n, salt=seed)
            mct = ref_mcts.MCT(eval_method="neural", nn=net)
            for step in range(n_moves):
                if board.is_game_over():
                    break
                mct.search(board, n_sim=n_sim)
                black, white = grid_to_bits(board.grid)
                N = np.full(A, -1, dtype=np.int64); Q = np.zeros(A); P = np.zeros(A, dtype=np.float32)
                for mv, node in mct.root.children.items():
                    a = move_to_action(mv, n)
                    N[a] = node.N; Q[a] = float(node.Q); P[a] = np.float32(node.P)
                best, _ = det_fair_max(mct.root.children.items(), key=lambda x: x[1].N)
                rows.append(dict(root=root_idx, n_sim=n_sim, step=step, black=black, white=white,
                                 player=board.player, N=N, Q=Q, P=P, rootN=mct.root.N, action=move_to_action(best, n),
                                 evals=net.n_evals))
                board.play_move(best)
                mct.change_root(best)
    out = dict(
        root=np.array([r["root"] for r in rows], np.int32), n_sim=np.array([r["n_sim"] for r in rows], np.int32),
        step=np.array([r["step"] for r in rows], np.int32), black=np.array([r["black"] for r in rows], np.uint64),
        white=np.array([r["white"] for r in rows], np.uint64), player=np.array([r["player"] for r in rows], np.int8),
        N=np.stack([r["N"] for r in rows]), Q=np.stack([r["Q"] for r in rows]), P=np.stack([r["P"] for r in rows]),
        rootN=np.array([r["rootN"] for r in rows], np.int64), action=np.array([r["action"] for r in rows], np.int16),
        evals=np.array([r["evals"] for r in rows], np.int64), salt=np.array([seed], np.uint64),
    )
    np.savez_compressed(os.path.join(HERE, f"mct_visits_n{n}.npz"), **out)
    print(f"mct n={n}: {len(rows)} searches")


if __name__ == "__main__":
    t0 = time.time()
    for n, cnt in ((4, 600), (6, 1500), (8, 3000)):
        positions(n, cnt, seed=1234 + n)
    table = {"8": [perft(OthelloBoard(n=8), d) for d in range(1, 7)],
             "6": [perft(OthelloBoard(n=6), d) for d in range(1, 7)],
             "4": [perft(OthelloBoard(n=4), d) for d in range(1, 9)],
             "known_8x8_to_depth_9": [4, 12, 56, 244, 1396, 8200, 55092, 390216, 3005288],
             "known_6x6_to_depth_7": [4, 12, 56, 244, 1364, 7604, 47740],
             "convention": "pass is a ply, a finished game counts as one leaf; entries '8','6','4' were computed by "
                           "the unmodified reference OthelloBoard, 'known_*' are SURVEY.md section 4 (reference-"
                           "reproduced there on 8 processes)"}
    json.dump(table, open(os.path.join(HERE, "perft.json"), "w"), indent=1)
    print("perft", table["8"], table["6"], table["4"])
    mct_fixture(8, n_roots=10, n_moves=3, sims_list=(1, 2, 50, 200), seed=7)
    mct_fixture(6, n_roots=8, n_moves=4, sims_list=(1, 3, 64), seed=11)
    print(f"done in {time.time() - t0:.1f}s")
