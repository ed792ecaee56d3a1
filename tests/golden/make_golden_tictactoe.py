"""Generates the TicTacToe golden fixtures by IMPORTING THE UNMODIFIED REFERENCE (/root/reference).

Run in the build container only (the reference does not exist on the GPU box):
    python tests/golden/make_golden_tictactoe.py
Outputs (committed):
    tictactoe_positions.npz   EVERY position reachable by legal play (5 478) of the reference TicTacToeBoard with, for
                              each: empty-cell mask (get_moves), is_game_over, get_score (inf -> 127), get_winner (when
                              over) and every legal move's resulting grid is covered by the successor being in the set;
                              plus one played move per live position and illegal-move probes.
    tictactoe_misc.json       perft(1..9) by the reference board (a finished game counts as one leaf).
    mct_visits_ttt.npz        alphazero.mcts.MCT("neural") visit counts / Q / P of the root's children under the
                              deterministic hash evaluator and the deterministic tie-break, with tree reuse.
    tictactoe_net.npz         a seeded reference TicTacToeNet: state_dict, all live positions as inputs, fp32 outputs.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ref_import import import_reference  # noqa: E402
from hash_eval import grid_to_bits, hash_eval  # noqa: E402

import_reference()
import alphazero.mcts as ref_mcts  # noqa: E402
from alphazero.games.tictactoe import TicTacToeBoard, TicTacToeNet  # noqa: E402


def act(move):
    return int(move[0]) * 8 + int(move[1])


def all_positions():
    seen, order, stack = {}, [], [TicTacToeBoard()]
    while stack:
        b = stack.pop()
        key = (tuple(b.grid.flatten().astype(int)), b.player)
        if key in seen:
            continue
        seen[key] = b
        order.append(b)
        if not b.is_game_over():
            for m in b.get_moves():
                c = b.clone(); c.play_move(m); stack.append(c)
    return order


def positions():
    rng = np.random.RandomState(33)
    rec = {k: [] for k in ("black", "white", "player", "legal", "over", "score", "winner", "action", "black2", "white2",
                           "player2", "illegal")}
    for b in all_positions():
        black, white = grid_to_bits(b.grid)
        legal = 0
        for m in b.get_moves():
            legal |= 1 << act(m)
        over = b.is_game_over()
        score = b.get_score()
        rec["black"].append(black); rec["white"].append(white); rec["player"].append(b.player)
        rec["legal"].append(legal); rec["over"].append(int(over)); rec["score"].append(127 if score == float("inf") else int(score))
        rec["winner"].append(int(b.get_winner()) if over else 0)
        occupied = [(r, c) for r in range(3) for c in range(3) if b.grid[r, c] != 0]
        rec["illegal"].append(act(occupied[rng.randint(len(occupied))]) if occupied else 255)
        moves = b.get_moves()
        if over or not moves:
            rec["action"].append(255); rec["black2"].append(black); rec["white2"].append(white); rec["player2"].append(b.player)
        else:
            mv = moves[rng.randint(len(moves))]
            c = b.clone(); c.play_move(mv)
            b2, w2 = grid_to_bits(c.grid)
            rec["action"].append(act(mv)); rec["black2"].append(b2); rec["white2"].append(w2); rec["player2"].append(c.player)
    u64 = ("black", "white", "legal", "black2", "white2")
    out = {k: np.array(v, dtype=np.uint64 if k in u64 else np.int16) for k, v in rec.items()}
    np.savez_compressed(os.path.join(HERE, "tictactoe_positions.npz"), **out)
    print(f"tictactoe positions: {len(out['black'])} ({int(out['over'].sum())} finished)")
    return out


def perft(b, d):
    if d == 0 or b.is_game_over():
        return 1
    t = 0
    for m in b.get_moves():
        c = b.clone(); c.play_move(m); t += perft(c, d - 1)
    return t


class HashNet:
    def __init__(self, salt):
        self.action_size = 9
        self.salt = salt
        self.n_evals = 0

    def evaluate(self, board):
        black, white = grid_to_bits(board.grid)
        own, opp = (black, white) if board.player == 1 else (white, black)
        probs, v_net = hash_eval(own, opp, self.action_size, self.salt)
        self.n_evals += 1
        return probs, board.player * float(np.float32(v_net))

    def get_normalized_probs(self, probs, legal_moves):
        return TicTacToeNet.get_normalized_probs(self, probs, legal_moves)


def det_fair_max(elements, key=lambda x: x):
    elements = list(elements)
    best = max(key(x) for x in elements)
    ties = [x for x in elements if key(x) == best]
    return min(ties, key=lambda x: act(x[0]))


def mct_fixture(n_roots, n_moves, sims_list, seed):
    ref_mcts.fair_max = det_fair_max
    rng = np.random.RandomState(seed)
    A = 65
    rows = []
    for root_idx in range(n_roots):
        b = TicTacToeBoard()
        for _ in range(0 if root_idx == 0 else rng.randint(0, 6)):
            if b.is_game_over():
                break
            mv = b.get_moves(); b.play_move(mv[rng.randint(len(mv))])
        if b.is_game_over():
            continue
        for n_sim in sims_list:
            board = b.clone()
            net = HashNet(salt=seed)
            mct = ref_mcts.MCT(eval_method="neural", nn=net)
            for step in range(n_moves):
                if board.is_game_over():
                    break
                mct.search(board, n_sim=n_sim)
                black, white = grid_to_bits(board.grid)
                N = np.full(A, -1, dtype=np.int64); Q = np.zeros(A); P = np.zeros(A, dtype=np.float32)
                for mv, node in mct.root.children.items():
                    N[act(mv)] = node.N; Q[act(mv)] = float(node.Q); P[act(mv)] = np.float32(node.P)
                best, _ = det_fair_max(mct.root.children.items(), key=lambda x: x[1].N)
                rows.append(dict(root=root_idx, n_sim=n_sim, step=step, black=black, white=white, player=board.player,
                                 N=N, Q=Q, P=P, rootN=mct.root.N, action=act(best), evals=net.n_evals))
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
    np.savez_compressed(os.path.join(HERE, "mct_visits_ttt.npz"), **out)
    print(f"tictactoe mct: {len(rows)} searches")


def net_fixture(pos, seed):
    torch.manual_seed(seed)
    net = TicTacToeNet(device="cpu")
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.running_mean.normal_(0.0, 0.2); m.running_var.uniform_(0.5, 1.5)
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0.0, 0.2)
    live = np.nonzero(pos["over"] == 0)[0]
    grids = []
    for i in live:
        g = np.zeros((3, 3))
        for r in range(3):
            for c in range(3):
                bit = r * 8 + c
                g[r, c] = ((int(pos["black"][i]) >> bit) & 1) - ((int(pos["white"][i]) >> bit) & 1)
        grids.append(int(pos["player"][i]) * g)
    x = torch.tensor(np.stack(grids), dtype=torch.float32)
    probs, v = net.predict(x)
    sd = {k: t.detach().cpu().numpy() for k, t in net.state_dict().items()}
    np.savez_compressed(os.path.join(HERE, "tictactoe_net.npz"), index=live.astype(np.int32), input=x.numpy(),
                        probs=probs.numpy(), value=v.numpy(), **{"sd." + k: a for k, a in sd.items()})
    print(f"tictactoe net: {len(live)} positions")


if __name__ == "__main__":
    pos = positions()
    table = {"perft_1_to_9": [perft(TicTacToeBoard(), d) for d in range(1, 10)],
             "convention": "a finished game counts as one leaf; computed by the unmodified reference TicTacToeBoard"}
    json.dump(table, open(os.path.join(HERE, "tictactoe_misc.json"), "w"), indent=1)
    print(table)
    mct_fixture(n_roots=12, n_moves=4, sims_list=(1, 2, 25, 100), seed=29)
    net_fixture(pos, seed=31)
