"""Generates the Connect4 golden fixtures by IMPORTING THE UNMODIFIED REFERENCE (/root/reference).

Run in the build container only (the reference does not exist on the GPU box):
    python tests/golden/make_golden_connect4.py
Outputs (committed):
    connect4_positions_{w}x{h}.npz  positions reached by random play of the reference Connect4Board with, for each:
                                    playable-column mask (get_moves), is_game_over, get_score, get_winner (when
                                    over), one played column and the resulting grid; plus illegal-column probes.
    perft_connect4.json             perft(1..d) of the reference Connect4Board (a finished game counts as one leaf).
    mct_visits_c4_{w}x{h}.npz       alphazero.mcts.MCT("neural") visit counts / Q / P of the root's children under
                                    the deterministic hash evaluator (tests/hash_eval.py) and the deterministic
                                    tie-break (fair_max -> lowest column), with tree reuse across moves.
    connect4_net_7x6.npz            a seeded reference Connect4Net: its state_dict, 256 input positions and the fp32
                                    outputs of PolicyValueNetwork.predict (pins the (w,h) view quirk, connect4.py:399).
"""
import json
import os
import sys
import time

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
from alphazero.games.connect4 import Connect4Board, Connect4Net  # noqa: E402

U64 = ("black", "white", "legal", "black2", "white2")


def positions(width, height, count, seed):
    rng = np.random.RandomState(seed)
    np.random.seed(seed)
    rec = {k: [] for k in ("black", "white", "player", "legal", "over", "score", "winner", "action",
                           "black2", "white2", "player2")}
    while len(rec["black"]) < count:
        b = Connect4Board(width=width, height=height)
        while True:
            black, white = grid_to_bits(b.grid)
            moves = [int(m) for m in b.get_moves()]
            legal = 0
            for m in moves:
                legal |= 1 << m
            over = b.is_game_over()
            keep = over or rng.rand() < 0.5
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
                rec["action"].append(mv); rec["black2"].append(b2); rec["white2"].append(w2)
                rec["player2"].append(b.player)
    out = {k: np.array(v[:count], dtype=np.uint64 if k in U64 else np.int16) for k, v in rec.items()}
    np.savez_compressed(os.path.join(HERE, f"connect4_positions_{width}x{height}.npz"), **out)
    print(f"connect4 positions {width}x{height}: {count} ({int(out['over'].sum())} finished)")


def perft(b, d):
    if d == 0 or b.is_game_over():
        return 1
    t = 0
    for m in b.get_moves():
        c = b.clone(); c.play_move(m); t += perft(c, d - 1)
    return t


class HashNet:
    """Fake PolicyValueNetwork for the reference MCT: evaluate() per alphazero/base.py:357-367 semantics."""

    def __init__(self, width, salt):
        self.action_size = width
        self.salt = salt
        self.n_evals = 0

    def evaluate(self, board):
        black, white = grid_to_bits(board.grid)
        own, opp = (black, white) if board.player == 1 else (white, black)
        probs, v_net = hash_eval(own, opp, self.action_size, self.salt)
        self.n_evals += 1
        return probs, board.player * float(np.float32(v_net))

    def get_normalized_probs(self, probs, legal_moves):
        return Connect4Net.get_normalized_probs(self, probs, legal_moves)


def det_fair_max(elements, key=lambda x: x):
    """Deterministic stand-in for alphazero.utils.fair_max: ties -> lowest column."""
    elements = list(elements)
    best = max(key(x) for x in elements)
    ties = [x for x in elements if key(x) == best]
    return min(ties, key=lambda x: int(x[0]))


def mct_fixture(width, height, n_roots, n_moves, sims_list, seed):
    ref_mcts.fair_max = det_fair_max
    rng = np.random.RandomState(seed)
    A = 65
    rows = []
    for root_idx in range(n_roots):
        b = Connect4Board(width=width, height=height)
        for _ in range(0 if root_idx == 0 else rng.randint(0, width * height - 6)):
            if b.is_game_over():
                break
            mv = b.get_moves(); b.play_move(mv[rng.randint(len(mv))])
        if b.is_game_over():
            continue
        for n_sim in sims_list:
            board = b.clone()
            net = HashNet(width, salt=seed)
            mct = ref_mcts.MCT(eval_method="neural", nn=net)
            for step in range(n_moves):
                if board.is_game_over():
                    break
                mct.search(board, n_sim=n_sim)
                black, white = grid_to_bits(board.grid)
                N = np.full(A, -1, dtype=np.int64); Q = np.zeros(A); P = np.zeros(A, dtype=np.float32)
                for mv, node in mct.root.children.items():
                    N[int(mv)] = node.N; Q[int(mv)] = float(node.Q); P[int(mv)] = np.float32(node.P)
                best, _ = det_fair_max(mct.root.children.items(), key=lambda x: x[1].N)
                rows.append(dict(root=root_idx, n_sim=n_sim, step=step, black=black, white=white,
                                 player=board.player, N=N, Q=Q, P=P, rootN=mct.root.N, action=int(best),
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
    np.savez_compressed(os.path.join(HERE, f"mct_visits_c4_{width}x{height}.npz"), **out)
    print(f"connect4 mct {width}x{height}: {len(rows)} searches")


def net_fixture(width, height, count, seed):
    torch.manual_seed(seed)
    np.random.seed(seed)
    net = Connect4Net(board_width=width, board_height=height, device="cpu")
    # make the BatchNorm statistics non-trivial so that folding them is actually exercised
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, (torch.nn.BatchNorm1d, torch.nn.BatchNorm2d)):
                m.running_mean.normal_(0.0, 0.2); m.running_var.uniform_(0.5, 1.5)
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0.0, 0.2)
    rng = np.random.RandomState(seed)
    blacks, whites, players, grids = [], [], [], []
    while len(grids) < count:
        b = Connect4Board(width=width, height=height)
        for _ in range(rng.randint(0, width * height)):
            if b.is_game_over():
                break
            mv = b.get_moves(); b.play_move(mv[rng.randint(len(mv))])
        if b.is_game_over():
            continue
        bl, wh = grid_to_bits(b.grid)
        blacks.append(bl); whites.append(wh); players.append(b.player); grids.append(b.player * b.grid)
    x = torch.tensor(np.stack(grids), dtype=torch.float32)
    probs, v = net.predict(x)
    # evaluate() on single boards must agree with the batched predict (same module, eval mode)
    sd = {k: t.detach().cpu().numpy() for k, t in net.state_dict().items()}
    np.savez_compressed(os.path.join(HERE, f"connect4_net_{width}x{height}.npz"),
                        black=np.array(blacks, np.uint64), white=np.array(whites, np.uint64),
                        player=np.array(players, np.int8), input=x.numpy(), probs=probs.numpy(), value=v.numpy(),
                        **{"sd." + k: a for k, a in sd.items()})
    print(f"connect4 net {width}x{height}: {count} positions, {sum(a.size for a in sd.values())} state_dict values")


if __name__ == "__main__":
    t0 = time.time()
    positions(7, 6, 3000, seed=4207)
    positions(4, 4, 400, seed=4204)
    positions(8, 8, 600, seed=4208)
    positions(5, 7, 400, seed=4257)
    table = {"7x6": [perft(Connect4Board(width=7, height=6), d) for d in range(1, 6)],
             "4x4": [perft(Connect4Board(width=4, height=4), d) for d in range(1, 8)],
             "5x4": [perft(Connect4Board(width=5, height=4), d) for d in range(1, 7)],
             "convention": "a finished game counts as one leaf; computed by the unmodified reference Connect4Board"}
    json.dump(table, open(os.path.join(HERE, "perft_connect4.json"), "w"), indent=1)
    print("perft", table)
    mct_fixture(7, 6, n_roots=10, n_moves=3, sims_list=(1, 2, 50, 200), seed=17)
    mct_fixture(5, 4, n_roots=6, n_moves=4, sims_list=(1, 3, 64), seed=19)
    net_fixture(7, 6, 256, seed=23)
    print(f"done in {time.time() - t0:.1f}s")
