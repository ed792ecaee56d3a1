"""Win / loss / draw counts of the UNMODIFIED reference Arena for a seeded AlphaZeroPlayer, the fixture behind the
arena win-rate parity test (BASELINE.json north_star: "arena win-rate parity must hold statistically").

    python tests/golden/make_golden_arena.py [games_per_match]      (build container only; minutes on 8 cores)

Matches (Othello 6x6, alphazero/arena.py:119-185 play_games with alternating starts, contests.py:27-42 style):
    AlphaZeroPlayer(n_sim=25, nn=seeded OthelloNet(n=6), no Dirichlet noise) vs RandomPlayer
    AlphaZeroPlayer(same)                                                    vs MCTSPlayer(n_sim=25)
The games are split over worker processes, each seeding NumPy's global RNG (the reference never seeds it, T7).
Output (committed): tests/golden/arena_othello6.json with the counts, the network seed and the SHA-256 of the
network's state_dict bytes -- the GPU test rebuilds the network from the seed and checks the hash before it plays.
"""
import hashlib
import json
import os
import sys
import time
from multiprocessing import Pool

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

NET_SEED, N_SIM, BOARD = 1234, 25, 6


def state_dict_sha256(sd) -> str:
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().numpy()).tobytes())
    return h.hexdigest()


def worker(args):
    opponent, n_rounds, seed = args
    from ref_import import import_reference
    import_reference()
    from alphazero.arena import Arena
    from alphazero.games.othello import OthelloBoard, OthelloNet
    from alphazero.players import AlphaZeroPlayer, MCTSPlayer, RandomPlayer
    torch.set_num_threads(1)
    torch.manual_seed(NET_SEED)
    net = OthelloNet(n=BOARD, device="cpu")
    net.eval()
    np.random.seed(seed)
    p1 = AlphaZeroPlayer(n_sim=N_SIM, nn=net)
    p2 = RandomPlayer() if opponent == "random" else MCTSPlayer(n_sim=N_SIM)
    stats = Arena(p1, p2, OthelloBoard(n=BOARD)).play_games(n_rounds, return_stats=True)
    return opponent, len(stats["player1"]), len(stats["player2"]), stats["draw"]


if __name__ == "__main__":
    games = int(sys.argv[1]) if len(sys.argv) > 1 else 240
    procs = os.cpu_count()
    per = max(2, (games // procs) // 2 * 2)
    t0 = time.time()
    jobs = [(opp, per, 1000 * (i + 1) + (0 if opp == "random" else 7)) for opp in ("random", "mcts") for i in range(procs)]
    with Pool(procs) as pool:
        res = pool.map(worker, jobs, chunksize=1)
    from ref_import import import_reference
    import_reference()
    from alphazero.games.othello import OthelloNet
    torch.manual_seed(NET_SEED)
    sha = state_dict_sha256(OthelloNet(n=BOARD, device="cpu").state_dict())
    out = {"board": BOARD, "n_sim": N_SIM, "net_seed": NET_SEED, "net_state_dict_sha256": sha, "matches": {}}
    for opp in ("random", "mcts"):
        rows = [r for r in res if r[0] == opp]
        out["matches"][opp] = {"games": per * procs, "player1_wins": sum(r[1] for r in rows),
                               "player2_wins": sum(r[2] for r in rows), "draws": sum(r[3] for r in rows)}
        print(opp, out["matches"][opp])
    json.dump(out, open(os.path.join(HERE, "arena_othello6.json"), "w"), indent=1)
    print(f"done in {time.time() - t0:.0f}s")
