"""TEST / MEASUREMENT INFRASTRUCTURE: times the UNMODIFIED Python reference (/root/reference) on this machine's cores.

    python oracle/time_reference.py [seconds_per_measurement]      (build container only: the reference cannot travel)

Writes profiles/r2_python_reference_cpu.json:
  * neural mode: MCT("neural", OthelloNet(n=8) random-init).search(board, n_sim=200) from the start position, one
    process and one process per core (alphazero/mcts.py:254-269) -- simulations / s;
  * BASELINE config 1: Arena(MCTSPlayer(n_sim=100), RandomPlayer(), OthelloBoard(8)) games (alphazero/arena.py:119-185,
    contests.py:27-42) for the time budget, one process per core -- simulations / s and games / h.
bench.py quotes these numbers beside the C-oracle baseline it can time on the GPU box (`cpu_baseline_python_reference`).
"""
import json
import os
import sys
import time
from multiprocessing import Pool

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)


def neural_worker(args):
    seed, seconds = args
    import numpy as np
    import torch
    from ref_import import import_reference
    import_reference()
    from alphazero.games.othello import OthelloBoard, OthelloNet
    from alphazero.mcts import MCT
    torch.set_num_threads(1)
    torch.manual_seed(0)
    np.random.seed(seed)
    net = OthelloNet(n=8, device="cpu")
    net.eval()
    sims, t0 = 0, time.time()
    while time.time() - t0 < seconds:
        mct = MCT(eval_method="neural", nn=net)
        mct.search(OthelloBoard(n=8), n_sim=200)
        sims += 200
    return sims, time.time() - t0


def arena_worker(args):
    seed, seconds = args
    import numpy as np
    from ref_import import import_reference
    import_reference()
    from alphazero.arena import Arena
    from alphazero.games.othello import OthelloBoard
    from alphazero.players import MCTSPlayer, RandomPlayer
    np.random.seed(seed)
    p1 = MCTSPlayer(n_sim=100)
    arena = Arena(p1, RandomPlayer(), OthelloBoard(n=8))
    games = sims = 0
    t0 = time.time()
    while time.time() - t0 < seconds:
        arena.play_game(player2_starts=bool(games % 2), return_results=True)
        games += 1
        sims += 100 * 30          # MCTSPlayer moves every other ply of a ~60-ply game
    return games, sims, time.time() - t0


if __name__ == "__main__":
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 30.0
    cores = os.cpu_count()
    out = {"host": "build container", "cores": cores, "seconds_per_measurement": seconds}
    s1, t1 = neural_worker((1, seconds))
    out["neural_1_process_sims_per_s"] = s1 / t1
    with Pool(cores) as pool:
        res = pool.map(neural_worker, [(i + 1, seconds) for i in range(cores)])
    out["neural_all_cores_sims_per_s"] = sum(r[0] / r[1] for r in res)
    with Pool(cores) as pool:
        res = pool.map(arena_worker, [(i + 1, seconds) for i in range(cores)])
    out["config1_games"] = sum(r[0] for r in res)
    out["config1_games_per_hour_all_cores"] = sum(r[0] / r[2] for r in res) * 3600
    out["config1_sims_per_s_all_cores_approx"] = sum(r[1] / r[2] for r in res)
    out["what"] = ("unmodified t0m1ab/alphazero: MCT('neural', OthelloNet(n=8)).search(n_sim=200) from the start position; "
                   "Arena(MCTSPlayer(n_sim=100), RandomPlayer(), OthelloBoard(8)) full games")
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "profiles", "r2_python_reference_cpu.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))
