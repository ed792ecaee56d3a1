"""TEST/BENCH INFRASTRUCTURE: the CPU baseline of the hot path, timed on the host cores.

What is timed is the oracle port of the reference's neural-mode search: the sequential pointer-tree MCT of
oracle/az_oracle.c (restating alphazero/mcts.py) with the leaf evaluation done exactly as the reference does it --
one batch-1 float32 PyTorch forward on the CPU per simulation (PolicyValueNetwork.evaluate, base.py:357-367),
reached from C through a ctypes callback.  The tree and board code is C here while the reference's is Python, so
this baseline is FASTER than the reference itself (127 sims/s/process for the Python path).
One process per core (torch threads = 1 each), every process runs a self-play game with tree reuse.
"""
import ctypes
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

EVAL_FN = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                           ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float))


def _make_eval(net, n):
    import torch
    A = n * n + 1

    def cb(ctx, own, opp, n_, probs, v):
        grid = np.zeros((n, n), dtype=np.float32)
        for r in range(n):
            for c in range(n):
                bit = r * 8 + c
                if (own >> bit) & 1:
                    grid[r, c] = 1.0
                elif (opp >> bit) & 1:
                    grid[r, c] = -1.0
        p, val = net.predict(torch.from_numpy(grid))
        p = p.reshape(-1).numpy()
        for a in range(A):
            probs[a] = float(p[a])
        v[0] = float(val.item())

    return EVAL_FN(cb)


def _worker(args):
    """Plays self-play moves (search n_sim, play the most visited move, re-root) until `moves` moves were searched
    `repeats` times; returns (sims, seconds) per repeat, measured around the searches only."""
    wid, n, n_sim, moves, repeats, seed = args
    import torch
    torch.set_num_threads(1)
    import oracle_lib as orc
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(0)
    net = OthelloNet(n=n)
    cb = _make_eval(net, n)
    L = orc.lib()
    out = []
    for rep in range(repeats):
        sims, t0, game = 0, time.time(), 0
        h = None
        for _ in range(moves):
            if h is None or orc.terminal(black, white, player, n)[0][0]:
                if h is not None:
                    L.orc_mct_free(h)
                game += 1
                h = L.orc_mct_new(n, 1, 1, 0, seed * 1000003 + wid * 101 + rep * 17 + game,
                                  ctypes.cast(cb, ctypes.c_void_p), None)
                L.orc_mct_set_dirichlet(h, 0.03, 0.25)
                b, w = orc.start(n)
                black, white, player = np.array([b], np.uint64), np.array([w], np.uint64), np.array([1], np.int8)
            L.orc_mct_search(h, int(black[0]), int(white[0]), int(player[0]), n_sim)
            sims += n_sim
            a = L.orc_mct_best_action(h)
            L.orc_mct_change_root(h, a)
            black, white, player, _, _ = orc.apply(black, white, player, np.array([a], np.uint8), n)
        out.append((sims, time.time() - t0))
        L.orc_mct_free(h)
    return out


def run(n: int = 8, n_sim: int = 200, moves: int = 4, repeats: int = 1, processes: int = None, seed: int = 0):
    """Returns a list over repeats of (total_sims, wall_seconds) with all processes running concurrently."""
    import oracle_lib
    oracle_lib.build()
    processes = processes or os.cpu_count()
    ctx = mp.get_context("spawn")
    with ctx.Pool(processes) as pool:
        t0 = time.time()
        res = pool.map(_worker, [(i, n, n_sim, moves, repeats, seed) for i in range(processes)])
        wall = time.time() - t0
    per_rep = []
    for rep in range(repeats):
        sims = sum(r[rep][0] for r in res)
        secs = max(r[rep][1] for r in res)
        per_rep.append((sims, secs))
    return per_rep, processes, wall


if __name__ == "__main__":
    reps, procs, wall = run(moves=2, repeats=2)
    for sims, secs in reps:
        print(f"{procs} processes: {sims} sims in {secs:.2f}s = {sims / secs:.1f} sims/s")
    print("wall", wall)
