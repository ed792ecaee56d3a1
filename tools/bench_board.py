"""BASELINE config 2: the board kernels timed alone on one B200, against the measured integer-ALU issue
rate (az_peak_int32) and the measured HBM bandwidth.

    python tools/bench_board.py > profiles/r1_bench_board.json

Prints one JSON object: device perft(1..11) of Othello 8x8 (checked against the known table), 1 M random playouts from
the start position (Philox4x32-10, seed 0x5EED), the movegen / apply / terminal kernels over a corpus of random-walk
positions, Connect4 twins, and the int32 peaks.  Instructions per unit come from the committed ncu capture
(profiles/r1_ncu_board_summary.json, `smsp__inst_executed.sum` per launch) when present.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes

import torch

from alphazero_b200 import _lib
from alphazero_b200 import boardops as bo
from alphazero_b200.gamespec import Connect4Spec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KNOWN_PERFT = [4, 12, 56, 244, 1396, 8200, 55092, 390216, 3005288, 24571284, 212258800]


def timed(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3      # seconds per call


def random_walk_corpus(count, plies_max, seed):
    """Positions after k ~ U{0..plies_max} uniformly random legal moves from the Othello 8x8 start (passes included)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    b, w = bo.start_position(8)
    own = torch.full((count,), bo.to_i64(b), dtype=torch.int64, device="cuda")
    opp = torch.full((count,), bo.to_i64(w), dtype=torch.int64, device="cuda")
    player = torch.ones(count, dtype=torch.int8, device="cuda")
    stop = torch.randint(0, plies_max + 1, (count,), generator=g, device="cuda")
    shifts = torch.arange(64, device="cuda")
    for ply in range(plies_max):
        legal = bo.legal(own, opp, 8)
        bits = ((legal.unsqueeze(1) >> shifts) & 1).to(torch.float32)
        has = bits.sum(1) > 0
        bits[~has, 0] = 1.0
        a = torch.multinomial(bits, 1, generator=g).flatten()
        action = torch.where(has, a, torch.full_like(a, 64)).to(torch.uint8)
        o2, p2, pl2, _ = bo.apply(own, opp, player, action, 8)
        term, _, _ = bo.terminal_score(own, opp, player, 8)
        move = (stop > ply) & (term == 0)
        own, opp, player = torch.where(move, o2, own), torch.where(move, p2, opp), torch.where(move, pl2, player)
    return own, opp, player


def ncu_inst(kernel_substr):
    path = os.path.join(ROOT, "profiles", "r1_ncu_board_summary.json")
    if not os.path.isfile(path):
        return None
    for k in json.load(open(path)):
        if kernel_substr in k["kernel"] and "smsp__inst_executed.sum" in k:
            return k["smsp__inst_executed.sum"]["value"]
    return None


def ncu_pass():
    """One launch of each board kernel between cudaProfilerStart/Stop, for
    `ncu --profile-from-start off --metrics ... python tools/bench_board.py --ncu` (same corpus as the timed run)."""
    base = random_walk_corpus(1 << 20, 56, 7)
    own, opp, player = (x.repeat(16) for x in base)
    legal1 = bo.legal(base[0], base[1], 8)
    first = ((legal1.unsqueeze(1) >> torch.arange(64, device="cuda")) & 1).to(torch.uint8).argmax(1)
    action = torch.where(legal1 != 0, first, torch.full_like(first, 64)).to(torch.uint8).repeat(16)
    b, w = bo.start_position(8)
    o1 = torch.tensor([bo.to_i64(b)], dtype=torch.int64, device="cuda")
    p1 = torch.tensor([bo.to_i64(w)], dtype=torch.int64, device="cuda")
    pl = torch.ones(1, dtype=torch.int8, device="cuda")
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    bo.legal(own, opp, 8)
    bo.apply(own, opp, player, action, 8)
    bo.terminal_score(own, opp, player, 8)
    bo.playouts(o1, p1, pl, 8, 1_000_000, 0x5EED)
    bo.perft(8, 9, dfs_depth=3)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()


def main():
    torch.cuda.init()
    L = _lib.lib()
    from tools.probes import build as _probes      # measurement-only kernels live outside the product library
    PL = _probes.lib()
    if "--ncu" in sys.argv:
        return ncu_pass()
    out = {"device": torch.cuda.get_device_name(0)}
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = None
    if os.path.isfile(peaks_path):
        pk = json.load(open(peaks_path))
        hbm = pk.get("hbm_gbps") or pk.get("hbm_GBps") or next((v for k, v in pk.items() if "hbm" in k.lower() and isinstance(v, (int, float))), None)
    out["hbm_peak_gbps"] = hbm

    # ---- integer issue-rate peaks
    sink = torch.empty(148 * 8 * 256, dtype=torch.int32, device="cuda")
    peaks = {}
    for kind, name in enumerate(("lop3", "shf", "iadd3", "shift_mask_or")):
        ops = ctypes.c_int64(0)
        call = lambda: _lib.check(PL.az_peak_int32(kind, 4096, sink.data_ptr(), sink.numel(), ctypes.byref(ops),
                                                  torch.cuda.current_stream().cuda_stream), "az_peak_int32")
        t = timed(call, iters=10)
        peaks[name] = {"warp_inst_per_s": sink.numel() / 32 * ops.value / t, "ms": t * 1e3}
    out["int32_peaks"] = peaks
    int_peak = peaks["lop3"]["warp_inst_per_s"]

    # ---- perft
    perft = {}
    for d in range(1, 12):
        n = bo.perft(8, d, dfs_depth=min(4, d))
        assert n == KNOWN_PERFT[d - 1], (d, n)
        perft[d] = n
    t = timed(lambda: bo.perft(8, 11, dfs_depth=4), iters=3, warmup=1)
    out["perft"] = {"table_1_to_11": perft, "depth11_seconds": t, "leaf_nodes_per_s": KNOWN_PERFT[10] / t,
                    "check": "identical to the known Othello perft table"}

    # ---- 1 M playouts from the start position
    b, w = bo.start_position(8)
    own = torch.tensor([bo.to_i64(b)], dtype=torch.int64, device="cuda")
    opp = torch.tensor([bo.to_i64(w)], dtype=torch.int64, device="cuda")
    pl = torch.ones(1, dtype=torch.int8, device="cuda")
    total = 1_000_000
    res = {}
    t = timed(lambda: res.update(r=bo.playouts(own, opp, pl, 8, total, 0x5EED)), iters=5)
    plies = float(res["r"][3].to(torch.float64).mean().item())
    out["playouts"] = {"count": total, "seconds": t, "playouts_per_s": total / t, "mean_plies": plies,
                       "plies_per_s": total * plies / t, "seed": "0x5EED"}

    # ---- movegen / apply / terminal over a random-walk corpus (16 B in, 8 B out per board for movegen)
    base = random_walk_corpus(1 << 20, 56, 7)
    rep = 16
    own, opp, player = (x.repeat(rep) for x in base)
    G = own.numel()
    legal1 = bo.legal(base[0], base[1], 8)
    first = ((legal1.unsqueeze(1) >> torch.arange(64, device="cuda")) & 1).to(torch.uint8).argmax(1)   # lowest legal cell
    action = torch.where(legal1 != 0, first, torch.full_like(first, 64)).to(torch.uint8).repeat(rep)
    kern = {}
    for name, fn, bytes_per in (
            ("k_othello_legal", lambda: bo.legal(own, opp, 8), 24),
            ("k_othello_apply", lambda: bo.apply(own, opp, player, action, 8), 16 + 2 + 16 + 2),
            ("k_othello_terminal", lambda: bo.terminal_score(own, opp, player, 8), 17 + 3)):
        t = timed(fn, iters=10)
        e = {"boards": G, "seconds": t, "boards_per_s": G / t, "hbm_gbps": G * bytes_per / t / 1e9}
        if hbm:
            e["hbm_frac"] = e["hbm_gbps"] / hbm
        inst = ncu_inst(name)
        if inst:      # the ncu capture ran this script: same corpus, same launch
            e["warp_inst_per_launch"] = inst
            e["warp_inst_per_board"] = inst * 32 / G
            e["issue_frac_of_int_peak"] = inst / t / int_peak
        kern[name] = e
    out["board_kernels"] = kern

    # ---- Connect4 7x6
    spec = Connect4Spec(7, 6)
    z = torch.zeros(G, dtype=torch.int64, device="cuda")
    t = timed(lambda: spec.legal(z, z), iters=10)
    z1 = torch.zeros(1, dtype=torch.int64, device="cuda")
    res = {}
    tp = timed(lambda: res.update(r=spec.playouts(z1, z1, pl, total, 0x5EED)), iters=5)
    out["connect4"] = {"legal_boards_per_s": G / t, "playouts_per_s": total / tp,
                       "mean_plies": float(res["r"][3].to(torch.float64).mean().item())}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
