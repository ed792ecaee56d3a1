"""Replays the random walks of tests/golden/make_golden_digest.py with any board backend (the C oracle or the CUDA
kernels) and returns the SHA-256 of the same per-position records the reference produced."""
import hashlib

import numpy as np

A_LCG, C_LCG = np.uint64(6364136223846793005), np.uint64(1442695040888963407)
GOLD = 0x9E3779B97F4A7C15
MASK = (1 << 64) - 1
REC = np.dtype([("black", "<u8"), ("white", "<u8"), ("player", "i1"), ("legal", "<u8"), ("over", "u1"),
                ("score", "<i2"), ("winner", "i1")])          # packed: 29 bytes, the generator's struct "<QQbQBhb"
assert REC.itemsize == 29


def _lcg(state: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        return state * A_LCG + C_LCG


def _nth_set_bit(mask: np.ndarray, k: np.ndarray) -> np.ndarray:
    """Index of the k-th (0-based) set bit of every 64-bit mask."""
    bits = ((mask[:, None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.int64)
    csum = np.cumsum(bits, axis=1)
    hit = (csum == (k[:, None] + 1)) & (bits == 1)
    return hit.argmax(axis=1)


def walk_digest(seed: int, games: int, start, legal_fn, apply_fn, terminal_fn, pass_action=64, max_plies=140):
    """start = (black, white) ints; legal_fn(b, w, p) -> uint64 action sets; apply_fn(b, w, p, action u8) -> (b, w, p);
    terminal_fn(b, w, p) -> (over u8, score, winner).  Returns (sha256 hex, positions)."""
    G = games
    black = np.full(G, start[0], np.uint64)
    white = np.full(G, start[1], np.uint64)
    player = np.ones(G, np.int8)
    state = _lcg(np.array([(seed + j * GOLD) & MASK for j in range(G)], dtype=np.uint64))
    alive = np.ones(G, bool)
    recs = np.zeros((max_plies, G), dtype=REC)
    valid = np.zeros((max_plies, G), bool)
    for ply in range(max_plies):
        if not alive.any():
            break
        legal = legal_fn(black, white, player)
        over, score, winner = terminal_fn(black, white, player)
        r = recs[ply]
        r["black"], r["white"], r["player"], r["legal"] = black, white, player, legal
        r["over"], r["score"] = over, score
        r["winner"] = np.where(over != 0, winner, 0)
        valid[ply] = alive
        go = alive & (over == 0)
        n_legal = np.array([bin(int(x)).count("1") for x in legal], dtype=np.uint64)
        nxt = _lcg(state)
        state = np.where(go, nxt, state)
        k = ((nxt >> np.uint64(33)) % np.maximum(n_legal, np.uint64(1))).astype(np.int64)
        action = np.where(n_legal > 0, _nth_set_bit(legal, k), pass_action).astype(np.uint8)
        nb, nw, npl = apply_fn(black, white, player, action)
        black, white, player = np.where(go, nb, black), np.where(go, nw, white), np.where(go, npl, player)
        alive = go
    assert not alive.any(), "walk did not finish"
    order = valid.T                                            # [G, plies]: game after game, ply after ply
    data = recs.T[order]
    return hashlib.sha256(data.tobytes()).hexdigest(), int(order.sum())
