"""Deterministic "hash evaluator" used for MCT visit-count identity tests.

It stands in for PolicyValueNetwork.evaluate (alphazero/base.py:357-367): from the canonical position
(own = discs of the side to move, opp = the other side; i.e. what the network sees as player*grid) it derives
  * raw probabilities p[a] = q/4096 with q in 1..32 for every network action index a (r*n+c, pass = n*n):
    exact in float32, and any partial sum is exact, so summation order cannot matter;
  * a network value v in {-256..255}/256 from the side to move's viewpoint (0 exercises the draw clamp).
The same function is implemented in oracle/az_oracle.c (orc_hash_eval) and in the CUDA library
(az_eval_hash); this numpy version feeds the unmodified Python reference when fixtures are generated.
"""
import numpy as np

M64 = (1 << 64) - 1


def mix64(z: int) -> int:
    z &= M64
    z ^= z >> 30
    z = (z * 0xBF58476D1CE4E5B9) & M64
    z ^= z >> 27
    z = (z * 0x94D049BB133111EB) & M64
    z ^= z >> 31
    return z


def board_hash(own: int, opp: int, salt: int = 0) -> int:
    return mix64((own * 0x9E3779B97F4A7C15 + mix64(opp + 0xD1B54A32D192ED03) + salt) & M64)


def hash_eval(own: int, opp: int, action_size: int, salt: int = 0):
    """Returns (probs float32[action_size], v float) for the canonical position."""
    h = board_hash(own, opp, salt)
    probs = np.empty(action_size, dtype=np.float32)
    for a in range(action_size):
        q = ((mix64((h + a * 0x632BE59BD9B4E019) & M64) >> 40) & 31) + 1
        probs[a] = np.float32(q) / np.float32(4096.0)
    v = (((h >> 20) & 0x1FF) - 256) / 256.0
    return probs, v


def grid_to_bits(grid: np.ndarray):
    """(black, white) bitboards on the 8-wide canvas from a reference grid (+1 black, -1 white)."""
    black = white = 0
    n_r, n_c = grid.shape
    for r in range(n_r):
        for c in range(n_c):
            if grid[r][c] == 1:
                black |= 1 << (r * 8 + c)
            elif grid[r][c] == -1:
                white |= 1 << (r * 8 + c)
    return black, white


def bits_to_grid(black: int, white: int, n: int) -> np.ndarray:
    g = np.zeros((n, n))
    for r in range(n):
        for c in range(n):
            b = r * 8 + c
            if (black >> b) & 1:
                g[r][c] = 1
            elif (white >> b) & 1:
                g[r][c] = -1
    return g
