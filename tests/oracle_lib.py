"""TEST INFRASTRUCTURE: builds (gcc) and binds oracle/az_oracle.c -- the CPU checker for the CUDA kernels."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "oracle", "az_oracle.c")
SO = os.path.join(ROOT, "oracle", "libaz_oracle.so")

_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.isfile(SO) or os.path.getmtime(SO) < os.path.getmtime(SRC):
        subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", SO, SRC, "-lm"], check=True)
    return SO


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        vp, i32, i64, u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64
        L.orc_othello_legal.argtypes = [vp, vp, vp, i32, i64, vp]
        L.orc_othello_apply.argtypes = [vp, vp, vp, vp, i32, i64, vp, vp, vp, vp, vp]
        L.orc_othello_terminal.argtypes = [vp, vp, vp, i32, i64, vp, vp, vp]
        L.orc_othello_perft.restype = u64
        L.orc_othello_perft.argtypes = [u64, u64, i32, i32, i32]
        L.orc_othello_start.argtypes = [i32, vp, vp]
        L.orc_philox4x32_10.argtypes = [vp, vp, vp]
        L.orc_othello_playouts.argtypes = [vp, vp, vp, i32, i64, i64, u64, u64, vp, vp, vp, vp]
        L.orc_hash_eval.argtypes = [u64, u64, i32, u64, vp, vp]
        L.orc_mct_new.restype = vp
        L.orc_mct_new.argtypes = [i32, i32, i32, u64, u64, vp, vp]
        L.orc_mct_set_dirichlet.argtypes = [vp, ctypes.c_double, ctypes.c_double]
        L.orc_mct_free.argtypes = [vp]
        L.orc_mct_search.argtypes = [vp, u64, u64, i32, i32]
        L.orc_mct_change_root.argtypes = [vp, i32]
        L.orc_mct_root_stats.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orc_mct_best_action.argtypes = [vp]
        L.orc_mct_best_action.restype = i32
        L.orc_connect4_legal.argtypes = [vp, vp, vp, i32, i32, i64, vp]
        L.orc_connect4_apply.argtypes = [vp, vp, vp, vp, i32, i32, i64, vp, vp, vp, vp]
        L.orc_connect4_terminal.argtypes = [vp, vp, vp, i32, i32, i64, vp, vp, vp]
        L.orc_connect4_perft.restype = u64
        L.orc_connect4_perft.argtypes = [u64, u64, i32, i32, i32, i32]
        L.orc_connect4_playouts.argtypes = [vp, vp, vp, i32, i32, i64, i64, u64, u64, vp, vp, vp, vp]
        L.orc_tictactoe_legal.argtypes = [vp, vp, vp, i64, vp]
        L.orc_tictactoe_apply.argtypes = [vp, vp, vp, vp, i64, vp, vp, vp, vp]
        L.orc_tictactoe_terminal.argtypes = [vp, vp, vp, i64, vp, vp, vp]
        L.orc_tictactoe_perft.restype = u64
        L.orc_tictactoe_perft.argtypes = [u64, u64, i32, i32]
        L.orc_tictactoe_playouts.argtypes = [vp, vp, vp, i64, i64, u64, u64, vp, vp, vp, vp]
        L.orc_mct_new_tictactoe.restype = vp
        L.orc_mct_new_tictactoe.argtypes = [i32, i32, u64, u64, vp, vp]
        L.orc_mct_new_connect4.restype = vp
        L.orc_mct_new_connect4.argtypes = [i32, i32, i32, i32, u64, u64, vp, vp]
        L.orc_mct_set_leaves.argtypes = [vp, i32]
        L.orc_greedy_moves.argtypes = [i32, vp, vp, vp, i32, i32, i64, vp, vp]
        _lib = L
    return _lib


def P(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def start(n):
    b, w = ctypes.c_uint64(), ctypes.c_uint64()
    lib().orc_othello_start(n, ctypes.byref(b), ctypes.byref(w))
    return int(b.value), int(w.value)


def legal(black, white, player, n):
    out = np.zeros(len(black), np.uint64)
    lib().orc_othello_legal(P(black), P(white), P(player), n, len(black), P(out))
    return out


def apply(black, white, player, action, n):
    G = len(black)
    ob, ow = np.zeros(G, np.uint64), np.zeros(G, np.uint64)
    op, st, fl = np.zeros(G, np.int8), np.zeros(G, np.uint8), np.zeros(G, np.uint64)
    lib().orc_othello_apply(P(black), P(white), P(player), P(action), n, G, P(ob), P(ow), P(op), P(fl), P(st))
    return ob, ow, op, fl, st


def terminal(black, white, player, n):
    G = len(black)
    t, s, w = np.zeros(G, np.uint8), np.zeros(G, np.int8), np.zeros(G, np.int8)
    lib().orc_othello_terminal(P(black), P(white), P(player), n, G, P(t), P(s), P(w))
    return t, s, w


def perft(n, depth, black=None, white=None, player=1):
    if black is None:
        black, white = start(n)
    return int(lib().orc_othello_perft(black, white, player, n, depth))


def playouts(black, white, player, n, per_root, seed, first_id=0):
    T = len(black) * per_root
    fb, fw = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
    win, pl = np.zeros(T, np.int8), np.zeros(T, np.uint8)
    lib().orc_othello_playouts(P(black), P(white), P(player), n, len(black), per_root, seed, first_id,
                               P(fb), P(fw), P(win), P(pl))
    return fb, fw, win, pl


def greedy_moves(game, black, white, player, n, h=None):
    """GreedyPlayer.get_move for every position: (lowest-index best action uint8 [G], set of best actions uint64 [G]);
    game 0 Othello (n), 1 Connect4 (n = width, h = height), 2 TicTacToe."""
    G = len(black)
    best, ties = np.zeros(G, np.uint8), np.zeros(G, np.uint64)
    lib().orc_greedy_moves(game, P(np.ascontiguousarray(black, np.uint64)), P(np.ascontiguousarray(white, np.uint64)),
                           P(np.ascontiguousarray(player, np.int8)), n, n if h is None else h, G, P(best), P(ties))
    return best, ties


def hash_eval(own, opp, action_size, salt):
    probs = np.zeros(action_size, np.float32)
    v = np.zeros(1, np.float32)
    lib().orc_hash_eval(own, opp, action_size, salt, P(probs), P(v))
    return probs, float(v[0])


# ---- Connect4 (boards are absolute black/white bitboards on the canvas, bit r*8+c, row 0 = top) ----

def c4_legal(black, white, player, width, height):
    out = np.zeros(len(black), np.uint64)
    lib().orc_connect4_legal(P(black), P(white), P(player), width, height, len(black), P(out))
    return out


def c4_apply(black, white, player, action, width, height):
    G = len(black)
    ob, ow = np.zeros(G, np.uint64), np.zeros(G, np.uint64)
    op, st = np.zeros(G, np.int8), np.zeros(G, np.uint8)
    lib().orc_connect4_apply(P(black), P(white), P(player), P(action), width, height, G, P(ob), P(ow), P(op), P(st))
    return ob, ow, op, st


def c4_terminal(black, white, player, width, height):
    G = len(black)
    t, s, w = np.zeros(G, np.uint8), np.zeros(G, np.int8), np.zeros(G, np.int8)
    lib().orc_connect4_terminal(P(black), P(white), P(player), width, height, G, P(t), P(s), P(w))
    return t, s, w


def c4_perft(width, height, depth, black=0, white=0, player=1):
    return int(lib().orc_connect4_perft(black, white, player, width, height, depth))


def c4_playouts(black, white, player, width, height, per_root, seed, first_id=0):
    T = len(black) * per_root
    fb, fw = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
    win, pl = np.zeros(T, np.int8), np.zeros(T, np.uint8)
    lib().orc_connect4_playouts(P(black), P(white), P(player), width, height, len(black), per_root, seed, first_id,
                                P(fb), P(fw), P(win), P(pl))
    return fb, fw, win, pl


# ---- TicTacToe (3 x 3 on the canvas, action = canvas cell r*8+c) ----

def ttt_legal(black, white, player):
    out = np.zeros(len(black), np.uint64)
    lib().orc_tictactoe_legal(P(black), P(white), P(player), len(black), P(out))
    return out


def ttt_apply(black, white, player, action):
    G = len(black)
    ob, ow = np.zeros(G, np.uint64), np.zeros(G, np.uint64)
    op, st = np.zeros(G, np.int8), np.zeros(G, np.uint8)
    lib().orc_tictactoe_apply(P(black), P(white), P(player), P(action), G, P(ob), P(ow), P(op), P(st))
    return ob, ow, op, st


def ttt_terminal(black, white, player):
    G = len(black)
    t, s, w = np.zeros(G, np.uint8), np.zeros(G, np.int8), np.zeros(G, np.int8)
    lib().orc_tictactoe_terminal(P(black), P(white), P(player), G, P(t), P(s), P(w))
    return t, s, w


def ttt_perft(depth, black=0, white=0, player=1):
    return int(lib().orc_tictactoe_perft(black, white, player, depth))


def ttt_playouts(black, white, player, per_root, seed, first_id=0):
    T = len(black) * per_root
    fb, fw = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
    win, pl = np.zeros(T, np.int8), np.zeros(T, np.uint8)
    lib().orc_tictactoe_playouts(P(black), P(white), P(player), len(black), per_root, seed, first_id,
                                 P(fb), P(fw), P(win), P(pl))
    return fb, fw, win, pl


class OracleMCT:
    """Sequential pointer-tree MCT of the oracle (restates alphazero/mcts.py).  `n` = Othello board size, or
    (width, height) with game="connect4"."""

    def __init__(self, n, mode="neural", tie_random=False, salt=0, seed=0, dirichlet=None, game="othello", leaves=None):
        if game == "tictactoe":
            self.h = lib().orc_mct_new_tictactoe(1 if mode == "neural" else 0, int(tie_random), salt, seed, None, None)
        elif game == "connect4":
            self.h = lib().orc_mct_new_connect4(n[0], n[1], 1 if mode == "neural" else 0, int(tie_random), salt, seed,
                                                None, None)
        else:
            self.h = lib().orc_mct_new(n, 1 if mode == "neural" else 0, int(tie_random), salt, seed, None, None)
        if dirichlet:
            lib().orc_mct_set_dirichlet(self.h, dirichlet[0], dirichlet[1])
        if leaves:      # throughput mode (extension): up to `leaves` leaves in flight per step under virtual loss
            lib().orc_mct_set_leaves(self.h, int(leaves))

    def search(self, black, white, player, n_sim):
        lib().orc_mct_search(self.h, int(black), int(white), int(player), int(n_sim))

    def change_root(self, action):
        lib().orc_mct_change_root(self.h, int(action))

    def stats(self):
        N, Q, Pp = np.zeros(65, np.int64), np.zeros(65), np.zeros(65)
        rn, ne, nr = np.zeros(1, np.int64), np.zeros(1, np.int64), np.zeros(1, np.int64)
        lib().orc_mct_root_stats(self.h, P(N), P(Q), P(Pp), P(rn), P(ne), P(nr))
        return N, Q, Pp, int(rn[0]), int(ne[0]), int(nr[0])

    def best_action(self):
        return int(lib().orc_mct_best_action(self.h))

    def __del__(self):
        try:
            lib().orc_mct_free(self.h)
        except Exception:
            pass
