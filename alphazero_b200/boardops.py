"""Batched Othello board operations on the GPU (thin wrappers over the C ABI, include/az_b200.h).

Positions are side-relative bitboards held in int64 CUDA tensors (bit pattern of the uint64; bit r*8+c):
`own` = discs of the side to move, `opp` = the other side, `player` (int8) = +1 black / -1 white.
Actions are uint8: 0..63 = cell (a // 8, a % 8), 64 = pass (the reference's (n, n) move).

Each function replaces the per-board Python loops of alphazero/games/othello.py named in its docstring.
"""
import ctypes

import numpy as np
import torch

from . import _lib

PASS = 64


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _check_dev(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("alphazero_b200 kernels need CUDA tensors (there is no CPU path)")


def start_position(n: int):
    """(black, white) bitboards of OthelloBoard.__get_init_board (othello.py:102-109) as Python ints."""
    if n % 2 != 0 or n < 4 or n > 8:
        raise ValueError(f"Board size must be even and in [4, 8] but got n={n}")
    h = n // 2
    black = (1 << ((h - 1) * 8 + (h - 1))) | (1 << (h * 8 + h))
    white = (1 << ((h - 1) * 8 + h)) | (1 << (h * 8 + (h - 1)))
    return black, white


def to_i64(x: int) -> int:
    """uint64 bit pattern -> Python int that fits a signed int64 tensor."""
    x &= (1 << 64) - 1
    return x - (1 << 64) if x >= (1 << 63) else x


def to_u64(x: int) -> int:
    return int(x) & ((1 << 64) - 1)


def legal(own: torch.Tensor, opp: torch.Tensor, n: int) -> torch.Tensor:
    """Legal placement mask per board (0 <=> only pass).  OthelloBoard.get_moves, othello.py:176-189."""
    _check_dev(own, opp)
    out = torch.empty_like(own)
    _lib.check(_lib.lib().az_othello_legal(_ptr(own), _ptr(opp), n, own.numel(), _ptr(out), _stream()), "az_othello_legal")
    return out


def flips(own, opp, action, n: int) -> torch.Tensor:
    """Discs flipped by `action` per board.  OthelloBoard.__get_flips, othello.py:141-153."""
    _check_dev(own, opp, action)
    out = torch.empty_like(own)
    _lib.check(_lib.lib().az_othello_flips(_ptr(own), _ptr(opp), _ptr(action), n, own.numel(), _ptr(out), _stream()),
               "az_othello_flips")
    return out


def apply(own, opp, player, action, n: int):
    """Plays `action` per board; returns (own, opp, player, status) with status 1 where the move is illegal
    (board unchanged; the reference raises ValueError).  OthelloBoard.play_move, othello.py:196-210."""
    _check_dev(own, opp, player, action)
    o2, p2, pl2 = torch.empty_like(own), torch.empty_like(opp), torch.empty_like(player)
    st = torch.empty(own.numel(), dtype=torch.uint8, device=own.device)
    _lib.check(_lib.lib().az_othello_apply(_ptr(own), _ptr(opp), _ptr(player), _ptr(action), n, own.numel(),
                                           _ptr(o2), _ptr(p2), _ptr(pl2), _ptr(st), _stream()), "az_othello_apply")
    return o2, p2, pl2, st


def terminal_score(own, opp, player, n: int):
    """(terminal uint8, score int8 = #own-#opp, winner int8 absolute id).  othello.py:133-135, 212-229."""
    _check_dev(own, opp, player)
    G = own.numel()
    term = torch.empty(G, dtype=torch.uint8, device=own.device)
    score = torch.empty(G, dtype=torch.int8, device=own.device)
    winner = torch.empty(G, dtype=torch.int8, device=own.device)
    _lib.check(_lib.lib().az_othello_terminal_score(_ptr(own), _ptr(opp), _ptr(player), n, G, _ptr(term),
                                                    _ptr(score), _ptr(winner), _stream()), "az_othello_terminal_score")
    return term, score, winner


def encode(own, opp, n: int, dtype=torch.float32) -> torch.Tensor:
    """player*grid network input [G, n, n].  PolicyValueNetwork.evaluate, base.py:363."""
    _check_dev(own, opp)
    G = own.numel()
    out = torch.empty((G, n, n), dtype=dtype, device=own.device)
    code = {torch.float32: 0, torch.bfloat16: 1}[dtype]
    _lib.check(_lib.lib().az_othello_encode(_ptr(own), _ptr(opp), n, G, _ptr(out), code, _stream()), "az_othello_encode")
    return out


def perft(n: int, depth: int, own: int = None, opp: int = None, dfs_depth: int = 3, device="cuda",
          scratch_entries: int = 1 << 22) -> int:
    """Leaf count of the game tree to `depth` (pass = ply, finished game = one leaf), computed on the device."""
    if own is None:
        own, opp = start_position(n)
    scratch = torch.empty(scratch_entries, dtype=torch.int64, device=device)
    count = ctypes.c_uint64(0)
    _lib.check(_lib.lib().az_othello_perft(to_u64(own), to_u64(opp), n, depth, min(dfs_depth, depth),
                                           _ptr(scratch), scratch.numel(), ctypes.byref(count), _stream()),
               "az_othello_perft")
    return int(count.value)


def playouts(own, opp, player, n: int, per_root: int, seed: int, first_id: int = 0):
    """Uniform random playouts: returns (final_black, final_white, winner int8, plies uint8), each
    [n_roots * per_root].  MCT.rollout + get_random_move, mcts.py:173-180 / othello.py:191-194."""
    _check_dev(own, opp, player)
    R = own.numel()
    T = R * per_root
    fb = torch.empty(T, dtype=torch.int64, device=own.device)
    fw = torch.empty(T, dtype=torch.int64, device=own.device)
    win = torch.empty(T, dtype=torch.int8, device=own.device)
    pl = torch.empty(T, dtype=torch.uint8, device=own.device)
    _lib.check(_lib.lib().az_othello_playouts(_ptr(own), _ptr(opp), _ptr(player), n, R, per_root, to_u64(seed),
                                              to_u64(first_id), _ptr(fb), _ptr(fw), _ptr(win), _ptr(pl), _stream()),
               "az_othello_playouts")
    return fb, fw, win, pl


def bits_from_numpy(a: np.ndarray, device="cuda") -> torch.Tensor:
    """uint64 numpy array -> int64 CUDA tensor with the same bits."""
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64)).to(device)


def bits_to_numpy(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().numpy().view(np.uint64)
