"""Game descriptors for the batched engine: which rule kernels (C ABI, include/az_b200.h) a TreeBatch /
SelfPlayEngine / batched Arena calls for its boards, and how device actions map to the reference's Python moves.

The reference expresses the same thing through the Board base class (alphazero/base.py:95-225); here a spec is a
small value object so that the batched code is written once for Othello (games/othello.py) and Connect4
(games/connect4.py).  Positions are always side-relative bitboards on the 8-wide canvas (bit r*8+c).
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


def _to_u64(x: int) -> int:
    return int(x) & ((1 << 64) - 1)


class GameSpec:
    """Common surface; subclasses bind the game's kernels."""

    game: str = None
    game_id: int = None
    n: int = None         # az_tree.n  (Othello: board size; Connect4: width)
    height: int = None    # az_tree.height (rows)
    rows: int = None
    cols: int = None
    action_size: int = None
    max_moves: int = None     # Board.max_moves (feeds the temperature schedulers)
    max_plies: int = None     # upper bound on the plies of one game (history buffers)
    mean_branching: float = None

    # ---- helpers shared by both games
    def _outs(self, own):
        G = own.numel()
        return (torch.empty(G, dtype=torch.uint8, device=own.device), torch.empty(G, dtype=torch.int8, device=own.device),
                torch.empty(G, dtype=torch.int8, device=own.device))

    def greedy_moves(self, own, opp, mask=None, seed: int = 0, ctr: int = 0, tie_random: bool = True, want_ties: bool = False):
        """GreedyPlayer.get_move for every position (players.py:98-123, az_greedy_move): uint8 actions (255 where
        mask is 0); with want_ties also the uint64 sets of the maximising actions."""
        G = own.numel()
        action = torch.empty(G, dtype=torch.uint8, device=own.device)
        ties = torch.empty(G, dtype=torch.int64, device=own.device) if want_ties else None
        m = None if mask is None else mask.to(torch.uint8).contiguous()
        _lib.check(_lib.lib().az_greedy_move(self.game_id, _ptr(own), _ptr(opp), _ptr(m), self.n, self.height, G,
                                             _to_u64(seed), int(ctr) & 0xFFFFFFFF, 1 if tie_random else 0, _ptr(action),
                                             _ptr(ties), _stream()), "az_greedy_move")
        return (action, ties) if want_ties else action

    def __eq__(self, other):
        return type(self) is type(other) and (self.n, self.height) == (other.n, other.height)

    def __hash__(self):
        return hash((self.game, self.n, self.height))


class OthelloSpec(GameSpec):
    """alphazero/games/othello.py: n x n, actions 0..63 = cell r*8+c, 64 = pass (the reference's (n, n))."""

    game, game_id = "othello", _lib.GAME_OTHELLO

    def __init__(self, n: int):
        if n % 2 != 0 or n < 4 or n > 8:
            raise ValueError(f"Board size must be even and in [4, 8] but got n={n}")
        self.n = self.height = self.rows = self.cols = int(n)
        self.action_size = n * n + 1
        self.max_moves = n * n - 4
        self.max_plies = 2 * n * n          # placements + passes
        self.mean_branching = 10.0

    def __repr__(self):
        return f"OthelloSpec(n={self.n})"

    def start_position(self):
        """(black, white) of OthelloBoard.__get_init_board (othello.py:102-109) as Python ints."""
        h = self.n // 2
        black = (1 << ((h - 1) * 8 + (h - 1))) | (1 << (h * 8 + h))
        white = (1 << ((h - 1) * 8 + h)) | (1 << (h * 8 + (h - 1)))
        return black, white

    def legal(self, own, opp):
        out = torch.empty_like(own)
        _lib.check(_lib.lib().az_othello_legal(_ptr(own), _ptr(opp), self.n, own.numel(), _ptr(out), _stream()),
                   "az_othello_legal")
        return out

    def apply(self, own, opp, player, action):
        o2, p2, pl2 = torch.empty_like(own), torch.empty_like(opp), torch.empty_like(player)
        st = torch.empty(own.numel(), dtype=torch.uint8, device=own.device)
        _lib.check(_lib.lib().az_othello_apply(_ptr(own), _ptr(opp), _ptr(player), _ptr(action), self.n, own.numel(),
                                               _ptr(o2), _ptr(p2), _ptr(pl2), _ptr(st), _stream()), "az_othello_apply")
        return o2, p2, pl2, st

    def terminal_score(self, own, opp, player):
        term, score, winner = self._outs(own)
        _lib.check(_lib.lib().az_othello_terminal_score(_ptr(own), _ptr(opp), _ptr(player), self.n, own.numel(),
                                                        _ptr(term), _ptr(score), _ptr(winner), _stream()),
                   "az_othello_terminal_score")
        return term, score, winner

    def encode(self, own, opp, dtype=torch.float32):
        out = torch.empty((own.numel(), self.n, self.n), dtype=dtype, device=own.device)
        code = {torch.float32: 0, torch.bfloat16: 1}[dtype]
        _lib.check(_lib.lib().az_othello_encode(_ptr(own), _ptr(opp), self.n, own.numel(), _ptr(out), code, _stream()),
                   "az_othello_encode")
        return out

    def legal_action_bits(self, legal):
        """[G, 64] 0/1 float matrix of the legal actions + a bool column 'has any' (pass rows have none)."""
        bits = ((legal.unsqueeze(1) >> torch.arange(64, device=legal.device)) & 1).to(torch.float32)
        return bits, PASS

    def action_to_move(self, a: int):
        return (self.n, self.n) if a == PASS else (a // 8, a % 8)

    def move_to_action(self, move) -> int:
        if tuple(move) == (self.n, self.n):
            return PASS
        r, c = int(move[0]), int(move[1])
        return r * 8 + c if 0 <= r < self.n and 0 <= c < self.n else 255


class Connect4Spec(GameSpec):
    """alphazero/games/connect4.py: width x height, actions = column index, no pass."""

    game, game_id = "connect4", _lib.GAME_CONNECT4

    def __init__(self, width: int, height: int):
        if width < 4 or height < 4:
            raise ValueError(f"Borad size must be at least 4x4, got {width}x{height}")
        if width > 8 or height > 8:
            raise ValueError("the bitboard engine supports Connect4 boards up to 8x8")
        self.n = self.cols = int(width)
        self.height = self.rows = int(height)
        self.action_size = int(width)
        self.max_moves = width * height
        self.max_plies = width * height
        self.mean_branching = float(width)

    def __repr__(self):
        return f"Connect4Spec(width={self.n}, height={self.height})"

    def start_position(self):
        return 0, 0

    def legal(self, own, opp):
        out = torch.empty_like(own)
        _lib.check(_lib.lib().az_connect4_legal(_ptr(own), _ptr(opp), self.n, self.height, own.numel(), _ptr(out),
                                                _stream()), "az_connect4_legal")
        return out

    def apply(self, own, opp, player, action):
        o2, p2, pl2 = torch.empty_like(own), torch.empty_like(opp), torch.empty_like(player)
        st = torch.empty(own.numel(), dtype=torch.uint8, device=own.device)
        _lib.check(_lib.lib().az_connect4_apply(_ptr(own), _ptr(opp), _ptr(player), _ptr(action), self.n, self.height,
                                                own.numel(), _ptr(o2), _ptr(p2), _ptr(pl2), _ptr(st), _stream()),
                   "az_connect4_apply")
        return o2, p2, pl2, st

    def terminal_score(self, own, opp, player):
        term, score, winner = self._outs(own)
        _lib.check(_lib.lib().az_connect4_terminal_score(_ptr(own), _ptr(opp), _ptr(player), self.n, self.height,
                                                         own.numel(), _ptr(term), _ptr(score), _ptr(winner), _stream()),
                   "az_connect4_terminal_score")
        return term, score, winner

    def encode(self, own, opp, dtype=torch.float32):
        out = torch.empty((own.numel(), self.height, self.n), dtype=dtype, device=own.device)
        code = {torch.float32: 0, torch.bfloat16: 1}[dtype]
        _lib.check(_lib.lib().az_connect4_encode(_ptr(own), _ptr(opp), self.n, self.height, own.numel(), _ptr(out), code,
                                                 _stream()), "az_connect4_encode")
        return out

    def legal_action_bits(self, legal):
        bits = ((legal.unsqueeze(1) >> torch.arange(64, device=legal.device)) & 1).to(torch.float32)
        return bits, 0

    def action_to_move(self, a: int):
        return np.int64(a)          # the reference's moves are np.int64 columns (connect4.py:162)

    def move_to_action(self, move) -> int:
        c = int(move)
        return c if 0 <= c < self.n else 255

    def playouts(self, own, opp, player, per_root: int, seed: int, first_id: int = 0):
        R = own.numel()
        T = R * per_root
        fb = torch.empty(T, dtype=torch.int64, device=own.device)
        fw = torch.empty(T, dtype=torch.int64, device=own.device)
        win = torch.empty(T, dtype=torch.int8, device=own.device)
        pl = torch.empty(T, dtype=torch.uint8, device=own.device)
        _lib.check(_lib.lib().az_connect4_playouts(_ptr(own), _ptr(opp), _ptr(player), self.n, self.height, R, per_root,
                                                   _to_u64(seed), _to_u64(first_id), _ptr(fb), _ptr(fw), _ptr(win),
                                                   _ptr(pl), _stream()), "az_connect4_playouts")
        return fb, fw, win, pl

    def perft(self, depth: int, own: int = 0, opp: int = 0, dfs_depth: int = 3, device="cuda",
              scratch_entries: int = 1 << 22) -> int:
        scratch = torch.empty(scratch_entries, dtype=torch.int64, device=device)
        count = ctypes.c_uint64(0)
        _lib.check(_lib.lib().az_connect4_perft(_to_u64(own), _to_u64(opp), self.n, self.height, depth,
                                                min(dfs_depth, depth), _ptr(scratch), scratch.numel(),
                                                ctypes.byref(count), _stream()), "az_connect4_perft")
        return int(count.value)


class TicTacToeSpec(GameSpec):
    """alphazero/games/tictactoe.py: 3 x 3, actions = canvas cell r*8+c (the reference's (r, c)), no pass."""

    game, game_id = "tictactoe", _lib.GAME_TICTACTOE

    def __init__(self):
        self.n = self.height = self.rows = self.cols = 3
        self.action_size = 9
        self.max_moves = 9
        self.max_plies = 9
        self.mean_branching = 5.0

    def __repr__(self):
        return "TicTacToeSpec()"

    def start_position(self):
        return 0, 0

    def legal(self, own, opp):
        out = torch.empty_like(own)
        _lib.check(_lib.lib().az_tictactoe_legal(_ptr(own), _ptr(opp), own.numel(), _ptr(out), _stream()),
                   "az_tictactoe_legal")
        return out

    def apply(self, own, opp, player, action):
        o2, p2, pl2 = torch.empty_like(own), torch.empty_like(opp), torch.empty_like(player)
        st = torch.empty(own.numel(), dtype=torch.uint8, device=own.device)
        _lib.check(_lib.lib().az_tictactoe_apply(_ptr(own), _ptr(opp), _ptr(player), _ptr(action), own.numel(),
                                                 _ptr(o2), _ptr(p2), _ptr(pl2), _ptr(st), _stream()), "az_tictactoe_apply")
        return o2, p2, pl2, st

    def terminal_score(self, own, opp, player):
        term, score, winner = self._outs(own)
        _lib.check(_lib.lib().az_tictactoe_terminal_score(_ptr(own), _ptr(opp), _ptr(player), own.numel(), _ptr(term),
                                                          _ptr(score), _ptr(winner), _stream()),
                   "az_tictactoe_terminal_score")
        return term, score, winner

    def encode(self, own, opp, dtype=torch.float32):
        out = torch.empty((own.numel(), 3, 3), dtype=dtype, device=own.device)
        code = {torch.float32: 0, torch.bfloat16: 1}[dtype]
        _lib.check(_lib.lib().az_tictactoe_encode(_ptr(own), _ptr(opp), own.numel(), _ptr(out), code, _stream()),
                   "az_tictactoe_encode")
        return out

    def legal_action_bits(self, legal):
        bits = ((legal.unsqueeze(1) >> torch.arange(64, device=legal.device)) & 1).to(torch.float32)
        return bits, 0

    def action_to_move(self, a: int):
        return (a // 8, a % 8)

    def move_to_action(self, move) -> int:
        r, c = int(move[0]), int(move[1])
        return r * 8 + c if 0 <= r < 3 and 0 <= c < 3 else 255

    def playouts(self, own, opp, player, per_root: int, seed: int, first_id: int = 0):
        R = own.numel()
        T = R * per_root
        fb = torch.empty(T, dtype=torch.int64, device=own.device)
        fw = torch.empty(T, dtype=torch.int64, device=own.device)
        win = torch.empty(T, dtype=torch.int8, device=own.device)
        pl = torch.empty(T, dtype=torch.uint8, device=own.device)
        _lib.check(_lib.lib().az_tictactoe_playouts(_ptr(own), _ptr(opp), _ptr(player), R, per_root, _to_u64(seed),
                                                    _to_u64(first_id), _ptr(fb), _ptr(fw), _ptr(win), _ptr(pl),
                                                    _stream()), "az_tictactoe_playouts")
        return fb, fw, win, pl


def spec_of(board) -> GameSpec:
    """Spec of a drop-in Board object (games/othello.py, games/connect4.py)."""
    game = getattr(board, "game", None)
    if game == "othello":
        return OthelloSpec(board.n)
    if game == "connect4":
        return Connect4Spec(board.width, board.height)
    if game == "tictactoe":
        return TicTacToeSpec()
    raise NotImplementedError(f"the B200 engine supports Othello, Connect4 and TicTacToe boards, got {game!r}")


def as_spec(n_or_spec) -> GameSpec:
    return n_or_spec if isinstance(n_or_spec, GameSpec) else OthelloSpec(int(n_or_spec))
