"""Othello for the B200 engine: OthelloConfig / OthelloBoard / OthelloNet with the reference's interface
(alphazero/games/othello.py:17-450).

OthelloBoard keeps the position as two 64-bit bitboards (bit r*8+c) and answers every rule query by calling the
CUDA kernels through the C ABI (alphazero_b200.boardops) -- the same kernels the batched engine uses; there is no
Python or NumPy implementation of the rules in this package.  OthelloNet is the reference's architecture with the
reference's parameter names, so its state_dicts load here and vice versa.
"""
import dataclasses

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import boardops as bo
from ..base import Board, Config, PolicyValueNetwork


@dataclasses.dataclass
class OthelloConfig(Config):
    """Defaults of alphazero/games/othello.py:17-47."""
    game: str = "othello"
    board_size: int = 6
    simulations: int = 100
    compute_time: float = None
    dirichlet_alpha: float = 0.03
    dirichlet_epsilon: float = 0.25
    temp_scheduler_type: str = "linear"
    temp_max_step: int = 4
    temp_min_step: int = 4
    iterations: int = 30
    episodes: int = 200
    epochs: int = 10
    batch_size: int = 64
    learning_rate: float = 0.01
    data_augmentation: bool = True
    device: str = "cpu"
    eval_opponent: str = "mcts"
    eval_episodes: int = 40
    do_eval: bool = True
    save: bool = True
    push: bool = False
    save_checkpoints: bool = True
    push_checkpoints: bool = False


def _dev1(x: int) -> torch.Tensor:
    return torch.tensor([bo.to_i64(x)], dtype=torch.int64, device="cuda")


class OthelloBoard(Board):
    """n x n Othello (n even, <= 8); +1 = black (moves first), -1 = white, moves are (row, col), pass is (n, n)."""

    CONFIG = OthelloConfig
    DIRECTIONS = [(1, 1), (1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1)]

    def __init__(self, n: int = None, grid: np.ndarray = None, player: int = 1, display_dir: str = None,
                 display_mode: str = None, config: Config = None):
        super().__init__(display_dir=display_dir, display_mode=display_mode)
        self.game = "othello"
        self.n = config.board_size if config is not None else n
        if self.n is None or self.n % 2 != 0:
            raise ValueError(f"Board size must be even but got n={self.n}")
        if self.n > 8:
            raise ValueError("the bitboard engine supports n <= 8")
        self.reset()
        if config is None:
            self.player = player
            if grid is not None:
                self.grid = grid

    # ---- state -------------------------------------------------------------------------------------------
    def reset(self) -> None:
        self.black, self.white = bo.start_position(self.n)
        self.player = 1
        self.pass_move = (self.n, self.n)
        self.max_moves = self.n * self.n - 4

    @property
    def grid(self) -> np.ndarray:
        """float64 [n, n] view of the bitboards (a fresh array; the reference's callers only read or copy it)."""
        g = np.zeros((self.n, self.n))
        for r in range(self.n):
            for c in range(self.n):
                bit = r * 8 + c
                if (self.black >> bit) & 1:
                    g[r, c] = 1
                elif (self.white >> bit) & 1:
                    g[r, c] = -1
        return g

    @grid.setter
    def grid(self, value: np.ndarray) -> None:
        value = np.asarray(value)
        if value.shape != (self.n, self.n):
            raise ValueError(f"grid must have shape {(self.n, self.n)}")
        self.black = self.white = 0
        for r in range(self.n):
            for c in range(self.n):
                if value[r, c] == 1:
                    self.black |= 1 << (r * 8 + c)
                elif value[r, c] == -1:
                    self.white |= 1 << (r * 8 + c)

    def __str__(self) -> str:
        return f"{type(self).__name__}{self.n}"

    def clone(self) -> "OthelloBoard":
        b = OthelloBoard.__new__(OthelloBoard)
        b.__dict__.update(self.__dict__)
        return b

    def get_board_shape(self):
        return (self.n, self.n)

    def get_n_cells(self) -> int:
        return self.n * self.n

    def get_action_size(self) -> int:
        return self.n * self.n + 1

    # ---- helpers -----------------------------------------------------------------------------------------
    def _side(self, player=None):
        """(own, opp) bitboards for `player` (anything but +1/-1 means the side to move, othello.py:143)."""
        p = player if player in (-1, 1) else self.player
        return (self.black, self.white) if p == 1 else (self.white, self.black)

    def _legal_mask(self, player=None) -> int:
        own, opp = self._side(player)
        return bo.to_u64(int(bo.legal(_dev1(own), _dev1(opp), self.n).item()))

    def _action(self, move) -> int:
        if move == self.pass_move:
            return bo.PASS
        r, c = int(move[0]), int(move[1])
        if not (0 <= r < self.n and 0 <= c < self.n):
            return 255
        return r * 8 + c

    # ---- rules (all through the CUDA kernels) ------------------------------------------------------------
    def get_score(self) -> int:
        own, opp = self._side()
        _, score, _ = bo.terminal_score(_dev1(own), _dev1(opp), torch.tensor([self.player], dtype=torch.int8, device="cuda"), self.n)
        return int(score.item())

    def is_legal_move(self, move, player: int = None) -> bool:
        mask = self._legal_mask(player)
        a = self._action(move)
        if a == bo.PASS:
            return mask == 0
        return a < 64 and bool((mask >> a) & 1)

    def get_moves(self, player: int = None) -> list:
        mask = self._legal_mask(player)
        moves = [(a // 8, a % 8) for a in range(64) if (mask >> a) & 1]
        return moves if moves else [self.pass_move]

    def get_random_move(self, player: int = None):
        moves = self.get_moves(player)
        return moves[np.random.choice(len(moves))]

    def play_move(self, move) -> None:
        a = self._action(move)
        own, opp = self._side()
        o2, p2, pl2, st = bo.apply(_dev1(own), _dev1(opp), torch.tensor([self.player], dtype=torch.int8, device="cuda"),
                                   torch.tensor([a], dtype=torch.uint8, device="cuda"), self.n)
        if int(st.item()) != 0:
            raise ValueError(f"Illegal move {move} for player {self.player}")
        o2, p2 = bo.to_u64(int(o2.item())), bo.to_u64(int(p2.item()))
        self.player = int(pl2.item())
        self.black, self.white = (o2, p2) if self.player == 1 else (p2, o2)

    def is_game_over(self) -> bool:
        own, opp = self._side()
        term, _, _ = bo.terminal_score(_dev1(own), _dev1(opp), torch.tensor([self.player], dtype=torch.int8, device="cuda"), self.n)
        return bool(term.item())

    def get_winner(self) -> int:
        own, opp = self._side()
        term, _, winner = bo.terminal_score(_dev1(own), _dev1(opp), torch.tensor([self.player], dtype=torch.int8, device="cuda"), self.n)
        if not bool(term.item()):
            raise ValueError("Game is not over yet...")
        return int(winner.item())

    def human_display(self, *args, **kwargs):
        raise NotImplementedError("board rendering (matplotlib PNG) is out of scope of the B200 engine")

    def pixel_display(self, *args, **kwargs):
        raise NotImplementedError("board rendering (PIL PNG) is out of scope of the B200 engine")


class OthelloNet(PolicyValueNetwork):
    """4 x (conv3x3, 32 ch, BN, ReLU) [pads 1,1,0,0] -> fc 1024 -> fc 512 (BN1d, ReLU, dropout .3) -> heads:
    log_softmax(n*n+1), tanh(1).  Same layer names as alphazero/games/othello.py:341-359."""

    CONFIG = OthelloConfig

    def __init__(self, n: int = None, device: str = None, config: Config = None):
        super().__init__()
        if config is not None:
            self.n = config.board_size
            self.device = PolicyValueNetwork.get_torch_device(config.device)
        else:
            self.n = n
            self.device = PolicyValueNetwork.get_torch_device(device)
        if self.n is None:
            raise ValueError("The board size must be a positive and even integer like 4, 6 or 8.")
        self.action_size = self.n * self.n + 1
        self.n_channels = 32
        self.dropout = 0.3
        ch, dev = self.n_channels, self.device
        self.conv1 = nn.Conv2d(1, ch, 3, stride=1, padding=1, device=dev)
        self.conv2 = nn.Conv2d(ch, ch, 3, stride=1, padding=1, device=dev)
        self.conv3 = nn.Conv2d(ch, ch, 3, stride=1, device=dev)
        self.conv4 = nn.Conv2d(ch, ch, 3, stride=1, device=dev)
        self.bn1 = nn.BatchNorm2d(ch, device=dev)
        self.bn2 = nn.BatchNorm2d(ch, device=dev)
        self.bn3 = nn.BatchNorm2d(ch, device=dev)
        self.bn4 = nn.BatchNorm2d(ch, device=dev)
        self.fc1_input_size = ch * (self.n - 4) * (self.n - 4)
        self.fc1 = nn.Linear(self.fc1_input_size, 1024, device=dev)
        self.fc_bn1 = nn.BatchNorm1d(1024, device=dev)
        self.fc2 = nn.Linear(1024, 512, device=dev)
        self.fc_bn2 = nn.BatchNorm1d(512, device=dev)
        self.fc_probs = nn.Linear(512, self.action_size, device=dev)
        self.fc_value = nn.Linear(512, 1, device=dev)

    def forward(self, input: torch.Tensor):
        x = input.view(-1, 1, self.n, self.n)
        x = F.relu(self.bn1(self.conv1(x)))
        x = F.relu(self.bn2(self.conv2(x)))
        x = F.relu(self.bn3(self.conv3(x)))
        x = F.relu(self.bn4(self.conv4(x)))
        x = x.reshape(-1, self.fc1_input_size)
        x = F.dropout(F.relu(self.fc_bn1(self.fc1(x))), p=self.dropout, training=self.training)
        x = F.dropout(F.relu(self.fc_bn2(self.fc2(x))), p=self.dropout, training=self.training)
        return F.log_softmax(self.fc_probs(x), dim=1), torch.tanh(self.fc_value(x))

    def _index(self, move) -> int:
        return self.action_size - 1 if tuple(move) == (self.n, self.n) else move[0] * self.n + move[1]

    def get_normalized_probs(self, probs: np.ndarray, legal_moves: list) -> dict:
        """Priors restricted to the legal moves and renormalised (uniform if their mass is < 1e-6),
        alphazero/games/othello.py:384-402."""
        picked = {move: probs[self._index(move)] for move in legal_moves}
        total = 0
        for p in picked.values():
            total += p
        if total < 1e-6:
            return {move: 1 / len(legal_moves) for move in legal_moves}
        return {move: p / total for move, p in picked.items()}

    def to_neural_output(self, move_probs: dict) -> np.ndarray:
        pi = np.zeros(self.action_size)
        for move, prob in move_probs.items():
            pi[self._index(move)] = prob
        return pi

    def _split(self, neural_output: np.ndarray):
        if neural_output.size != self.action_size:
            raise ValueError(f"Neural output should have size {self.action_size}, but has size {neural_output.size}")
        return neural_output[:-1].reshape(self.n, self.n), neural_output[-1:]

    def reflect_neural_output(self, neural_output: np.ndarray, axis: int) -> np.ndarray:
        board, tail = self._split(neural_output)
        return np.concatenate([np.flip(board, axis=axis).reshape(-1), tail])

    def rotate_neural_output(self, neural_output: np.ndarray, angle: int) -> np.ndarray:
        board, tail = self._split(neural_output)
        return np.concatenate([np.rot90(board, k=angle // 90).reshape(-1), tail])
