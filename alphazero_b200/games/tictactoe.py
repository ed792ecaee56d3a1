"""TicTacToe for the B200 engine: TicTacToeConfig / TicTacToeBoard / TicTacToeNet with the reference's interface
(alphazero/games/tictactoe.py:17-368).

TicTacToeBoard keeps the position as two bitboards on the engine's canvas (bit r*8+c, r, c < 3) and answers every rule
query through the CUDA kernels (gamespec.TicTacToeSpec); there is no Python or NumPy implementation of the rules here.
TicTacToeNet is the reference's architecture with the reference's parameter names (state_dicts are interchangeable).
"""
import dataclasses

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..base import Board, Config, PolicyValueNetwork
from ..boardops import to_i64, to_u64
from ..gamespec import TicTacToeSpec


@dataclasses.dataclass
class TicTacToeConfig(Config):
    """Defaults of alphazero/games/tictactoe.py:17-47."""
    game: str = "tictactoe"
    board_size: int = 3
    simulations: int = 100
    compute_time: float = None
    dirichlet_alpha: float = 0.03
    dirichlet_epsilon: float = 0.25
    temp_scheduler_type: str = "linear"
    temp_max_step: int = 2
    temp_min_step: int = 2
    iterations: int = 30
    episodes: int = 100
    epochs: int = 10
    batch_size: int = 64
    learning_rate: float = 0.01
    data_augmentation: bool = True
    device: str = "cpu"
    eval_opponent: str = "mcts"
    eval_episodes: int = 100
    do_eval: bool = True
    save: bool = True
    push: bool = False
    save_checkpoints: bool = True
    push_checkpoints: bool = False


def _dev1(x: int) -> torch.Tensor:
    return torch.tensor([to_i64(x)], dtype=torch.int64, device="cuda")


def _p1(player: int) -> torch.Tensor:
    return torch.tensor([player], dtype=torch.int8, device="cuda")


class TicTacToeBoard(Board):
    """3 x 3 board; +1 (X) moves first, moves are (row, col) tuples, there is no pass move."""

    CONFIG = TicTacToeConfig
    DIRECTIONS = [(1, 1), (1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1)]
    _spec = None

    def __init__(self, grid: np.ndarray = None, player: int = 1, display_dir: str = None, display_mode: str = None,
                 config: Config = None):
        super().__init__(display_dir=display_dir, display_mode=display_mode)
        self.game = "tictactoe"
        if TicTacToeBoard._spec is None:
            TicTacToeBoard._spec = TicTacToeSpec()
        self.reset()
        if config is None:
            self.player = player
            if grid is not None:
                self.grid = grid

    def reset(self) -> None:
        self.black = self.white = 0
        self.player = 1
        self.max_moves = 9

    @property
    def grid(self) -> np.ndarray:
        g = np.zeros((3, 3))
        for r in range(3):
            for c in range(3):
                bit = r * 8 + c
                g[r, c] = ((self.black >> bit) & 1) - ((self.white >> bit) & 1)
        return g

    @grid.setter
    def grid(self, value: np.ndarray) -> None:
        value = np.asarray(value)
        if value.shape != (3, 3):
            raise ValueError("grid must have shape (3, 3)")
        self.black = self.white = 0
        for r in range(3):
            for c in range(3):
                if value[r, c] == 1:
                    self.black |= 1 << (r * 8 + c)
                elif value[r, c] == -1:
                    self.white |= 1 << (r * 8 + c)

    def __str__(self) -> str:
        return f"{type(self).__name__}"

    def clone(self) -> "TicTacToeBoard":
        b = TicTacToeBoard.__new__(TicTacToeBoard)
        b.__dict__.update(self.__dict__)
        return b

    def get_board_shape(self):
        return (3, 3)

    def get_n_cells(self) -> int:
        return 9

    def get_action_size(self) -> int:
        return 9

    def _side(self):
        return (self.black, self.white) if self.player == 1 else (self.white, self.black)

    def _legal_mask(self) -> int:
        own, opp = self._side()
        return to_u64(int(self._spec.legal(_dev1(own), _dev1(opp)).item()))

    def _terminal(self):
        own, opp = self._side()
        term, score, winner = self._spec.terminal_score(_dev1(own), _dev1(opp), _p1(self.player))
        return bool(term.item()), int(score.item()), int(winner.item())

    def get_score(self):
        """float("inf") when the side to move can complete an alignment, else 0 (tictactoe.py:127-133)."""
        return float("inf") if self._terminal()[1] == 127 else 0

    def is_legal_move(self, move, player: int = None) -> bool:
        r, c = int(move[0]), int(move[1])
        return 0 <= r < 3 and 0 <= c < 3 and bool((self._legal_mask() >> (r * 8 + c)) & 1)

    def get_moves(self, player: int = None) -> list:
        mask = self._legal_mask()
        return [(r, c) for r in range(3) for c in range(3) if (mask >> (r * 8 + c)) & 1]

    def get_random_move(self, player: int = None):
        moves = self.get_moves(player)
        return moves[np.random.choice(len(moves))]

    def play_move(self, move) -> None:
        own, opp = self._side()
        action = self._spec.move_to_action(move)
        o2, p2, pl2, st = self._spec.apply(_dev1(own), _dev1(opp), _p1(self.player),
                                           torch.tensor([action], dtype=torch.uint8, device="cuda"))
        if int(st.item()) != 0:
            raise ValueError(f"Illegal move {move} for player {self.player}")
        o2, p2 = to_u64(int(o2.item())), to_u64(int(p2.item()))
        self.player = int(pl2.item())
        self.black, self.white = (o2, p2) if self.player == 1 else (p2, o2)

    def is_game_over(self) -> bool:
        return self._terminal()[0]

    def get_winner(self) -> int:
        term, _, winner = self._terminal()
        if not term:
            raise ValueError("Game is not over yet...")
        return winner

    def human_display(self, *args, **kwargs):
        raise NotImplementedError("board rendering (matplotlib PNG) is out of scope of the B200 engine")

    def pixel_display(self, *args, **kwargs):
        raise NotImplementedError("board rendering (PIL PNG) is out of scope of the B200 engine")


class TicTacToeNet(PolicyValueNetwork):
    """flatten -> fc 9 -> fc 9 (BatchNorm1d, ReLU) -> heads: log_softmax(9), tanh(1); the layer names of
    alphazero/games/tictactoe.py:289-296."""

    CONFIG = TicTacToeConfig

    def __init__(self, device: str = None, config: Config = None):
        super().__init__()
        self.device = PolicyValueNetwork.get_torch_device(config.device if config is not None else device)
        self.action_size = 9
        dev = self.device
        self.fc1 = nn.Linear(9, 9, device=dev)
        self.fc2 = nn.Linear(9, 9, device=dev)
        self.fc_probs = nn.Linear(9, self.action_size, device=dev)
        self.fc_value = nn.Linear(9, 1, device=dev)
        self.flatten = nn.Flatten()
        self.bn1 = nn.BatchNorm1d(9, device=dev)
        self.bn2 = nn.BatchNorm1d(9, device=dev)

    def forward(self, input: torch.Tensor):
        if input.ndim == 2:
            input = input.unsqueeze(0)
        x = self.flatten(input)
        x = F.relu(self.bn1(self.fc1(x)))
        x = F.relu(self.bn2(self.fc2(x)))
        return F.log_softmax(self.fc_probs(x), dim=1), torch.tanh(self.fc_value(x))

    def get_normalized_probs(self, probs: np.ndarray, legal_moves: list) -> dict:
        """Priors of the empty cells, renormalised (uniform if their mass is < 1e-6), tictactoe.py:318-334."""
        total = 0
        legal_probs = {}
        for move in legal_moves:
            legal_probs[move] = probs[3 * move[0] + move[1]]
            total += legal_probs[move]
        if total < 1e-6:
            return {move: 1 / len(legal_moves) for move in legal_moves}
        return {move: prob / total for move, prob in legal_probs.items()}

    def to_neural_output(self, move_probs: dict) -> np.ndarray:
        pi = np.zeros(9)
        for move, prob in move_probs.items():
            pi[3 * move[0] + move[1]] = prob
        return pi

    def reflect_neural_output(self, neural_output: np.ndarray, axis: int) -> np.ndarray:
        if not neural_output.size == self.action_size:
            raise ValueError(f"Neural output should have size {self.action_size}, but has size {neural_output.size}")
        return np.flip(neural_output.reshape((3, 3)), axis=axis).flatten()

    def rotate_neural_output(self, neural_output: np.ndarray, angle: int) -> np.ndarray:
        if not neural_output.size == self.action_size:
            raise ValueError(f"Neural output should have size {self.action_size}, but has size {neural_output.size}")
        return np.rot90(neural_output.reshape((3, 3)), k=angle // 90).flatten()
