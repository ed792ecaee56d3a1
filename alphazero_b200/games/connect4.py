"""Connect4 for the B200 engine: Connect4Config / Connect4Board / Connect4Net with the reference's interface
(alphazero/games/connect4.py:17-445).

Connect4Board keeps the position as two 64-bit bitboards on the engine's canvas (bit r*8+c, row 0 = top) and answers
every rule query by calling the CUDA kernels through the C ABI (gamespec.Connect4Spec) -- the same kernels the batched
engine uses; there is no Python or NumPy implementation of the rules in this package.  Connect4Net is the reference's
architecture with the reference's parameter names, so its state_dicts load here and vice versa.
"""
import dataclasses

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..base import Board, Config, PolicyValueNetwork
from ..boardops import to_i64, to_u64
from ..gamespec import Connect4Spec


@dataclasses.dataclass
class Connect4Config(Config):
    """Defaults of alphazero/games/connect4.py:17-48."""
    game: str = "connect4"
    board_width: int = 7
    board_height: int = 6
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
    return torch.tensor([to_i64(x)], dtype=torch.int64, device="cuda")


def _p1(player: int) -> torch.Tensor:
    return torch.tensor([player], dtype=torch.int8, device="cuda")


class Connect4Board(Board):
    """width x height Connect4 (4 <= width, height <= 8); +1 moves first, moves are column indices (np.int64), there is
    no pass move (pass_move = None); row 0 is the top row, discs land on row free_rows[col] - 1."""

    CONFIG = Connect4Config
    DIRECTIONS = [(1, 1), (1, 0), (1, -1), (0, -1), (-1, -1), (-1, 0), (-1, 1), (0, 1)]

    def __init__(self, width: int = None, height: int = None, grid: np.ndarray = None, player: int = 1,
                 display_dir: str = None, display_mode: str = None, config: Config = None):
        super().__init__(display_dir=display_dir, display_mode=display_mode)
        self.game = "connect4"
        if config is not None:
            self.width, self.height = config.board_width, config.board_height
        else:
            self.width, self.height = width, height
        if self.width is None or self.height is None or self.width < 4 or self.height < 4:
            raise ValueError(f"Borad size must be at least 4x4, got {self.width}x{self.height}")
        self._spec = Connect4Spec(self.width, self.height)
        self.reset()
        if config is None:
            self.player = player
            if grid is not None:
                self.grid = grid

    # ---- state -------------------------------------------------------------------------------------------
    def reset(self) -> None:
        self.black = self.white = 0
        self.player = 1
        self.max_moves = self.width * self.height

    @property
    def grid(self) -> np.ndarray:
        """float64 [height, width] view of the bitboards (a fresh array)."""
        g = np.zeros((self.height, self.width))
        for r in range(self.height):
            for c in range(self.width):
                bit = r * 8 + c
                if (self.black >> bit) & 1:
                    g[r, c] = 1
                elif (self.white >> bit) & 1:
                    g[r, c] = -1
        return g

    @grid.setter
    def grid(self, value: np.ndarray) -> None:
        value = np.asarray(value)
        if value.shape != (self.height, self.width):
            raise ValueError(f"grid must have shape {(self.height, self.width)}")
        self.black = self.white = 0
        for r in range(self.height):
            for c in range(self.width):
                if value[r, c] == 1:
                    self.black |= 1 << (r * 8 + c)
                elif value[r, c] == -1:
                    self.white |= 1 << (r * 8 + c)

    @property
    def free_rows(self) -> np.ndarray:
        """Number of empty cells per column (connect4.py:111-113)."""
        occ = self.black | self.white
        return np.array([self.height - sum((occ >> (r * 8 + c)) & 1 for r in range(self.height))
                         for c in range(self.width)], dtype=np.int64)

    def __str__(self) -> str:
        return f"{type(self).__name__}{self.width}x{self.height}"

    def clone(self) -> "Connect4Board":
        b = Connect4Board.__new__(Connect4Board)
        b.__dict__.update(self.__dict__)
        return b

    def get_board_shape(self):
        return (self.height, self.width)

    def get_n_cells(self) -> int:
        return self.width * self.height

    def get_action_size(self) -> int:
        return self.width

    # ---- rules (all through the CUDA kernels) ------------------------------------------------------------
    def _side(self):
        return (self.black, self.white) if self.player == 1 else (self.white, self.black)

    def _legal_mask(self) -> int:
        own, opp = self._side()
        return to_u64(int(self._spec.legal(_dev1(own), _dev1(opp)).item()))

    def _terminal(self):
        own, opp = self._side()
        term, score, winner = self._spec.terminal_score(_dev1(own), _dev1(opp), _p1(self.player))
        return bool(term.item()), int(score.item()), int(winner.item())

    def get_score(self) -> int:
        return self._terminal()[1]

    def is_legal_move(self, move, player: int = None) -> bool:
        if not 0 <= move < self.width:
            raise ValueError(f"Column index must be in [0, {self.width-1}], got {move}")
        return bool((self._legal_mask() >> int(move)) & 1)

    def get_moves(self, player: int = None) -> list:
        mask = self._legal_mask()
        return [np.int64(c) for c in range(self.width) if (mask >> c) & 1]

    def get_random_move(self, player: int = None):
        moves = self.get_moves(player)
        return moves[np.random.choice(len(moves))]

    def play_move(self, move) -> None:
        if not 0 <= move < self.width:
            raise ValueError(f"Column index must be in [0, {self.width-1}], got {move}")
        own, opp = self._side()
        o2, p2, pl2, st = self._spec.apply(_dev1(own), _dev1(opp), _p1(self.player),
                                           torch.tensor([int(move)], dtype=torch.uint8, device="cuda"))
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


class Connect4Net(PolicyValueNetwork):
    """4 x (conv3x3, 32 ch, BN, ReLU) [pads 1,1,0,0] -> fc 64 -> fc 32 (BN1d, ReLU, dropout .3) -> heads:
    log_softmax(width), tanh(1).  Same layer names as alphazero/games/connect4.py:369-390.  The input is VIEWED as
    (width, height) although the grid is (height, width) -- the reference's reinterpretation (connect4.py:399), kept."""

    CONFIG = Connect4Config

    def __init__(self, board_width: int = None, board_height: int = None, device: str = None, config: Config = None):
        super().__init__()
        if config is not None:
            self.width, self.height = config.board_width, config.board_height
            self.device = PolicyValueNetwork.get_torch_device(config.device)
        else:
            self.width, self.height = board_width, board_height
            self.device = PolicyValueNetwork.get_torch_device(device)
        if self.width < 4 or self.height < 4:
            raise ValueError(f"Borad size must be at least 4x4, got {self.width}x{self.height}")
        self.action_size = self.width
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
        self.fc1_input_size = ch * (self.width - 4) * (self.height - 4)
        self.fc1 = nn.Linear(self.fc1_input_size, 64, device=dev)
        self.fc_bn1 = nn.BatchNorm1d(64, device=dev)
        self.fc2 = nn.Linear(64, 32, device=dev)
        self.fc_bn2 = nn.BatchNorm1d(32, device=dev)
        self.fc_probs = nn.Linear(32, self.action_size, device=dev)
        self.fc_value = nn.Linear(32, 1, device=dev)

    def forward(self, input: torch.Tensor):
        x = input.reshape(-1, 1, self.width, self.height)
        x = F.relu(self.bn1(self.conv1(x)))
        x = F.relu(self.bn2(self.conv2(x)))
        x = F.relu(self.bn3(self.conv3(x)))
        x = F.relu(self.bn4(self.conv4(x)))
        x = x.reshape(-1, self.fc1_input_size)
        x = F.dropout(F.relu(self.fc_bn1(self.fc1(x))), p=self.dropout, training=self.training)
        x = F.dropout(F.relu(self.fc_bn2(self.fc2(x))), p=self.dropout, training=self.training)
        return F.log_softmax(self.fc_probs(x), dim=1), torch.tanh(self.fc_value(x))

    def get_normalized_probs(self, probs: np.ndarray, legal_moves: list) -> dict:
        """Priors restricted to the playable columns and renormalised (uniform if their mass is < 1e-6),
        alphazero/games/connect4.py:414-428."""
        mask = np.zeros(self.action_size, dtype=bool)
        mask[legal_moves] = True
        total = np.sum(probs[mask])
        if total < 1e-6:
            return {move: 1 / len(legal_moves) for move in legal_moves}
        return {move: probs[move] / total for move in legal_moves}

    def to_neural_output(self, move_probs: dict) -> np.ndarray:
        pi = np.zeros(self.action_size)
        for move, prob in move_probs.items():
            pi[move] = prob
        return pi

    def reflect_neural_output(self, neural_output: np.ndarray, axis: int) -> np.ndarray:
        """Left-right mirror whatever `axis` says: the board has no vertical symmetry (connect4.py:437-445)."""
        if not neural_output.size == self.action_size:
            raise ValueError(f"Neural output should have size {self.action_size}, but has size {neural_output.size}")
        return np.flip(neural_output)
