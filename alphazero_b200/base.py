"""Contracts of the drop-in surface: the same names, signatures and error behaviour as alphazero/base.py
(Board :95-225, Player :228-264, PolicyValueNetwork :267-397, TemperatureScheduler :400-421, Config :60-92,
enums :24-57), written for this package.  Everything above these classes in the reference (players, arena,
trainer) only talks to this surface, so code written against the reference runs against these classes.
"""
import copy
import dataclasses
import enum
import json
import os

import numpy as np
import torch
from torch import nn

DEFAULT_MODELS_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "models")
DEFAULT_OUTPUTS_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "outputs")


class dotdict(dict):
    """dict with attribute access (reference: alphazero/utils.py:22-25)."""

    def __getattr__(self, name):
        return self[name]


class Action:
    """Type tag for moves: (row, col) tuples for Othello, the pass move is (n, n)."""


class _ValueEnum(enum.Enum):
    @classmethod
    def to_dict(cls):
        return {m.value: m for m in cls}


class MoveFormat(_ValueEnum):
    ROW_COL = "row_col"
    ROW = "row"
    COL = "col"


class TreeEval(_ValueEnum):
    ROLLOUT = "rollout"
    NEURAL = "neural"


class DisplayMode(_ValueEnum):
    HUMAN = "human"
    PIXEL = "pixel"


class DataTransf(_ValueEnum):
    REFLECT_H = "reflection_horizontal"
    REFLECT_V = "reflection_vertical"
    ROTATE_90 = "rotation_90"
    ROTATE_180 = "rotation_180"
    ROTATE_270 = "rotation_270"


@dataclasses.dataclass
class Config:
    """Training / playing configuration; field names and defaults follow alphazero/base.py:60-92 so the
    reference's config.json files load unchanged."""
    game: str = None
    simulations: int = None
    compute_time: float = None
    dirichlet_alpha: float = 0.03
    dirichlet_epsilon: float = 0.25
    temp_scheduler_type: str = "linear"
    temp_max_step: int = 15
    temp_min_step: int = 20
    iterations: int = None
    episodes: int = None
    epochs: int = None
    batch_size: int = None
    learning_rate: float = None
    data_augmentation: bool = False
    device: str = None
    eval_opponent: str = "mcts"
    eval_episodes: int = 10
    do_eval: bool = False
    save: bool = True
    push: bool = False
    save_checkpoints: bool = True
    push_checkpoints: bool = False

    def to_dict(self) -> dotdict:
        return dotdict(copy.deepcopy(dataclasses.asdict(self)))


class Board:
    """Game-state contract (alphazero/base.py:95-225).  Attributes callers read: game, grid, player, pass_move,
    max_moves, display_dir, display_mode."""

    CONFIG = Config

    def __init__(self, display_dir: str = None, display_mode=None):
        self.display_dir = DEFAULT_OUTPUTS_PATH if display_dir is None else display_dir
        if display_mode is None:
            self.display_mode = DisplayMode.HUMAN
        else:
            try:
                self.display_mode = DisplayMode(display_mode)
            except ValueError:
                raise ValueError(f"Unknown display mode: {display_mode}")
        self.game = None
        self.player = None
        self.pass_move = None
        self.max_moves = None

    def __str__(self) -> str:
        return type(self).__name__

    def _abstract(self, *a, **k):
        raise NotImplementedError

    reset = clone = get_board_shape = get_n_cells = get_action_size = get_score = _abstract
    is_legal_move = get_moves = get_random_move = play_move = is_game_over = get_winner = _abstract
    human_display = pixel_display = _abstract

    def display(self, show_indexes: bool = True, infos: dict = None, filename: str = None, mode: str = None) -> None:
        if mode is None:
            mode = self.display_mode
        else:
            try:
                mode = DisplayMode(mode)
            except ValueError:
                raise ValueError(f"Unknown display mode: {mode}")
        if mode == DisplayMode.HUMAN:
            self.human_display(show_indexes, infos, filename)
        else:
            self.pixel_display(show_indexes, filename)


class Player:
    """Agent contract (alphazero/base.py:228-264)."""

    def __init__(self, verbose: bool = False) -> None:
        self.verbose = verbose

    def __str__(self) -> str:
        return type(self).__name__

    def clone(self) -> "Player":
        raise NotImplementedError

    def reset(self) -> None:
        pass

    def apply_move(self, move, player: int = None) -> None:
        pass

    def get_move(self, board: Board, temp: float = None):
        """-> (move, action_probs | None, visit_counts | None, prior_probs | None)"""
        raise NotImplementedError

    def get_stats_after_move(self) -> dict:
        return {}


class PolicyValueNetwork(nn.Module):
    """Policy-value network contract (alphazero/base.py:267-397).  forward(input) -> (log_probs [B, A], v [B, 1]);
    the state_dict format is the reference's, so checkpoints are interchangeable."""

    CONFIG = Config

    def __str__(self) -> str:
        return type(self).__name__

    def clone(self) -> "PolicyValueNetwork":
        return copy.deepcopy(self)

    def get_parameters_count(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def save_model(self, model_name: str, model_path: str = None, verbose: bool = False) -> None:
        model_name = model_name.split(".")[0]
        model_path = os.path.join(DEFAULT_MODELS_PATH, model_name) if model_path is None else model_path
        os.makedirs(model_path, exist_ok=True)
        torch.save(self.state_dict(), os.path.join(model_path, f"{model_name}.pt"))
        if verbose:
            print(f"{self} saved in: {model_path}")

    @classmethod
    def from_pretrained(cls, model_name: str, models_path: str = None, verbose: bool = False):
        models_path = DEFAULT_MODELS_PATH if models_path is None else models_path
        model_name = model_name.split(".")[0]
        model_dir = os.path.join(models_path, model_name)
        config_path = os.path.join(model_dir, "config.json")
        pt_path = os.path.join(model_dir, f"{model_name}.pt")
        if not os.path.isfile(config_path):
            raise ValueError(f"Config file not found: {config_path}")
        if not os.path.isfile(pt_path):
            raise ValueError(f"Model file not found: {pt_path}")
        with open(config_path) as f:
            model = cls(config=cls.CONFIG(**json.load(f)))
        model.load_state_dict(torch.load(pt_path, map_location=model.device))
        if verbose:
            print(f"{model} loaded from: {model_dir}")
        return model

    @staticmethod
    def get_torch_device(device: str) -> torch.device:
        if device is None:
            device = "cpu"
        elif device == "mps" and not torch.backends.mps.is_available():
            raise ValueError("MPS not available...")
        elif device == "cuda" and not torch.cuda.is_available():
            raise ValueError("CUDA not available...")
        return torch.device(device)

    def predict(self, input: torch.Tensor):
        """(probs, v) in eval mode without grad (base.py:350-355)."""
        self.eval()
        with torch.no_grad():
            log_probs, v = self.forward(input)
        return torch.exp(log_probs), v

    def evaluate(self, board: Board):
        """Single-position evaluation from the side to move's view; v is returned as an ABSOLUTE value
        (> 0: player +1 ahead), base.py:357-367.  The batched engine does not use this per-leaf call."""
        x = torch.tensor(board.player * board.grid, dtype=torch.float, device=self.device)
        probs, v = self.predict(x)
        return probs.cpu().numpy().reshape(-1), board.player * v.cpu().item()

    def get_normalized_probs(self, probs: np.ndarray, legal_moves: list) -> dict:
        raise NotImplementedError

    def to_neural_output(self, move_probs: dict) -> np.ndarray:
        raise NotImplementedError

    def reflect_neural_output(self, neural_output: np.ndarray, axis: int) -> np.ndarray:
        raise NotImplementedError

    def rotate_neural_output(self, neural_output: np.ndarray, angle: int) -> np.ndarray:
        raise NotImplementedError


class TemperatureScheduler:
    """step -> temperature (alphazero/base.py:400-421)."""

    def __init__(self, temp_max_step: int, temp_min_step: int, max_steps: int) -> None:
        self.temp_max_step = temp_max_step
        self.temp_min_step = temp_min_step
        self.max_steps = max_steps

    def compute_temperature(self, step: int) -> float:
        raise NotImplementedError

    def __getitem__(self, step: int) -> float:
        return self.compute_temperature(step)

    def __iter__(self):
        for step in range(self.max_steps + 1):
            yield self.compute_temperature(step)
