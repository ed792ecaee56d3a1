"""`Sample` and the self-play half of `AlphaZeroTrainer` (alphazero/trainer.py:30-118, 215-286).

self_play() produces exactly what the reference's loop produces -- normalised Samples (state*player, pi, z*player)
of `config.episodes` games plus the reflection/rotation twins of every sample with move_idx >= 2 -- but the
episodes are played concurrently by SelfPlayEngine on the GPU.  The optimisation / evaluation / checkpoint half of
the trainer (trainer.py:288-572) is plain PyTorch and outside the hot path; it is not
re-implemented here."""
import json
from typing import Callable

import numpy as np

from .base import Config, DataTransf, dotdict
from .games.registers import BOARDS_REGISTER, CONFIGS_REGISTER, DATA_AUGMENT_STRATEGIES, NETWORKS_REGISTER
from .players import AlphaZeroPlayer
from .schedulers import TEMP_SCHEDULERS
from .gamespec import spec_of
from .selfplay import SampleBatch, SelfPlayEngine


class Sample:
    """One training sample (trainer.py:30-78)."""

    def __init__(self, state: np.ndarray, pi: np.ndarray, player: int, outcome: int = None, episode_idx: int = None,
                 move_idx: int = None, transformation: str = None) -> None:
        self.state = state
        self.pi = pi
        self.player = player
        self.outcome = outcome
        self.episode_idx = episode_idx
        self.move_idx = move_idx
        self.transformation = transformation

    def __str__(self) -> str:
        return (f"Sample(state={self.state}, pi={self.pi}, player={self.player}, outcome={self.outcome}, "
                f"episode_idx={self.episode_idx}, move_idx={self.move_idx}, transformation={self.transformation})")

    def normalize(self) -> None:
        self.state = self.state * self.player
        self.outcome = self.outcome * self.player
        self.player = 1

    def _twin(self, state, pi, mode) -> "Sample":
        tag = mode.value if self.transformation is None else f"{self.transformation}+{mode.value}"
        return Sample(state=state, pi=pi, player=self.player, outcome=self.outcome, episode_idx=self.episode_idx,
                      move_idx=self.move_idx, transformation=tag)

    def create_reflection_twin(self, reflection: Callable, mode: DataTransf) -> "Sample":
        if mode not in (DataTransf.REFLECT_H, DataTransf.REFLECT_V):
            raise ValueError(f"Reflection mode must be either {DataTransf.REFLECT_H} or {DataTransf.REFLECT_V}.")
        axis = 1 if mode == DataTransf.REFLECT_H else 0
        return self._twin(np.flip(self.state.copy(), axis=axis), reflection(self.pi.copy(), axis=axis), mode)

    def create_rotation_twin(self, rotation: Callable, mode: DataTransf) -> "Sample":
        angles = {DataTransf.ROTATE_90: 90, DataTransf.ROTATE_180: 180, DataTransf.ROTATE_270: 270}
        if mode not in angles:
            raise ValueError(f"Rotation mode must be one of {list(angles)}.")
        return self._twin(np.rot90(self.state.copy(), k=angles[mode] // 90), rotation(self.pi.copy(), angle=angles[mode]), mode)


def samples_from_batch(batch: SampleBatch) -> list:
    """SampleBatch (device/host tensors) -> the reference's list[Sample], ordered by episode then move."""
    b = batch.cpu()
    states = b.states()
    pi = b.pi.numpy().astype(np.float64)
    z, ep, mv = b.z.numpy(), b.episode_idx.numpy(), b.move_idx.numpy()
    return [Sample(state=states[i], pi=pi[i], player=1, outcome=int(z[i]), episode_idx=int(ep[i]), move_idx=int(mv[i]))
            for i in range(len(b))]


class AlphaZeroTrainer:
    """Self-play producer with the reference trainer's attributes (trainer.py:121-138)."""

    DEFAULT_EXP_NAME = "alphazero-undefined"

    def __init__(self, verbose: bool = False):
        self.game = None
        self.config = None
        self.board = None
        self.nn = None
        self.az_player = None
        self.temp_scheduler = None
        self.data_augment_strategy = None
        self.memory = None
        self.verbose = verbose
        self._engine = None

    def print(self, log: str):
        if self.verbose:
            print(log)

    @staticmethod
    def load_config_from_json(game: str, json_config_file: str):
        if json_config_file is None:
            return CONFIGS_REGISTER[game]()
        with open(json_config_file) as f:
            cfg = dotdict(json.load(f))
        return CONFIGS_REGISTER[cfg.game](**cfg)

    def setup(self, game: str = "othello", config: Config = None, nn=None) -> None:
        """The object construction part of train() (trainer.py:488-516)."""
        self.config = config if config is not None else CONFIGS_REGISTER[game]()
        self.game = self.config.game
        self.board = BOARDS_REGISTER[self.game](config=self.config)
        self.nn = nn if nn is not None else NETWORKS_REGISTER[self.game](config=self.config)
        self.az_player = AlphaZeroPlayer(n_sim=self.config.simulations, compute_time=self.config.compute_time,
                                         nn=self.nn, dirichlet_alpha=self.config.dirichlet_alpha,
                                         dirichlet_epsilon=self.config.dirichlet_epsilon)
        self.temp_scheduler = TEMP_SCHEDULERS[self.config.temp_scheduler_type](
            self.config.temp_max_step, self.config.temp_min_step, self.board.max_moves)
        self.data_augment_strategy = DATA_AUGMENT_STRATEGIES[self.game]
        self._engine = None

    def update_network(self, nn) -> None:
        """trainer.py:383-387: swap the network used by self-play."""
        self.nn = nn
        self.az_player.mct.nn = nn
        if self._engine is not None:
            self._engine.set_network(nn)

    def self_play(self, iter_idx: int = 0, seed: int = None) -> None:
        """trainer.py:215-286 with all `config.episodes` games played concurrently on the device."""
        cfg = self.config
        if cfg.simulations is None:
            raise ValueError("batched self-play needs config.simulations (compute_time is wall-clock bound per game)")
        if self._engine is None or self._engine.G != cfg.episodes or self._engine.n_sim != cfg.simulations:
            self._engine = SelfPlayEngine(self.nn, cfg.episodes, n=spec_of(self.board), n_sim=cfg.simulations,
                                          dirichlet_alpha=cfg.dirichlet_alpha, dirichlet_epsilon=cfg.dirichlet_epsilon,
                                          temp_scheduler_type=cfg.temp_scheduler_type, temp_max_step=cfg.temp_max_step,
                                          temp_min_step=cfg.temp_min_step,
                                          seed=int(np.random.randint(1 << 31)) if seed is None else seed)
        self.memory = samples_from_batch(self._engine.play())
        if cfg.data_augmentation:
            strat = self.data_augment_strategy
            extra = []
            for s in self.memory:
                if s.move_idx >= 2:
                    refl = s.create_reflection_twin(self.nn.reflect_neural_output, mode=strat.reflection)
                    extra.append(refl)
                    for rot in strat.rotations:
                        extra.append(s.create_rotation_twin(self.nn.rotate_neural_output, mode=rot))
                        extra.append(refl.create_rotation_twin(self.nn.rotate_neural_output, mode=rot))
            self.memory += extra
        self.print(f"Total number of samples: {len(self.memory)}")
