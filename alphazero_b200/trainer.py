"""`Sample` and `AlphaZeroTrainer` (alphazero/trainer.py:30-118, 121-572).

self_play() produces exactly what the reference's loop produces -- normalised Samples (state*player, pi, z*player)
of `config.episodes` games plus the reflection/rotation twins of every sample with move_idx >= 2 -- but the
episodes are played concurrently by SelfPlayEngine on the GPU and the twins are made by a CUDA kernel
(SampleBatch.augment, csrc/sample_kernels.cu); the memory stays on the device as one SampleBatch
(`memory_batch`), the reference's `memory` list of Sample objects is materialised only when somebody reads it.

optimize_network() is the reference's optimisation step (twin network, SGD momentum .9 / weight decay 1e-4,
ExponentialLR .9, loss = -sum(pi * log p)/B + sum((v - z)^2)/B) with the mini-batches assembled on the device
(SampleBatch.batch); the network forward/backward is the nn.Module's own PyTorch code.  Under torch.distributed
every rank plays its shard of the episodes, the samples are all-gathered, rank 0 optimises and the new weights are
broadcast (distributed.broadcast_network) -- SURVEY.md 8e / 8f N2.  Hub upload and plotting are out of scope."""
import json
import os
from pathlib import Path
from typing import Callable

import numpy as np
import torch

from . import distributed as azdist
from .arena import Arena
from .base import DEFAULT_MODELS_PATH, Config, DataTransf, dotdict
from .games.registers import BOARDS_REGISTER, CONFIGS_REGISTER, DATA_AUGMENT_STRATEGIES, NETWORKS_REGISTER
from .players import PLAYERS_REGISTER, AlphaZeroPlayer
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
    """SampleBatch (device/host tensors) -> the reference's list[Sample], in the batch's order (self-play: by episode
    then move; augmented: originals, then the twins of each eligible sample -- trainer.py:272-284)."""
    b = batch.cpu()
    states = b.states()
    pi = b.pi.numpy().astype(np.float64)
    z, ep, mv = b.z.numpy(), b.episode_idx.numpy(), b.move_idx.numpy()
    tr = None if b.transformation is None else b.transformation.numpy()
    tags = SampleBatch.TRANSFORMATIONS
    return [Sample(state=states[i], pi=pi[i], player=1, outcome=int(z[i]), episode_idx=int(ep[i]), move_idx=int(mv[i]),
                   transformation=None if tr is None else tags[tr[i]])
            for i in range(len(b))]


def batch_from_samples(samples: list, game: str, device="cuda") -> SampleBatch:
    """The reference's list[Sample] (normalised: player == 1, trainer.py:74-78) -> a SampleBatch on `device`: what
    `trainer.memory = [...]` turns the caller's samples into, so that optimize_network trains on THEM."""
    if len(samples) == 0:
        raise ValueError("an empty memory cannot be turned into a SampleBatch")
    rows, cols = samples[0].state.shape
    own, opp = np.zeros(len(samples), np.uint64), np.zeros(len(samples), np.uint64)
    for i, sm in enumerate(samples):
        st = np.asarray(sm.state) * (sm.player if sm.player is not None else 1)
        for r in range(rows):
            for c in range(cols):
                if st[r, c] > 0:
                    own[i] |= np.uint64(1) << np.uint64(r * 8 + c)
                elif st[r, c] < 0:
                    opp[i] |= np.uint64(1) << np.uint64(r * 8 + c)
    tags = list(SampleBatch.TRANSFORMATIONS)
    tr = None
    if any(sm.transformation is not None for sm in samples):
        tr = torch.tensor([tags.index(sm.transformation) if sm.transformation in tags else 0 for sm in samples], dtype=torch.uint8)
    t = lambda a, dt: torch.from_numpy(np.asarray(a)).to(dt).to(device)
    z = [(sm.outcome or 0) * (sm.player if sm.player is not None else 1) for sm in samples]
    return SampleBatch(t(own.view(np.int64), torch.int64), t(opp.view(np.int64), torch.int64),
                       t(np.stack([np.asarray(sm.pi, np.float32) for sm in samples]), torch.float32), t(z, torch.int8),
                       t([sm.episode_idx or 0 for sm in samples], torch.int32), t([sm.move_idx or 0 for sm in samples], torch.int32),
                       cols, rows, None if tr is None else tr.to(device), game)


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
        self.memory_batch = None      # the memory as one device SampleBatch (this engine's native form)
        self._memory = None           # list[Sample], built on demand
        self.nn_twin = None
        self.loss_values = dict()
        self.eval_results = None
        self.verbose = verbose
        self._engine = None

    def __str__(self) -> str:
        return f"{type(self).__name__}{self.game.capitalize()}" if self.game is not None else f"{type(self).__name__}"

    @staticmethod
    def print_config(config: Config, verbose: bool = True) -> None:
        if verbose:
            for pname, value in config.to_dict().items():
                print(f"- {pname}: {value}")

    @property
    def memory(self) -> list:
        """The reference's `memory: list[Sample]` (trainer.py:219), materialised from the device batch when read."""
        if self._memory is None and self.memory_batch is not None:
            self._memory = samples_from_batch(self.memory_batch)
        return self._memory

    @memory.setter
    def memory(self, value) -> None:
        """Assigning a list of Samples (as callers of the reference's trainer.memory can) replaces the training data:
        the device batch is rebuilt from the list, never left pointing at stale self-play samples."""
        self._memory = value
        if value is None or len(value) == 0:
            self.memory_batch = None
        else:
            dev = "cuda" if torch.cuda.is_available() else "cpu"
            self.memory_batch = batch_from_samples(list(value), self.game, device=dev)

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
        self.data_augment_strategy = DATA_AUGMENT_STRATEGIES[self.game] if self.config.data_augmentation else None
        self._init_evaluator()
        self.loss_values = dict()
        self._engine = None

    def _init_evaluator(self) -> None:
        """trainer.py:389-407: validate the evaluation opponent."""
        if not self.config.do_eval:
            return
        if self.config.eval_opponent == "human":
            raise ValueError("Evaluation against a HumanPlayer during training is not allowed.")
        if self.config.eval_opponent not in PLAYERS_REGISTER:
            raise ValueError(f"Opponent player '{self.config.eval_opponent}' not found in the players register.")
        if self.config.eval_opponent == "alphazero":
            raise ValueError("Evaluation against another AlphaZeroPlayer during training is not yet implemented.")
        self.eval_results = {"eval_opponent": self.config.eval_opponent, "eval_episodes": self.config.eval_episodes,
                             "results": dict()}

    # ---- self-play --------------------------------------------------------------------------------------
    def self_play(self, iter_idx: int = 0, seed: int = None) -> None:
        """trainer.py:215-286 with the episodes played concurrently on the device and the symmetry twins made by
        SampleBatch.augment.  Under torch.distributed each rank plays its shard of `config.episodes` and every rank
        ends up with the all-gathered memory (episode indices global, rank order)."""
        cfg = self.config
        if cfg.simulations is None:
            raise ValueError("batched self-play needs config.simulations (compute_time is wall-clock bound per game)")
        rank, world = azdist.rank_world()
        lo, hi = azdist.shard_range(cfg.episodes, rank, world)
        G = hi - lo
        if G < 1:
            raise ValueError(f"config.episodes={cfg.episodes} is smaller than the number of ranks ({world})")
        if self._engine is None or self._engine.G != G or self._engine.n_sim != cfg.simulations:
            self._engine = SelfPlayEngine(self.nn, G, n=spec_of(self.board), n_sim=cfg.simulations,
                                          dirichlet_alpha=cfg.dirichlet_alpha, dirichlet_epsilon=cfg.dirichlet_epsilon,
                                          temp_scheduler_type=cfg.temp_scheduler_type, temp_max_step=cfg.temp_max_step,
                                          temp_min_step=cfg.temp_min_step,
                                          seed=(int(np.random.randint(1 << 31)) if seed is None else seed) + 7919 * rank)
        batch = self._engine.play()
        if world > 1:
            batch = azdist.gather_samples(batch, episode_offset=lo)
        if cfg.data_augmentation:
            batch = batch.augment(min_move_idx=2)
        self.memory_batch, self._memory = batch, None
        self.print(f"Total number of samples: {len(batch)}")

    # ---- optimisation -----------------------------------------------------------------------------------
    def _batch_generator(self, n_batches: int):
        """trainer.py:288-322: a shuffled pass over the memory, `batch_size` samples at a time, assembled on the
        device (az_samples_batch).  The permutation comes from torch's device generator instead of NumPy's."""
        mem, B = self.memory_batch, self.config.batch_size
        perm = self._permutation(len(mem)).to(mem.own.device)[: n_batches * B].reshape(n_batches, B)
        for idx in perm:
            yield mem.batch(idx)

    def _permutation(self, count: int) -> torch.Tensor:
        """The epoch's sample order (np.random.shuffle in trainer.py:292-293)."""
        return torch.randperm(count, device=self.memory_batch.own.device)

    def optimize_network(self, iter_idx: int) -> None:
        """trainer.py:324-381."""
        cfg = self.config
        if self.memory_batch is None:
            raise ValueError("optimize_network needs the memory of a self_play() call")
        dev = self.memory_batch.own.device
        self.nn_twin = self.nn.clone()
        self.nn_twin.to(dev)
        self.nn_twin.device = dev
        self.nn_twin.train()
        optimizer = torch.optim.SGD(self.nn_twin.parameters(), lr=cfg.learning_rate, momentum=0.9, weight_decay=0.0001)
        lr_scheduler = torch.optim.lr_scheduler.ExponentialLR(optimizer, gamma=0.9)
        self.loss_values[iter_idx] = dict()
        self.print(f"\nNetwork optimization ({cfg.epochs} epochs)")
        for epoch_idx in range(cfg.epochs):
            n_batches = len(self.memory_batch) // cfg.batch_size
            if n_batches == 0:
                raise ValueError(f"Too few samples in the memory ({len(self.memory_batch)}) to create a batch with "
                                 f"batch_size = {cfg.batch_size}")
            pi_losses, v_losses = [], []
            for x, pi, z in self._batch_generator(n_batches):
                optimizer.zero_grad()
                log_probs, v = self.nn_twin(x)
                loss_pi = -torch.sum(pi * log_probs) / cfg.batch_size
                loss_v = torch.sum((v - z) ** 2) / cfg.batch_size
                (loss_pi + loss_v).backward()
                optimizer.step()
                pi_losses.append(loss_pi.detach())
                v_losses.append(loss_v.detach())
            # one device->host read per epoch instead of two per batch
            self.loss_values[iter_idx][epoch_idx] = {"pi": torch.stack(pi_losses).cpu().tolist(),
                                                     "v": torch.stack(v_losses).cpu().tolist()}
            lr_scheduler.step()
        self.memory_batch.check_batches()

    def update_network(self, iter_idx=None) -> None:
        """trainer.py:383-387: the optimised twin becomes the network of self-play.  Called with a network instead of
        an iteration index it installs that network.  Under torch.distributed rank 0's weights are broadcast."""
        if isinstance(iter_idx, torch.nn.Module):
            self.nn = iter_idx
        else:
            if self.nn_twin is None:
                raise ValueError("update_network needs the twin network of an optimize_network() call")
            self.nn = self.nn_twin.clone()
            self.nn_twin = None
        self.nn.eval()
        if azdist.rank_world()[1] > 1:
            azdist.broadcast_network(self.nn, src=0)
        self.az_player.mct.nn = self.nn
        if self._engine is not None:
            self._engine.set_network(self.nn)

    # ---- evaluation -------------------------------------------------------------------------------------
    def evaluate(self, iter_idx: int) -> None:
        """trainer.py:409-446; the games of the match are one device batch when both players have a device move rule
        (Arena.play_games_in_parallel), else they go through the Player API one by one."""
        cfg = self.config
        if not cfg.do_eval:
            return
        kwargs = {"n_sim": cfg.simulations} if cfg.eval_opponent == "mcts" else {}
        eval_player = AlphaZeroPlayer(n_sim=cfg.simulations, compute_time=cfg.compute_time, nn=self.nn)
        opponent_player = PLAYERS_REGISTER[cfg.eval_opponent](**kwargs)
        arena = Arena(player1=eval_player, player2=opponent_player, board=BOARDS_REGISTER[self.game](config=cfg))
        self.print(f"\nAlphaZeroPlayer evaluation against {opponent_player} ({cfg.eval_episodes} episodes)")
        batched = cfg.eval_opponent in ("mcts", "random", "greedy") and cfg.simulations is not None and \
            cfg.eval_episodes % 2 == 0
        if batched:
            stats = arena.play_games_in_parallel(n_rounds=cfg.eval_episodes, n_process=2, return_stats=True)
        else:
            stats = arena.play_games(n_rounds=cfg.eval_episodes, return_stats=True)
        if self.verbose:
            Arena.print_stats_results(eval_player, opponent_player, stats)
        for key in ("player1", "player2", "draw"):
            stats.pop(key, None)
        self.eval_results["results"][iter_idx] = stats
        self.eval_results["player1"] = f"{eval_player}"
        self.eval_results["player2"] = f"{opponent_player}"

    # ---- persistence ------------------------------------------------------------------------------------
    def save_player_pt(self, model_name: str, path: str = None) -> None:
        path = os.path.join(DEFAULT_MODELS_PATH, model_name) if path is None else path
        Path(path).mkdir(parents=True, exist_ok=True)
        self.nn.save_model(model_name, path, verbose=False)

    def save_player_config(self, model_name: str, path: str = None) -> None:
        path = os.path.join(DEFAULT_MODELS_PATH, model_name) if path is None else path
        Path(path).mkdir(parents=True, exist_ok=True)
        with open(os.path.join(path, "config.json"), "w") as f:
            json.dump(self.config.to_dict(), f, indent=4)

    def save_training_stats(self, model_name: str, path: str = None) -> None:
        path = os.path.join(DEFAULT_MODELS_PATH, model_name) if path is None else path
        Path(path).mkdir(parents=True, exist_ok=True)
        with open(os.path.join(path, "loss.json"), "w") as f:
            json.dump(self.loss_values, f, indent=4)
        if self.config.do_eval:
            with open(os.path.join(path, "eval.json"), "w") as f:
                json.dump(self.eval_results, f, indent=4)

    # ---- the loop ---------------------------------------------------------------------------------------
    def train(self, game: str = None, experiment_name: str = None, json_config_file: str = None, plot: bool = False,
              verbose: bool = None, config: Config = None, models_path: str = None) -> None:
        """trainer.py:475-572.  `config` (a Config object instead of a JSON file) and `models_path` (where the
        experiment directory goes, default DEFAULT_MODELS_PATH) are additions.  Under torch.distributed only rank 0
        writes files and evaluates."""
        if game is None and json_config_file is None and config is None:
            raise ValueError("The name of the game or a JSON configuration file must be provided to launch a training.")
        cfg = config if config is not None else AlphaZeroTrainer.load_config_from_json(game, json_config_file)
        if game is not None and game != cfg.game:
            raise ValueError(f"Game '{game}' and game '{cfg.game}' in the configuration file do not match.")
        if plot:
            raise NotImplementedError("plotting (matplotlib) is out of scope of the B200 engine")
        if cfg.push or cfg.push_checkpoints:
            raise NotImplementedError("pushing to the Hugging Face Hub is out of scope of the B200 engine")
        experiment_name = experiment_name if experiment_name is not None else AlphaZeroTrainer.DEFAULT_EXP_NAME
        self.verbose = verbose if verbose is not None else self.verbose
        cfg.save = cfg.save or cfg.save_checkpoints
        root = os.path.join(DEFAULT_MODELS_PATH if models_path is None else models_path, experiment_name)
        rank, world = azdist.rank_world()
        self.print("[1] Initializing the Board, PolicyValueNetwork and Player...")
        self.setup(config=cfg)
        if world > 1:
            azdist.broadcast_network(self.nn, src=0)      # every rank starts from rank 0's random initialisation
        if rank == 0:
            self.save_player_config(experiment_name, path=root)
        self.print("\n[2] Training...")
        for iter_idx in range(cfg.iterations):
            self.print(f"\n----- Iteration {iter_idx+1}/{cfg.iterations} -----")
            self.self_play(iter_idx)
            if rank == 0:
                self.optimize_network(iter_idx)
                self.update_network(iter_idx)
            else:
                self.update_network(self.nn)              # receives rank 0's broadcast
            if rank == 0:
                self.evaluate(iter_idx)
                self.save_training_stats(experiment_name, path=root)
                if cfg.save_checkpoints:
                    self.save_player_pt(model_name=f"{experiment_name}-chkpt-{iter_idx+1}",
                                        path=os.path.join(root, "checkpoints"))
                    self.print(f"\n{self} checkpoint {iter_idx+1}/{cfg.iterations} successfully saved.")
        if cfg.save and rank == 0:
            self.save_player_pt(experiment_name, path=root)
            self.print(f"\n[3] {self} successfully saved!")
