"""Generates the training-side golden fixtures by IMPORTING THE UNMODIFIED REFERENCE (/root/reference).

Run in the build container only (the reference does not exist on the GPU box):
    python tests/golden/make_golden_train.py
Outputs (committed):
    augment_{game}.npz   random normalised samples (state, pi, z, episode_idx, move_idx) and the memory the data
                         augmentation block of AlphaZeroTrainer.self_play (trainer.py:275-284) builds from them with
                         the reference's own Sample.create_reflection_twin / create_rotation_twin, the networks'
                         reflect/rotate_neural_output and DATA_AUGMENT_STRATEGIES; transformation tags included.
    train_{game}.npz     a seeded reference network (dropout set to 0 so CPU and GPU agree), a fixed memory, the
                         index permutations AlphaZeroTrainer._batch_generator drew, and what the UNMODIFIED
                         AlphaZeroTrainer.optimize_network (trainer.py:324-381) produced from them: loss_pi / loss_v
                         of every batch and the trained twin (Connect4Net: full state_dict; OthelloNet 6x6: per-tensor
                         sums and L2 norms, the state_dict would be 2.7 MB).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from ref_import import import_reference  # noqa: E402

import_reference()
import alphazero.trainer as ref_trainer  # noqa: E402
from alphazero.games.connect4 import Connect4Board, Connect4Config, Connect4Net  # noqa: E402
from alphazero.games.othello import OthelloBoard, OthelloConfig, OthelloNet  # noqa: E402
from alphazero.games.registers import DATA_AUGMENT_STRATEGIES  # noqa: E402
from alphazero.games.tictactoe import TicTacToeNet  # noqa: E402

TAGS = [None, "reflection_horizontal", "rotation_90", "reflection_horizontal+rotation_90", "rotation_180",
        "reflection_horizontal+rotation_180", "rotation_270", "reflection_horizontal+rotation_270"]


def random_samples(rng, count, rows, cols, action_size, n_episodes=5):
    """Normalised samples as self_play leaves them in memory (player = 1): random discs, a random sparse pi."""
    out = []
    for i in range(count):
        state = rng.choice([0.0, 1.0, -1.0], size=(rows, cols), p=[0.4, 0.3, 0.3])
        pi = rng.rand(action_size) * (rng.rand(action_size) < 0.4)
        if pi.sum() == 0:
            pi[rng.randint(action_size)] = 1.0
        pi = pi / pi.sum()
        out.append(ref_trainer.Sample(state=state, pi=pi, player=1, outcome=int(rng.randint(-1, 2)),
                                      episode_idx=int(i % n_episodes), move_idx=int(rng.randint(0, 9))))
    return out


def augment(memory, net, strategy):
    """The statements of trainer.py:275-284, on the reference's objects."""
    augmented_samples = []
    for sample in memory:
        if sample.move_idx >= 2:
            reflected_sample = sample.create_reflection_twin(net.reflect_neural_output, mode=strategy.reflection)
            augmented_samples.append(reflected_sample)
            for rot_mode in strategy.rotations:
                augmented_samples.append(sample.create_rotation_twin(net.rotate_neural_output, mode=rot_mode))
                augmented_samples.append(reflected_sample.create_rotation_twin(net.rotate_neural_output, mode=rot_mode))
    return memory + augmented_samples


def pack(memory):
    return dict(state=np.stack([s.state for s in memory]), pi=np.stack([s.pi for s in memory]),
                z=np.array([s.outcome for s in memory], dtype=np.int8),
                episode_idx=np.array([s.episode_idx for s in memory], dtype=np.int32),
                move_idx=np.array([s.move_idx for s in memory], dtype=np.int32),
                transformation=np.array([TAGS.index(s.transformation) for s in memory], dtype=np.uint8))


def make_augment(name, game, net, rows, cols, seed):
    rng = np.random.RandomState(seed)
    memory = random_samples(rng, 48, rows, cols, net.action_size)
    src = pack(memory)
    full = pack(augment(memory, net, DATA_AUGMENT_STRATEGIES[game]))
    np.savez_compressed(os.path.join(HERE, f"augment_{name}.npz"), **{f"in_{k}": v for k, v in src.items()},
                        **{f"out_{k}": v for k, v in full.items()})
    print(name, len(memory), "->", len(full["z"]))


def make_train(name, game, config, net_cls, board_cls, rows, cols, seed, full_state):
    torch.manual_seed(seed)
    rng = np.random.RandomState(seed)
    net = net_cls(config=config)
    net.dropout = 0.0                      # F.dropout(p=0) is the identity: removes the torch RNG from the comparison
    init = {k: v.clone().numpy() for k, v in net.state_dict().items()}
    memory = random_samples(rng, 80, rows, cols, net.action_size)
    tr = ref_trainer.AlphaZeroTrainer()
    tr.game, tr.config, tr.nn, tr.board = game, config, net, board_cls(config=config)
    tr.memory, tr.loss_values = memory, dict()
    perms = []
    orig_shuffle = np.random.shuffle

    def recording_shuffle(x):
        orig_shuffle(x)
        perms.append(np.array(x).copy())

    np.random.seed(seed)
    np.random.shuffle = recording_shuffle
    try:
        tr.optimize_network(0)             # UNMODIFIED reference code
    finally:
        np.random.shuffle = orig_shuffle
    twin = tr.nn_twin.state_dict()
    if full_state:
        out = {f"init_{k}": v for k, v in init.items()}
    else:      # the same torch.manual_seed(seed) + constructor reproduces the initialisation; its checksums pin that
        out = {f"initsum_{k}": np.array([np.float64(v.astype(np.float64).sum()), np.linalg.norm(v.astype(np.float64).ravel())])
               for k, v in init.items()}
        out["seed"] = np.array([seed])
    if full_state:
        out.update({f"final_{k}": v.numpy() for k, v in twin.items()})
    else:
        out.update({f"finalsum_{k}": np.array([v.double().sum().item(), v.double().norm().item()]) for k, v in twin.items()})
    src = pack(memory)
    out.update({f"mem_{k}": v for k, v in src.items()})
    out["perms"] = np.stack(perms)
    out["loss_pi"] = np.array([tr.loss_values[0][e]["pi"] for e in range(config.epochs)])
    out["loss_v"] = np.array([tr.loss_values[0][e]["v"] for e in range(config.epochs)])
    out["hyper"] = np.array([config.batch_size, config.epochs, config.learning_rate])
    np.savez_compressed(os.path.join(HERE, f"train_{name}.npz"), **out)
    print(name, "loss_pi", out["loss_pi"][:, 0], "->", out["loss_pi"][:, -1], "loss_v", out["loss_v"][:, -1])


def main():
    make_augment("othello_n8", "othello", OthelloNet(n=8), 8, 8, 11)
    make_augment("othello_n6", "othello", OthelloNet(n=6), 6, 6, 12)
    make_augment("connect4_7x6", "connect4", Connect4Net(board_width=7, board_height=6), 6, 7, 13)
    make_augment("tictactoe", "tictactoe", TicTacToeNet(), 3, 3, 14)
    make_train("connect4_7x6", "connect4", Connect4Config(batch_size=16, epochs=3, learning_rate=0.01), Connect4Net,
               Connect4Board, 6, 7, 21, full_state=True)
    make_train("othello_n6", "othello", OthelloConfig(board_size=6, batch_size=16, epochs=2, learning_rate=0.01),
               OthelloNet, OthelloBoard, 6, 6, 22, full_state=False)


if __name__ == "__main__":
    main()
