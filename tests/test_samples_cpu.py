"""CPU: the training-side oracle (oracle/samples_oracle.py) against the fixtures the unmodified reference produced
(tests/golden/make_golden_train.py): augmentation order / boards / pi / tags, and the optimisation loop."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLDEN = os.path.join(HERE, "golden")

import samples_oracle as so  # noqa: E402


@pytest.mark.parametrize("name,game", [("othello_n8", "othello"), ("othello_n6", "othello"), ("connect4_7x6", "connect4"),
                                       ("tictactoe", "tictactoe")])
def test_oracle_augmentation_matches_reference_fixture(name, game):
    d = np.load(os.path.join(GOLDEN, f"augment_{name}.npz"))
    got = so.augment(game, d["in_state"], d["in_pi"], d["in_z"], d["in_episode_idx"], d["in_move_idx"])
    for k in ("state", "pi", "z", "episode_idx", "move_idx", "transformation"):
        assert np.array_equal(got[k], d[f"out_{k}"]), k
    eligible = int((d["in_move_idx"] >= 2).sum())
    assert len(got["z"]) == len(d["in_z"]) + eligible * (1 if game == "connect4" else 7)


def test_state_bits_round_trip():
    rng = np.random.RandomState(3)
    state = rng.choice([0.0, 1.0, -1.0], size=(50, 6, 7))
    own, opp = so.state_to_bits(state)
    assert not np.any(own & opp)
    back = np.zeros_like(state)
    for r in range(6):
        for c in range(7):
            b = np.uint64(r * 8 + c)
            back[:, r, c] = ((own >> b) & np.uint64(1)).astype(float) - ((opp >> b) & np.uint64(1)).astype(float)
    assert np.array_equal(back, state)


def _connect4_net(d):
    from alphazero_b200.games.connect4 import Connect4Config, Connect4Net
    net = Connect4Net(config=Connect4Config())
    net.load_state_dict({k[5:]: torch.tensor(d[k]) for k in d.files if k.startswith("init_")})
    net.dropout = 0.0
    return net


def test_oracle_optimisation_reproduces_reference_connect4():
    d = np.load(os.path.join(GOLDEN, "train_connect4_7x6.npz"))
    B, E, lr = int(d["hyper"][0]), int(d["hyper"][1]), float(d["hyper"][2])
    twin, lp, lv = so.optimize(_connect4_net(d), d["mem_state"], d["mem_pi"], d["mem_z"], d["perms"], B, E, lr)
    assert np.allclose(lp, d["loss_pi"], rtol=0, atol=1e-6) and np.allclose(lv, d["loss_v"], rtol=0, atol=1e-6)
    for k, v in twin.state_dict().items():
        assert np.allclose(v.numpy(), d[f"final_{k}"], rtol=0, atol=1e-6), k


def test_oracle_optimisation_reproduces_reference_othello6():
    from alphazero_b200.games.othello import OthelloConfig, OthelloNet
    d = np.load(os.path.join(GOLDEN, "train_othello_n6.npz"))
    torch.manual_seed(int(d["seed"][0]))
    net = OthelloNet(config=OthelloConfig(board_size=6))
    for k, v in net.state_dict().items():       # the seeded constructor reproduces the reference's initialisation
        assert np.allclose([v.double().sum().item(), v.double().norm().item()], d[f"initsum_{k}"], rtol=1e-12, atol=1e-12), k
    net.dropout = 0.0
    B, E, lr = int(d["hyper"][0]), int(d["hyper"][1]), float(d["hyper"][2])
    twin, lp, lv = so.optimize(net, d["mem_state"], d["mem_pi"], d["mem_z"], d["perms"], B, E, lr)
    assert np.allclose(lp, d["loss_pi"], rtol=0, atol=1e-5) and np.allclose(lv, d["loss_v"], rtol=0, atol=1e-5)
    for k, v in twin.state_dict().items():
        got = [v.double().sum().item(), v.double().norm().item()]
        assert np.allclose(got, d[f"finalsum_{k}"], rtol=1e-5, atol=1e-4), k
