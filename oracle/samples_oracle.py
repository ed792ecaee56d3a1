"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy / plain fp32 PyTorch) of the training-side steps of the
reference: symmetry augmentation, mini-batch assembly, the optimisation loop.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import this; the product never does.

Pinned by tests/golden/augment_*.npz and train_*.npz, which tests/golden/make_golden_train.py produced by running the
UNMODIFIED reference (tests/test_samples_cpu.py checks this file against them).

Follows (file:line relative to /root/reference):
  alphazero/trainer.py:80-118     Sample.create_reflection_twin / create_rotation_twin
  alphazero/trainer.py:275-284    the augmentation loop of AlphaZeroTrainer.self_play
  alphazero/games/othello.py:414-450, connect4.py:437-445, tictactoe.py:343-368   reflect/rotate_neural_output
  alphazero/games/registers.py:37-50   DATA_AUGMENT_STRATEGIES
  alphazero/trainer.py:288-322    _batch_generator
  alphazero/trainer.py:324-381    optimize_network
"""
import copy

import numpy as np
import torch

REFLECT, ROT90, ROT180, ROT270 = "reflect_h", 90, 180, 270
# transformation codes in the order the reference appends a sample's twins (trainer.py:278-283)
CODES = (None, "reflection_horizontal", "rotation_90", "reflection_horizontal+rotation_90", "rotation_180",
         "reflection_horizontal+rotation_180", "rotation_270", "reflection_horizontal+rotation_270")


def _reflect_pi(game, pi, rows, cols):
    if game == "connect4":                      # connect4.py:445: np.flip of the whole vector
        return pi[::-1].copy()
    if game == "tictactoe":                     # tictactoe.py:353: all 9 entries are cells
        return np.flip(pi.reshape(rows, cols), axis=1).flatten()
    out = np.zeros_like(pi)                     # othello.py:424-428: flip the board part along axis 1, keep the tail
    out[:-1] = np.flip(pi[:-1].reshape(rows, cols), axis=1).flatten()
    out[-1:] = pi[-1:]
    return out


def _rotate_pi(game, pi, n, angle):
    if game == "tictactoe":                     # tictactoe.py:368
        return np.rot90(pi.reshape(n, n), k=angle // 90).flatten()
    out = np.zeros_like(pi)                     # othello.py:444-448
    out[:-1] = np.rot90(pi[:-1].reshape(n, n), k=angle // 90).flatten()
    out[-1:] = pi[-1:]
    return out


def augment(game, state, pi, z, episode_idx, move_idx, min_move_idx=2):
    """state [S, rows, cols], pi [S, A], z/episode_idx/move_idx [S] -> the augmented memory in the reference's order
    (originals, then per eligible sample: reflection, then for each rotation the rotated sample and the rotated
    reflection) as a dict of arrays with a `transformation` code per sample."""
    rows, cols = state.shape[1:]
    rotations = (ROT90, ROT180, ROT270) if game in ("othello", "tictactoe") else ()
    out = [(state[i], pi[i], z[i], episode_idx[i], move_idx[i], 0) for i in range(len(z))]
    for i in range(len(z)):
        if move_idx[i] < min_move_idx:
            continue
        rs, rp = np.flip(state[i], axis=1), _reflect_pi(game, pi[i], rows, cols)
        out.append((rs, rp, z[i], episode_idx[i], move_idx[i], 1))
        for k, angle in enumerate(rotations):
            out.append((np.rot90(state[i], k=angle // 90), _rotate_pi(game, pi[i], cols, angle), z[i], episode_idx[i],
                        move_idx[i], 2 + 2 * k))
            out.append((np.rot90(rs, k=angle // 90), _rotate_pi(game, rp, cols, angle), z[i], episode_idx[i], move_idx[i],
                        3 + 2 * k))
    return dict(state=np.stack([o[0] for o in out]), pi=np.stack([o[1] for o in out]),
                z=np.array([o[2] for o in out], dtype=np.int8),
                episode_idx=np.array([o[3] for o in out], dtype=np.int32),
                move_idx=np.array([o[4] for o in out], dtype=np.int32),
                transformation=np.array([o[5] for o in out], dtype=np.uint8))


def batch(state, pi, z, index):
    """trainer.py:309-320: the three float32 tensors of one mini-batch."""
    return (torch.tensor(state[index], dtype=torch.float32), torch.tensor(pi[index], dtype=torch.float32),
            torch.tensor(z[index].reshape(-1, 1), dtype=torch.float32))


def optimize(net, state, pi, z, perms, batch_size, epochs, learning_rate, device="cpu"):
    """trainer.py:324-381 with the permutation of each epoch given (the reference draws it with np.random.shuffle).
    Returns (trained twin, loss_pi [epochs][n_batches], loss_v [epochs][n_batches])."""
    twin = copy.deepcopy(net).to(device)
    twin.train()
    opt = torch.optim.SGD(twin.parameters(), lr=learning_rate, momentum=0.9, weight_decay=0.0001)
    sched = torch.optim.lr_scheduler.ExponentialLR(opt, gamma=0.9)
    loss_pi, loss_v = [], []
    for e in range(epochs):
        n_batches = len(z) // batch_size
        idx = np.asarray(perms[e])[: n_batches * batch_size].reshape(n_batches, batch_size)
        lp, lv = [], []
        for b in idx:
            x, p, t = (a.to(device) for a in batch(state, pi, z, b))
            opt.zero_grad()
            log_probs, v = twin(x)
            l_pi = -torch.sum(p * log_probs) / batch_size
            l_v = torch.sum((v - t) ** 2) / batch_size
            (l_pi + l_v).backward()
            opt.step()
            lp.append(l_pi.item())
            lv.append(l_v.item())
        loss_pi.append(lp)
        loss_v.append(lv)
        sched.step()
    return twin, np.array(loss_pi), np.array(loss_v)


def state_to_bits(state):
    """[S, rows, cols] grids of 0/+1/-1 -> (own, opp) uint64 bitboards on the 8-wide canvas (bit r*8+c)."""
    S, rows, cols = state.shape
    own = np.zeros(S, dtype=np.uint64)
    opp = np.zeros(S, dtype=np.uint64)
    for r in range(rows):
        for c in range(cols):
            bit = np.uint64(1) << np.uint64(r * 8 + c)
            own |= np.where(state[:, r, c] == 1, bit, np.uint64(0))
            opp |= np.where(state[:, r, c] == -1, bit, np.uint64(0))
    return own, opp
