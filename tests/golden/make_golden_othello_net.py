"""Golden fixture for the headline network: a seeded REFERENCE OthelloNet(n=8) (alphazero/games/othello.py:327-382),
256 mid-game positions and the reference's own fp32 predict() outputs on them.

    python tests/golden/make_golden_othello_net.py      (build container only)

Output (committed): tests/golden/othello_net_n8.npz -- black / white / player, input (player * grid, float32
[256, 8, 8]), probs [256, 65], value [256, 1], and the full state_dict ("sd.<name>") so that the GPU tests load exactly
these weights into the engine.  BatchNorm statistics are made non-trivial so that folding them is exercised.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ref_import import import_reference  # noqa: E402
from hash_eval import grid_to_bits  # noqa: E402

import_reference()
from alphazero.games.othello import OthelloBoard, OthelloNet  # noqa: E402


def main(n=8, count=256, seed=29):
    torch.manual_seed(seed)
    np.random.seed(seed)
    net = OthelloNet(n=n, device="cpu")
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, (torch.nn.BatchNorm1d, torch.nn.BatchNorm2d)):
                m.running_mean.normal_(0.0, 0.2); m.running_var.uniform_(0.5, 1.5)
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0.0, 0.2)
        net.fc_probs.weight.mul_(3.0)        # a policy that is not nearly uniform
    net.eval()
    rng = np.random.RandomState(seed)
    blacks, whites, players, grids = [], [], [], []
    while len(grids) < count:
        b = OthelloBoard(n=n)
        for _ in range(rng.randint(0, n * n - 6)):
            if b.is_game_over():
                break
            mv = b.get_moves(); b.play_move(mv[rng.randint(len(mv))])
        if b.is_game_over():
            continue
        bl, wh = grid_to_bits(b.grid)
        blacks.append(bl); whites.append(wh); players.append(b.player); grids.append(b.player * b.grid)
    x = torch.tensor(np.stack(grids), dtype=torch.float32)
    probs, v = net.predict(x)
    # evaluate() of single boards (base.py:357-367) agrees with the batched predict
    b = OthelloBoard(n=n, grid=np.stack(grids)[0] * players[0], player=players[0])
    p0, v0 = net.evaluate(b)
    assert np.allclose(p0, probs[0].numpy(), atol=1e-6) and abs(v0 - players[0] * float(v[0])) < 1e-6
    sd = {k: t.detach().cpu().numpy() for k, t in net.state_dict().items()}
    np.savez_compressed(os.path.join(HERE, f"othello_net_n{n}.npz"),
                        black=np.array(blacks, np.uint64), white=np.array(whites, np.uint64),
                        player=np.array(players, np.int8), input=x.numpy(), probs=probs.numpy(), value=v.numpy(),
                        **{"sd." + k: a for k, a in sd.items()})
    print(f"othello net n={n}: {count} positions, {sum(a.size for a in sd.values())} state_dict values, "
          f"max prob {probs.max().item():.3f}")


if __name__ == "__main__":
    main()
