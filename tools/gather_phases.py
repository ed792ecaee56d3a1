"""Times the host-visible phases around the sample all-gather on ONE GPU, at the sizes 8 ranks produce: packing a
rank's batch into wire records, unpacking 8 ranks' records, the device -> host read of the gathered batch.

    python tools/gather_phases.py [games] [n_sim]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200.distributed import _pack_records, _unpack_records
from alphazero_b200.games.othello import OthelloNet
from alphazero_b200.selfplay import SelfPlayEngine

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n_sim = int(sys.argv[2]) if len(sys.argv) > 2 else 16
torch.manual_seed(0)
eng = SelfPlayEngine(OthelloNet(n=8), G, n=8, n_sim=n_sim, dirichlet_alpha=0.03, dirichlet_epsilon=0.25,
                     temp_scheduler_type="linear", temp_max_step=4, temp_min_step=4, seed=1, dtype=torch.bfloat16, device="cuda")
batch = eng.play()
print(f"{len(batch)} samples from {G} games")


def timed(name, fn, reps=3):
    out = None
    for i in range(reps):
        torch.cuda.synchronize()
        t0 = time.time()
        out = fn()
        torch.cuda.synchronize()
        print(f"  {name:34s} run {i}: {(time.time() - t0) * 1e3:8.2f} ms")
    return out


rec, dense = timed("pack one rank's batch", lambda: _pack_records(batch, 0))
print(f"records {rec.numel() / 1e6:.1f} MB, dense policies {dense.numel() * 4 / 1e6:.1f} MB ({dense.shape[0]} samples)")
for world in (1, 2, 8):
    recs, denses = torch.cat([rec] * world), torch.cat([dense] * world)
    big = timed(f"unpack {world} ranks' records", lambda: _unpack_records(recs, denses, batch.pi.shape[1], batch))
    timed(f"device -> host of {world} ranks' batch", lambda: big.cpu())
