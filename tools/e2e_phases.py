"""Wall / device time of one self-play batch from resident start positions, from host start positions, and of the
device -> host read of its samples.    python tools/e2e_phases.py [othello|connect4] [games] [n_sim]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200 import boardops as bo
from alphazero_b200.selfplay import HostStaging, SelfPlayEngine

game = sys.argv[1] if len(sys.argv) > 1 else "othello"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
n_sim = int(sys.argv[3]) if len(sys.argv) > 3 else 200
torch.manual_seed(0)
if game == "connect4":
    from alphazero_b200.games.connect4 import Connect4Net
    from alphazero_b200.gamespec import Connect4Spec
    spec, net = Connect4Spec(7, 6), Connect4Net(7, 6)
else:
    from alphazero_b200.games.othello import OthelloNet
    from alphazero_b200.gamespec import OthelloSpec
    spec, net = OthelloSpec(8), OthelloNet(n=8)
eng = SelfPlayEngine(net, G, n=spec, n_sim=n_sim, dirichlet_alpha=0.03, dirichlet_epsilon=0.25, temp_scheduler_type="linear",
                     temp_max_step=4, temp_min_step=4, seed=5, dtype=torch.bfloat16, device="cuda")
black, white = spec.start_position()
h_own = torch.full((G,), bo.to_i64(black), dtype=torch.int64).pin_memory()
h_opp = torch.full((G,), bo.to_i64(white), dtype=torch.int64).pin_memory()
h_player = torch.ones(G, dtype=torch.int8).pin_memory()
st = HostStaging()


def run(name, fn, reps=3):
    for i in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"  {name:40s} run {i}: wall {(time.time() - t0) * 1e3:8.1f} ms  device {e0.elapsed_time(e1):8.1f} ms  sims {eng.total_sims}")
    return out


b = run("play()", lambda: eng.play())
b = run("play(host own, opp, player)", lambda: eng.play(h_own.cuda(non_blocking=True), h_opp.cuda(non_blocking=True), h_player.cuda(non_blocking=True)))
run("batch.cpu(staging)", lambda: b.cpu(st))
run("play() again", lambda: eng.play())
