import os, sys, time
sys.path.insert(0, "/root/repo")
import torch
from alphazero_b200.games.othello import OthelloNet
from alphazero_b200.selfplay import SelfPlayEngine
leaves = int(sys.argv[1])
torch.manual_seed(0)
eng = SelfPlayEngine(OthelloNet(n=8), 4096, n=8, n_sim=200, dirichlet_alpha=0.03, dirichlet_epsilon=0.25,
                     temp_scheduler_type="linear", temp_max_step=4, temp_min_step=4, seed=0x5EED, dtype=torch.bfloat16,
                     device=torch.device("cuda"), leaves=leaves if leaves > 1 else None)
for i in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.play(); torch.cuda.synchronize()
    print(f"leaves {leaves} rows1={os.environ.get('AZ_B200_TRUNK_ROWS1')}: batch {i}: {1000*(time.perf_counter()-t0):.1f} ms, {eng.total_sims/(time.perf_counter()-t0)/1e6:.2f} M sims/s", flush=True)
