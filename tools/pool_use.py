"""High-water mark of the per-game node pool over whole self-play batches, for several seeds (sizing `capacity`)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200.selfplay import SelfPlayEngine
from alphazero_b200.games.othello import OthelloNet

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n_sim = int(sys.argv[2]) if len(sys.argv) > 2 else 200
torch.manual_seed(0)
net = OthelloNet(n=8)
for rank in range(8):
    eng = SelfPlayEngine(net, G, n=8, n_sim=n_sim, seed=0x5EED + 7919 * rank, capacity=40000)
    eng.play()
    print(f"seed of rank {rank}: max nodes used in one pool half {eng.max_nodes_used}  (default capacity {48 * (n_sim + 1)})", flush=True)
    del eng
    torch.cuda.empty_cache()
