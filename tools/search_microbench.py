"""Device time of one whole search step (leaf evaluation + tree step) inside TreeBatch.search, per simulation step:
eager loop vs the whole-search CUDA graph.  AZ_B200_NO_PDL=1 switches programmatic dependent launch off."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200.engine import TreeBatch
from alphazero_b200.fused_net import FusedNetEvaluator
from alphazero_b200.games.othello import OthelloNet

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n_sim = int(sys.argv[2]) if len(sys.argv) > 2 else 200
torch.manual_seed(0)
net = OthelloNet(n=8)
for graph in (False, True):
    tree = TreeBatch(G, n=8, mode="neural", n_sim_hint=n_sim, graph_search=graph, dirichlet_alpha=0.03, dirichlet_epsilon=0.25)
    ev = FusedNetEvaluator(net, tree)
    times = []
    for rep in range(6):
        tree.reset_roots()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tree.search(n_sim, ev, check=False)
        e1.record()
        torch.cuda.synchronize()
        times.append(1000 * e0.elapsed_time(e1) / (n_sim + 1))
    tree.check_status()
    print(f"graph_search={graph!s:5s} pdl={'off' if os.environ.get('AZ_B200_NO_PDL') else 'on '}  us per simulation step: "
          + " ".join(f"{t:6.2f}" for t in times))
