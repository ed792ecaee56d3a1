"""The trunk kernels on the same batch in one process (same GPU, same clocks: the boxes differ by +-5 %): us per
launch in a CUDA graph.      python tools/trunk_ab.py [G]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200.engine import TreeBatch
from alphazero_b200.fused_net import FusedOthelloEvaluator
from alphazero_b200.games.othello import OthelloNet

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
torch.manual_seed(0)
net = OthelloNet(n=8)
tree = TreeBatch(G, n=8, mode="neural", capacity=64)
tree.leaf_own.random_(0, 2**62); tree.leaf_opp.copy_(~tree.leaf_own & torch.randint(0, 2**62, (G,), device="cuda"))


def timed(ev):
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        ev.trunk()
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        for _ in range(20):
            ev.trunk()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return 1000 * e0.elapsed_time(e1) / 200


evs = {v: FusedOthelloEvaluator(net, tree, use_cuda_graph=False, trunk_variant=v) for v in (1, 2, 0)}
for rep in range(2):
    print(f"{G} positions, us per launch: k_net_trunk {timed(evs[1]):.2f} | k_net_trunk_rows {timed(evs[2]):.2f} | az_net_trunk's choice (k_net_trunk_rows2 above 2072 positions) {timed(evs[0]):.2f}")
