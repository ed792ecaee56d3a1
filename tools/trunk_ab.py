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

# Connect4Net's 7 x 6 image: az_connect4_image_bits + the trunk (k_net_trunk<7, 6> | k_net_trunk_rows2<7, 6> above 2664 boards)
from alphazero_b200.engine import TreeBatch as _TB
from alphazero_b200.fused_net import FusedNetEvaluator
from alphazero_b200.games.connect4 import Connect4Net
from alphazero_b200.gamespec import Connect4Spec

c4 = _TB(G, n=Connect4Spec(7, 6), mode="neural", capacity=64)
c4.leaf_own.random_(0, 2**40); c4.leaf_opp.copy_(~c4.leaf_own & torch.randint(0, 2**40, (G,), device="cuda"))
cnet = Connect4Net(board_width=7, board_height=6)
cev = {v: FusedNetEvaluator(cnet, c4, use_cuda_graph=False, trunk_variant=v) for v in (1, 0)}
for rep in range(2):
    print(f"{G} Connect4 positions, us per (image bits + trunk): k_net_trunk<7, 6> {timed(cev[1]):.2f} | az_net_trunk's choice {timed(cev[0]):.2f}")
