"""Small run of every round-2 kernel for compute-sanitizer (memcheck / racecheck): the fc stack on ragged tiles, a
throughput-mode search (virtual loss, red.global backup), self-play through az_tree_move with a tier switch, the greedy
move rule.  Sizes are tiny on purpose: the tools slow kernels down by two to three orders of magnitude.

    compute-sanitizer --tool memcheck  python tools/sanitize_target.py
    compute-sanitizer --tool racecheck python tools/sanitize_target.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200.engine import TreeBatch
from alphazero_b200.evaluators import HashEvaluator
from alphazero_b200.fused_net import FusedNetEvaluator
from alphazero_b200.games.othello import OthelloNet
from alphazero_b200.gamespec import OthelloSpec
from alphazero_b200.selfplay import SelfPlayEngine

torch.manual_seed(0)
net = OthelloNet(n=6)
# fc stack + trunk on 130 positions (two row tiles, the second almost empty)
tree = TreeBatch(130, n=6, mode="neural", capacity=64)
ev = FusedNetEvaluator(net, tree, use_cuda_graph=False)
ev(tree)
torch.cuda.synchronize()
print("fc stack ok", float(tree.probs.sum()))
# the 8 x 8 trunk kernels of large batches: k_net_trunk_rows (1030 positions: a partial last group) and k_net_trunk_rows2
# (4200 positions: two CTAs per SM, a few CTAs with two groups), each followed by the fc stack
net8 = OthelloNet(n=8)
for G8 in (1030, 4200):
    t8 = TreeBatch(G8, n=8, mode="neural", capacity=64)
    t8.leaf_own.copy_(t8.root_own); t8.leaf_opp.copy_(t8.root_opp)
    ev8 = FusedNetEvaluator(net8, t8, use_cuda_graph=False)
    ev8(t8)
    torch.cuda.synchronize()
    print("8 x 8 trunk ok", G8, float(t8.probs.sum()))
# throughput mode: 4 leaves in flight, hash evaluator, tree reuse over two moves
t4 = TreeBatch(16, n=6, mode="neural", tie_random=False, capacity=2048, leaves=4, graph_search=False)
h = HashEvaluator(3, t4.leaf_view)
for _ in range(2):
    t4.search(24, h)
    a, _ = t4.choose(None)
    t4.advance(a)
t4.check_status()
print("throughput mode ok", int(t4.sims_done.sum()))
# self-play through az_tree_move (fused network, 8 games, n_sim 8)
eng = SelfPlayEngine(net, 8, n=6, n_sim=8, seed=1)
b = eng.play()
print("self-play ok", len(b), eng.total_sims)
# greedy move rule
spec = OthelloSpec(6)
own = torch.full((64,), spec.start_position()[0], dtype=torch.int64, device="cuda")
opp = torch.full((64,), spec.start_position()[1], dtype=torch.int64, device="cuda")
print("greedy ok", spec.greedy_moves(own, opp, seed=1, ctr=1)[:4].tolist())
