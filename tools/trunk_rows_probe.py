"""Cycle stamps of k_net_trunk_rows (debug_layer 200) and its time in a CUDA graph.
   python tools/trunk_rows_probe.py [G] [mode]     mode 201 / 202 / 203: the timing-only modes of the kernel (stamps only)"""
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
ev = FusedOthelloEvaluator(net, tree, use_cuda_graph=False)
n_ctas = min(148, (G + 13) // 14)
W = 64
dbg = torch.zeros(n_ctas * W, dtype=torch.int64, device="cuda")
MODE = int(sys.argv[2]) if len(sys.argv) > 2 else 200
for _ in range(3):
    ev.trunk(debug=dbg.view(torch.bfloat16), debug_layer=MODE)
torch.cuda.synchronize()
d = dbg.reshape(n_ctas, W).double()
print(f"stamps, mode {MODE}")
names = {0: "setup done (after pdl wait)", 1: "im2col of group 0 built", 8: "group 0: conv1 issued",
         9: "group 0: conv2 issued", 10: "group 0: conv3 issued", 3: "group 0 issued (conv4)", 11: "exit",
         12: "mma: total wait im2col", 5: "mma: total wait ready", 6: "mma: total wait for conv4's items + fences", 7: "mma: total issue + commit",
         13: "epilogue team 0: total wait full", 14: "epilogue team 0: total work"}
for i, nm in names.items():
    print(f"    {nm:40s} median {d[:, i].median().item():8.0f}  min {d[:, i].min().item():8.0f}  max {d[:, i].max().item():8.0f} cycles")
if G > 14 * n_ctas:      # the per-job stamps are of the second group of a CTA
    med = lambda i: int(d[:, i].median().item())
    print("    second group, conv3 jobs of warp 0 (tiles 4, 3, 2, 1, 0): issue starts", [med(32 + j) for j in range(5)], "ends", [med(37 + j) for j in range(5)])
    print("    second group, conv2 items (tiles 3, 4, 2, 5, 1, 6, 0, 7): accumulator complete seen", [med(42 + j) for j in range(8)])
    print("                                                              tile ready               ", [med(50 + j) for j in range(8)])
    print("    warp 0, second group: wait for the input tile of job 0..13:", [int(d[:, 16 + j].median().item()) for j in range(14)])
if MODE != 200:
    sys.exit(0)
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
print(f"k_net_trunk_rows, {G} positions: {1000 * e0.elapsed_time(e1) / 200:.2f} us per launch in a graph")
