"""Times the pieces of the fused network forward in isolation (CUDA events, G leaf positions)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200 import _lib
from alphazero_b200.engine import TreeBatch
from alphazero_b200.fused_net import FusedOthelloEvaluator
from alphazero_b200.games.othello import OthelloNet

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 200
torch.manual_seed(0)
net = OthelloNet(n=8)
tree = TreeBatch(G, n=8, mode="neural", capacity=64)
tree.leaf_own.random_(0, 2**62); tree.leaf_opp.copy_(~tree.leaf_own & torch.randint(0, 2**62, (G,), device="cuda"))
ev = FusedOthelloEvaluator(net, tree, heads="library")      # the first version: trunk + three cuBLASLt GEMMs
ev_stack = FusedOthelloEvaluator(net, tree)                 # the product: trunk + az_net_fcstack


def timeit(fn, name):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:28s} {1000 * e0.elapsed_time(e1) / iters:8.2f} us")


def timeit_graph(fn, name, reps=20):
    """Same, with `reps` launches captured in one CUDA graph: removes the ctypes/launch floor (~12 us) of the loop above."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(max(1, iters // reps)):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:28s} {1000 * e0.elapsed_time(e1) / (max(1, iters // reps) * reps):8.2f} us  (in a graph)")


timeit(ev.trunk, "trunk (tcgen05)")
timeit(lambda: torch._addmm_activation(ev.b1, ev.features, ev.w1t, out=ev.h1), "fc1+relu (cuBLASLt)")
timeit(lambda: torch._addmm_activation(ev.b2, ev.h1, ev.w2t, out=ev.h2), "fc2+relu (cuBLASLt)")
timeit(lambda: torch.addmm(ev.bh, ev.h2, ev.wht, out=ev.logits), "heads gemm (cuBLASLt)")
timeit(lambda: _lib.check(_lib.lib().az_net_heads(ev.logits.data_ptr(), ev.ld, ev.A, G, tree.probs.data_ptr(), tree.value.data_ptr(), torch.cuda.current_stream().cuda_stream)), "heads softmax/tanh")
timeit_graph(lambda: torch._addmm_activation(ev.b1, ev.features, ev.w1t, out=ev.h1), "fc1+relu (cuBLASLt)")
timeit_graph(lambda: torch._addmm_activation(ev.b2, ev.h1, ev.w2t, out=ev.h2), "fc2+relu (cuBLASLt)")
timeit_graph(lambda: torch.addmm(ev.bh, ev.h2, ev.wht, out=ev.logits), "heads gemm (cuBLASLt)")
timeit_graph(lambda: _lib.check(_lib.lib().az_net_heads(ev.logits.data_ptr(), ev.ld, ev.A, G, tree.probs.data_ptr(), tree.value.data_ptr(), torch.cuda.current_stream().cuda_stream)), "heads softmax/tanh")
timeit_graph(lambda: _lib.check(_lib.lib().az_net_headgemm(ev.h2.data_ptr(), ev.h2_features, ev.head_blob.data_ptr(), ev.NH, ev.A, G, tree.probs.data_ptr(), tree.value.data_ptr(), torch.cuda.current_stream().cuda_stream)), "heads fused (tcgen05)")
def _fcstack():
    e = ev_stack
    _lib.check(_lib.lib().az_net_fcstack(e.features.data_ptr(), e.stack_blob.data_ptr(), e.h1.data_ptr(), e.h2.data_ptr(),
                                         e.logits.data_ptr(), e.ld, e.A + 1, G, e.K1, e.N1, e.N2, e.RT, e.NS,
                                         torch.cuda.current_stream().cuda_stream), "az_net_fcstack")


timeit_graph(_fcstack, f"fc1+fc2+heads az_net_fcstack RT={ev_stack.RT} NS={ev_stack.NS}")
# absolute CTA start / end times (globaltimer) of the fc stack when it runs right after the trunk, as in the search graph
_n_ctas = ev_stack.tiles * ev_stack.NS
_dbg = torch.zeros(_n_ctas * 32, dtype=torch.int64, device="cuda")
_lib.lib().az_net_fcstack_debug(_dbg.data_ptr())
for _ in range(3):
    ev_stack.enqueue()
torch.cuda.synchronize()
_lib.lib().az_net_fcstack_debug(None)
_d = _dbg.reshape(_n_ctas, 32).double()
_t0 = _d[:, 25].min()
_starts, _ends = (_d[:, 25] - _t0) / 1e3, (_d[:, 19] - _t0) / 1e3
print(f"fcstack after trunk: CTA entry 0 .. {_starts.max().item():.2f} us (median {_starts.median().item():.2f}), "
      f"CTA exit {_ends.min().item():.2f} .. {_ends.max().item():.2f} us; entries later than 5 us: {int((_starts > 5).sum())} of {_n_ctas}")
print("   entry histogram (us):", torch.histc(_starts.float(), bins=10, min=0, max=max(1.0, _starts.max().item())).int().tolist())
timeit_graph(ev_stack.trunk, "trunk, packed features")
timeit_graph(ev.trunk, "trunk, row-major features")
timeit_graph(ev_stack.enqueue, "trunk + fcstack (what the search graph holds)")
timeit_graph(ev.enqueue, "trunk + 3 cuBLASLt GEMMs (first version)")
timeit(ev_stack, "whole forward (graph, stack)")
timeit(ev, f"whole forward (graph, {ev.heads})")
ev_fused = FusedOthelloEvaluator(net, tree, heads="fused")
timeit(ev_fused, "whole forward (graph, fused)")
timeit(lambda: tree.step(200, True, True), "tree step (idle leaves)")

# cycle accounting of the trunk's MMA warp / epilogue warps (debug_layer = 100)
dbg = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
for _ in range(3):      # back to back, like the search loop: the last launch is reported
    _lib.check(_lib.lib().az_net_trunk(tree.leaf_own.data_ptr(), tree.leaf_opp.data_ptr(), 8, 8, G, ev.wblob.data_ptr(),
                                       ev.features.data_ptr(), 0, dbg.data_ptr(), 100, torch.cuda.current_stream().cuda_stream), "trunk timing")
torch.cuda.synchronize()
d = dbg.reshape(148, 16).double()
names = ["mma: wait ready", "mma: wait empty", "mma: issue+commit", "mma: loop total", "mma: wait im2col", "epi team0: wait full",
         "epi team1: wait full", "epi warp0: work", "entry -> setup done", "entry -> weights landed", None, None, "entry -> exit"]
for i, nm in enumerate(names):
    if nm:
        print(f"{nm:24s} median {d[:, i].median().item():9.0f} cycles   max {d[:, i].max().item():9.0f}")
g0, g1 = d[:, 10], d[:, 11]
print(f"globaltimer: first CTA entry -> last CTA entry {(g0.max() - g0.min()).item() / 1e3:6.2f} us, "
      f"first entry -> last exit {(g1.max() - g0.min()).item() / 1e3:6.2f} us")
