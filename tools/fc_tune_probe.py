"""Does cuBLASLt algorithm tuning (PyTorch TunableOp) speed up the three library GEMMs of the fused forward?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
G = 4096
x = torch.randn(G, 512, device="cuda", dtype=torch.bfloat16)
w1 = torch.randn(512, 1024, device="cuda", dtype=torch.bfloat16); b1 = torch.randn(1024, device="cuda", dtype=torch.bfloat16)
w2 = torch.randn(1024, 512, device="cuda", dtype=torch.bfloat16); b2 = torch.randn(512, device="cuda", dtype=torch.bfloat16)
wh = torch.randn(512, 72, device="cuda", dtype=torch.bfloat16); bh = torch.randn(72, device="cuda", dtype=torch.bfloat16)
h1 = torch.empty(G, 1024, device="cuda", dtype=torch.bfloat16); h2 = torch.empty(G, 512, device="cuda", dtype=torch.bfloat16)
lg = torch.empty(G, 72, device="cuda", dtype=torch.bfloat16)

def fwd():
    torch._addmm_activation(b1, x, w1, out=h1)
    torch._addmm_activation(b2, h1, w2, out=h2)
    torch.addmm(bh, h2, wh, out=lg)

def graph_time(label):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(5):
            fwd()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fwd()
    for _ in range(20):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(500):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{label:40s} {e0.elapsed_time(e1) / 500 * 1e3:7.2f} us per forward (3 GEMMs)")

graph_time("default heuristics")
try:
    import torch.cuda.tunable as tn
    tn.enable(True); tn.tuning_enable(True); tn.set_max_tuning_duration(200); tn.set_max_tuning_iterations(50)
    fwd(); torch.cuda.synchronize()
    tn.tuning_enable(False)
    graph_time("TunableOp tuned")
    print(tn.get_results())
except Exception as e:
    print("tunable op unavailable:", repr(e))
# single fused GEMM shapes for reference: what would one (4096 x 512) x (512 x 1024) GEMM cost alone
for (a, b, name) in ((x, w1, "fc1"), (h1, w2, "fc2"), (h2, wh, "heads")):
    out = torch.empty(a.shape[0], b.shape[1], device="cuda", dtype=torch.bfloat16)
    g = torch.cuda.CUDAGraph()
    torch.mm(a, b, out=out); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(20):
            torch.mm(a, b, out=out)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 1000 * 1e3:6.2f} us per GEMM back to back")
