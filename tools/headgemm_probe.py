"""az_net_headgemm timing sweep over K (features), NH (padded outputs) and G, each as 20 launches in one CUDA graph.
With the argument `ncu`: a few plain launches (target for ncu -k regex:headgemm)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200 import _lib

L = _lib.lib()
dev = "cuda"


def run(G, K, NH, A, reps=20, iters=10):
    h2 = torch.randn(G, K, device=dev).to(torch.bfloat16)
    blob = torch.zeros(int(L.az_net_headgemm_blob_bytes(K, NH)), dtype=torch.uint8, device=dev)
    probs = torch.empty(G, A, device=dev); value = torch.empty(G, device=dev)
    fn = lambda: _lib.check(L.az_net_headgemm(h2.data_ptr(), K, blob.data_ptr(), NH, A, G, probs.data_ptr(), value.data_ptr(),
                                              torch.cuda.current_stream().cuda_stream))
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return 1000 * e0.elapsed_time(e1) / (reps * iters)


if len(sys.argv) > 1 and sys.argv[1] == "ncu":
    run(4096, 512, 80, 65, reps=1, iters=3)
else:
    for G in (128, 4096, 148 * 128):
        for K, NH, A in ((32, 16, 7), (128, 16, 7), (512, 16, 7), (32, 80, 65), (128, 80, 65), (256, 80, 65), (512, 80, 65)):
            print(f"G={G:6d} K={K:4d} NH={NH:3d} A={A:3d}  {run(G, K, NH, A):7.2f} us", flush=True)
