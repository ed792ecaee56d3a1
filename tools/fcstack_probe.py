"""az_net_fcstack on random operands: checks it against float64 torch and times it inside a CUDA graph, for several
cluster shapes, next to the three cuBLASLt GEMMs it replaces.

    python tools/fcstack_probe.py [G] [iters]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200 import _lib
from alphazero_b200.fused_net import pack_fcstack, pack_rows

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 400
dev = "cuda"


def time_graph(fn, reps=20):
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
    n = max(1, iters // reps)
    e0.record()
    for _ in range(n):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return 1000 * e0.elapsed_time(e1) / (n * reps)


def run(K1, N1, N2, n_out, shapes, G):
    torch.manual_seed(K1 + n_out)
    w1 = torch.randn(N1, K1, dtype=torch.float64) / K1 ** 0.5
    w2 = torch.randn(N2, N1, dtype=torch.float64) / N1 ** 0.5
    wh = torch.randn(n_out, N2, dtype=torch.float64) / N2 ** 0.5
    b1, b2, bh = (torch.randn(n, dtype=torch.float64) * 0.1 for n in (N1, N2, n_out))
    x = torch.randn(G, K1).to(torch.bfloat16)
    bf = lambda t: t.float().to(torch.bfloat16)
    xd = x.to(dev)
    # reference: float64 products of the bf16-rounded operands, bf16 rounding of h1 / h2 like the kernel's epilogues
    h1 = bf(torch.relu(xd.double() @ bf(w1).to(dev).double().t() + b1.float().double().to(dev)))
    h2 = bf(torch.relu(h1.double() @ bf(w2).to(dev).double().t() + b2.float().double().to(dev)))
    want = h2.double() @ bf(wh).to(dev).double().t() + bh.float().double().to(dev)
    ld = ((n_out + 7) // 8) * 8
    print(f"--- {K1} -> {N1} -> {N2} -> {n_out}, G = {G}")
    for RT, NS in shapes:
        blob, plan = pack_fcstack(w1, b1, w2, b2, wh, bh, NS, RT)
        blob = blob.to(dev)
        tiles = -(-((G + 127) // 128) // RT) * RT
        xp = torch.zeros(tiles * 128, K1, dtype=torch.bfloat16)
        xp[:G] = x
        xp = pack_rows(xp, plan["kch"][0]).to(dev)
        s1 = torch.zeros(tiles * 128 * N1, dtype=torch.bfloat16, device=dev)
        s2 = torch.zeros(tiles * 128 * N2, dtype=torch.bfloat16, device=dev)
        logits = torch.full((G, ld), -77.0, dtype=torch.bfloat16, device=dev)

        def call():
            _lib.check(_lib.lib().az_net_fcstack(xp.data_ptr(), blob.data_ptr(), s1.data_ptr(), s2.data_ptr(),
                                                 logits.data_ptr(), ld, n_out, G, K1, N1, N2, RT, NS,
                                                 torch.cuda.current_stream().cuda_stream), "az_net_fcstack")
        call()
        torch.cuda.synchronize()
        err = (logits[:, :n_out].double() - want).abs().max().item()
        pad_ok = bool((logits[:, n_out:] == -77.0).all().item())
        scale = want.abs().max().item()
        if os.environ.get("FCSTACK_STAMPS"):
            n_ctas = tiles * NS
            dbg = torch.zeros(n_ctas * 32, dtype=torch.int64, device=dev)
            _lib.lib().az_net_fcstack_debug(dbg.data_ptr())
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            _lib.lib().az_net_fcstack_debug(None)
            d = dbg.reshape(n_ctas, 32).double()
            names = {1: "setup done (cluster sync, pdl wait)", 26: "producer: first wait L1", 2: "L1 first stage full",
                     3: "L1 last MMA issued", 4: "L1 accumulator complete (epilogue wakes)", 5: "L1 epilogue done",
                     27: "producer: first wait L2", 8: "L2 first stage full", 9: "L2 last MMA issued", 10: "L2 accumulator complete",
                     11: "L2 epilogue done", 28: "producer: first wait L3", 14: "L3 first stage full", 15: "L3 last MMA issued",
                     16: "L3 accumulator complete", 17: "L3 epilogue done", 24: "exit (after teardown cluster sync)",
                     29: "L1 epilogue: first block in registers", 6: "L1 epilogue: first block stored", 7: "L1 epilogue: first block signalled",
                     30: "L2 epilogue: first block in registers", 12: "L2 epilogue: first block stored", 13: "L2 epilogue: first block signalled",
                     31: "L3 epilogue: accumulator in registers",
                     20: "producer: total wait empty", 21: "producer: total wait ready", 22: "mma: total wait full",
                     23: "mma: total issue"}
            if plan["variant"] >= 2:      # stamp slots of k_net_fcstack2 / 3 (csrc/fcstack2_kernels.cu, fcstack3_kernels.cu)
                names = {1: "setup done (cluster sync, pdl wait)", 2: "fc1 first stage full", 3: "fc1 pass 0 issued",
                         4: "fc1 pass 1 issued", 5: "pass 0 accumulator seen by the epilogue", 6: "pass 0 blocks published",
                         7: "pass 1 accumulator seen", 8: "pass 1 blocks published", 9: "fc2 first stage full",
                         10: "fc2 middle stage full", 11: "fc2 last MMA issued", 12: "fc2 accumulator seen",
                         13: "h2 in shared memory", 14: "heads: operands ready", 15: "heads issued",
                         16: "heads accumulator seen", 17: "partial sums sent", 18: "after the cluster barrier",
                         24: "exit (logits stored)"}
                if plan["variant"] == 3:
                    names.update({18: "owner: peers' partial sums seen", 20: "leader: total wait own full",
                                  21: "leader: total wait peer full", 22: "leader: total issue + commit",
                                  26: "owner: logits stored", 23: "before the pair barrier"})
            for i, nm in names.items():
                col = d[:, i]
                if plan["variant"] == 3 and i in (2, 3, 4, 9, 10, 11, 14, 15, 20, 21, 22):
                    col = d[0::2, i]      # stamps of the control warp of rank 0
                print(f"    {nm:44s} median {col.median().item():8.0f}  min {col.min().item():8.0f}  max {col.max().item():8.0f} cycles")
        us = time_graph(call)
        flops = 2.0 * G * (K1 * N1 + N1 * N2 + N2 * n_out)
        print(f"RT={RT} NS={NS} variant {plan['variant']}: max |err| {err:.4f} (logit scale {scale:.2f}, bf16 ulp {scale * 2 ** -8:.4f}) pad untouched {pad_ok}"
              f" | {us:7.2f} us in a graph = {flops / us / 1e6:7.1f} TFLOP/s")
    # the library path it replaces
    w1t, w2t, wht = (bf(w).t().contiguous().to(dev) for w in (w1, w2, torch.cat([wh, torch.zeros(ld - n_out, N2, dtype=torch.float64)])))
    bb1, bb2, bbh = bf(b1).to(dev), bf(b2).to(dev), bf(torch.cat([bh, torch.zeros(ld - n_out, dtype=torch.float64)])).to(dev)
    o1 = torch.empty(G, N1, dtype=torch.bfloat16, device=dev); o2 = torch.empty(G, N2, dtype=torch.bfloat16, device=dev)
    o3 = torch.empty(G, ld, dtype=torch.bfloat16, device=dev)

    def lib3():
        torch._addmm_activation(bb1, xd, w1t, out=o1)
        torch._addmm_activation(bb2, o1, w2t, out=o2)
        torch.addmm(bbh, o2, wht, out=o3)
    with torch.no_grad():
        us = time_graph(lib3)
    print(f"cuBLASLt fc1 + fc2 + heads: {us:7.2f} us in a graph")


if __name__ == "__main__":
    which = os.environ.get("FCSTACK_PROBE", "all")
    if which in ("all", "othello8"):
        run(512, 1024, 512, 66, [(2, 4), (1, 4), (1, 8)], G)
    if which == "othello8_14":
        run(512, 1024, 512, 66, [(1, 4)], G)
    if which == "v3":
        run(512, 1024, 512, 66, [(2, 4), (1, 4)], G)
        run(512, 1024, 512, 66, [(2, 4)], 300)
        run(128, 1024, 512, 38, [(2, 4)], 1000)
    if which == "v2":
        run(512, 1024, 512, 66, [(1, 4)], G)
        run(512, 1024, 512, 66, [(1, 4)], 300)
        run(128, 1024, 512, 38, [(1, 4)], 1000)
        run(192, 512, 256, 80, [(1, 4)], 333)
    if which in ("all", "small"):
        run(512, 1024, 512, 66, [(2, 4), (1, 4)], 300)
        run(128, 1024, 512, 38, [(2, 4), (1, 8)], 1000)
        run(192, 64, 32, 8, [(1, 1), (2, 1)], 2048)
