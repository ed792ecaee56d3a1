"""cycles per tcgen05.mma on this GPU for the shapes the trunk could use (az_peak_mma); prints a table."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200 import _lib

from tools.probes import build as _probes
L = _probes.lib()
cyc = torch.zeros(148, dtype=torch.int64, device="cuda")


def run(m, n, ks, regions, flags):
    reps = 768 // ks
    _lib.check(L.az_peak_mma(m, n, ks, reps, regions, flags, cyc.data_ptr(), torch.cuda.current_stream().cuda_stream), "probe")
    torch.cuda.synchronize()
    return float(cyc.median().item()) / reps


print("# tcgen05.mma.cta_group::1.kind::f16 (bf16 SS, K = 16, no-swizzle K-major), median over 148 SMs")
print("# flags: 1 = all MMAs accumulate, 2 = commit per chain, 4 = concurrent tcgen05.ld readers")
print(f"{'m':>4} {'n':>4} {'ksteps':>6} {'regions':>7} {'flags':>5} {'cyc/chain':>9} {'cyc/mma':>8} {'MAC/cyc':>8}")
cases = []
for n in (32, 96, 192, 256):
    for ks in (1, 2, 3, 6, 8):
        cases.append((128, n, ks, 2, 0))
for flags in (1, 2, 3, 4, 6, 7):
    for ks in (1, 6):
        cases.append((128, 96, ks, 2, flags))
cases += [(64, 96, 6, 2, 0), (128, 96, 6, 1, 2)]
for m, n, ks, regions, flags in cases:
    c = run(m, n, ks, regions, flags)
    print(f"{m:4d} {n:4d} {ks:6d} {regions:7d} {flags:5d} {c:9.1f} {c / ks:8.1f} {m * n * 16 * ks / c:8.0f}")
