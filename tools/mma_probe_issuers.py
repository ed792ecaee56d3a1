import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200 import _lib
from tools.probes import build as _probes
L = _probes.lib()
cyc = torch.zeros(148, dtype=torch.int64, device="cuda")
def run(m, n, ks, flags):
    reps = 768 // ks
    _lib.check(L.az_peak_mma(m, n, ks, reps, 1, flags, cyc.data_ptr(), torch.cuda.current_stream().cuda_stream), "probe")
    torch.cuda.synchronize()
    return float(cyc.median().item()) / (reps * ks)
print("# cycles per MMA seen by issuer 0 (each of W warps issues the same number of MMAs into its own accumulator)")
print("# => pipe cycles per MMA = value / W")
for n in (32, 64, 96, 128):
    for ks in (2, 6, 8):
        row = [run(128, n, ks, (w - 1) << 4) for w in (1, 2, 3, 4)]
        print(f"N={n:3d} ks={ks}: " + "  ".join(f"W={w}: {c:6.1f} ({c / w:5.1f}/mma)" for w, c in zip((1, 2, 3, 4), row)))
