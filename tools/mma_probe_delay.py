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
for delay in (0, 16, 32, 64, 96, 128, 192, 256):
    f = 8 | 2 | (delay << 8)
    print(f"delay {delay:4d} ALU instr: ks=6 cyc/chain {run(128, 96, 6, 2, f):8.1f}   ks=1 {run(128, 96, 1, 2, f):8.1f}   ks=6 N=256 {run(128, 256, 6, 2, f):8.1f}")
