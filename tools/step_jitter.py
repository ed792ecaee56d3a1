"""Per-step times of SelfPlayEngine.play() over many batches, with the garbage collector's pauses beside them: where do
outlier steps come from?     python tools/step_jitter.py [steps] [games] [smi] [e2e]
smi: an `nvidia-smi -lms 200` poller runs beside the steps (as in bench.py); e2e: every other step reads its samples back
to pinned host memory (as bench.py's end-to-end steps do)."""
import gc
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from alphazero_b200.games.othello import OthelloNet
from alphazero_b200.selfplay import SelfPlayEngine

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
G = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
SMI, E2E = "smi" in sys.argv[3:], "e2e" in sys.argv[3:]
torch.manual_seed(0)
eng = SelfPlayEngine(OthelloNet(n=8), G, n=8, n_sim=200, dirichlet_alpha=0.03, dirichlet_epsilon=0.25,
                     temp_scheduler_type="linear", temp_max_step=4, temp_min_step=4, seed=0x5EED, dtype=torch.bfloat16,
                     device=torch.device("cuda"))
pauses = []
t_gc = [0.0]


def cb(phase, info):
    if phase == "start":
        t_gc[0] = time.perf_counter()
    else:
        pauses.append((info["generation"], 1000 * (time.perf_counter() - t_gc[0])))


gc.callbacks.append(cb)
poller = None
if SMI:
    import subprocess
    poller = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                               "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                               "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    time.sleep(2.0)
from alphazero_b200.selfplay import HostStaging
staging = HostStaging()
for _ in range(3):
    eng.play()
torch.cuda.synchronize()
for i in range(steps):
    del pauses[:]
    t0 = time.perf_counter()
    b = eng.play()
    host = b.cpu(staging) if (E2E and i % 2) else None
    torch.cuda.synchronize()
    ms = 1000 * (time.perf_counter() - t0)
    big = [(g, round(p, 1)) for g, p in pauses if p > 2.0]
    print(f"step {i:3d}: {ms:7.1f} ms, {len(pauses)} collections ({sum(p for _, p in pauses):6.1f} ms), over 2 ms: {big}", flush=True)
    del b, host
if poller:
    poller.terminate()
