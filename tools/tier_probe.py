"""Per-ply search time by batch size (main batch / tail tiers) over one self-play batch."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200.selfplay import SelfPlayEngine

game = sys.argv[1] if len(sys.argv) > 1 else "othello"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
torch.manual_seed(0)
if game == "connect4":
    from alphazero_b200.games.connect4 import Connect4Net
    from alphazero_b200.gamespec import Connect4Spec
    net, spec = Connect4Net(board_width=7, board_height=6), Connect4Spec(7, 6)
else:
    from alphazero_b200.games.othello import OthelloNet
    net, spec = OthelloNet(n=8), 8
for tiers in ((), "auto"):
    eng = SelfPlayEngine(net, G, n=spec, n_sim=200, seed=1, tail_tiers=tiers)
    for _ in range(3):
        eng.play()
    eng.search_events = []
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.play(); e1.record(); torch.cuda.synchronize()
    by = collections.defaultdict(list)
    for a, b, k, size in eng.search_events:
        by[size].append(a.elapsed_time(b))
    print(f"{game} tiers={eng.tail_tiers}: batch {e0.elapsed_time(e1):.1f} ms, sims {eng.total_sims}, "
          f"{eng.total_sims / e0.elapsed_time(e1) / 1e3:.2f} M sims/s")
    for size, v in sorted(by.items(), reverse=True):
        print(f"   batch size {size:5d}: {len(v):3d} plies, {sum(v) / len(v):7.2f} ms per search, total {sum(v):7.1f} ms")
    del eng
    torch.cuda.empty_cache()
