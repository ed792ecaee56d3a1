"""Where a self-play move's device time went besides the search in ROUND 1: the ply loop SelfPlayEngine.play had then
(history copies, search, bookkeeping, choose, advance, terminal test + masks, the host sync as separate launches), with
CUDA events around each phase -- profiles/r1_move_phases_*.txt.  The shipped loop is one az_tree_move launch per ply
(DESIGN 5.2); its share is bench.py's `search_share_of_batch`."""
import sys, os, time, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from alphazero_b200.selfplay import SelfPlayEngine
from alphazero_b200.games.othello import OthelloNet

G = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n_sim = int(sys.argv[2]) if len(sys.argv) > 2 else 200
game = sys.argv[3] if len(sys.argv) > 3 else "othello"
torch.manual_seed(0)
if game == "connect4":
    from alphazero_b200.games.connect4 import Connect4Net
    from alphazero_b200.gamespec import Connect4Spec
    eng = SelfPlayEngine(Connect4Net(board_width=7, board_height=6), G, n=Connect4Spec(7, 6), n_sim=n_sim, seed=1, tail_tiers=())
else:
    eng = SelfPlayEngine(OthelloNet(n=8), G, n=8, n_sim=n_sim, seed=1, tail_tiers=())
for _ in range(2):
    eng.play()
tree, spec = eng.tree, eng.spec
marks = []
def mark(name):
    e = torch.cuda.Event(enable_timing=True); e.record(); marks.append((name, e))
tree.reset_roots()
term, _, win = spec.terminal_score(tree.root_own, tree.root_opp, tree.root_player)
live = term == 0
tree.active.copy_(live.to(torch.uint8))
temp_t = torch.empty(G, dtype=torch.float32, device=tree.device)
sims = torch.zeros((), dtype=torch.int64, device=tree.device)
torch.cuda.synchronize()
t = 0; host_wait = 0.0; live_counts = []
mark("start")
while True:
    eng.h_own[t].copy_(tree.root_own); eng.h_opp[t].copy_(tree.root_opp)
    eng.h_player[t].copy_(tree.root_player); eng.h_live[t].copy_(live)
    mark("history")
    tree.search(n_sim, eng.evaluator, extra_steps=1 if t == 0 else 0, check=False)
    mark("search")
    sims += tree.sims_done.to(torch.int64)[live].sum()
    mark("count")
    temp_t.fill_(float(eng.temp[t]))
    action, pi = tree.choose(temp_t)
    eng.h_pi[t].copy_(pi)
    mark("choose")
    tree.advance(action)
    mark("advance")
    term, _, win = spec.terminal_score(tree.root_own, tree.root_opp, tree.root_player)
    live = live & (term == 0)
    tree.active.copy_(live.to(torch.uint8))
    mark("terminal")
    t += 1
    h0 = time.time()
    n_live = int(live.sum().item()); alive = n_live > 0
    live_counts.append(n_live)
    host_wait += time.time() - h0
    mark("sync")
    if not alive:
        break
torch.cuda.synchronize()
tot = collections.defaultdict(float)
for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
    tot[n1] += e0.elapsed_time(e1)
total = sum(tot.values())
print(f"G={G} n_sim={n_sim} plies={t} total {total:.1f} ms; host blocked in .item() {1000 * host_wait:.1f} ms")
for k, v in tot.items():
    print(f"  {k:9s} {v:8.2f} ms  {100 * v / total:5.2f} %  {1000 * v / t:8.1f} us per ply")
print("live games after each ply:", live_counts)
