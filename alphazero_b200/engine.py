"""TreeBatch: G concurrent MCTS trees on one GPU (host side of the tree kernels, include/az_b200.h).

This is the batched counterpart of alphazero.mcts.MCT (mcts.py:49-269): one object owns the node pools, root
boards and per-step leaf buffers of G games; `search` runs MCT.search for all of them at once, one simulation per
game per step (the reference's sequential order per game), with the leaf evaluation delegated to an evaluator:

    evaluator(tree)  must fill  tree.probs [G, n*n+1] float32 (network probabilities, pass last)
                                tree.value [G] float32      (network value, side to move's view)
                     from       tree.enc   [G, n, n]         (player*grid, written by the select kernel)
                     or         tree.leaf_own / tree.leaf_opp (side-relative bitboards of the leaves).

Evaluators: alphazero_b200.evaluators.{NetEvaluator, HashEvaluator}; rollout mode needs none (az_tree_rollout).
"""
import ctypes
import itertools
import time

import torch

from . import _lib
from .boardops import _ptr, _stream, to_i64
from .gamespec import as_spec

NODE_BYTES = 32
_EVALUATOR_UIDS = itertools.count(1)      # process-wide: id() of a dropped evaluator can be handed out again


def _evaluator_uid(evaluator) -> int:
    uid = getattr(evaluator, "_az_uid", None)
    if uid is None:
        uid = next(_EVALUATOR_UIDS)
        evaluator._az_uid = uid
    return uid


class LeafView:
    """What an evaluator sees of a throughput-mode TreeBatch: the flat batch of G * leaves leaf slots (slot l*G + g =
    descent l of game g) with the attributes evaluators read from a tree."""

    def __init__(self, tree: "TreeBatch"):
        self.tree = tree
        self.G = tree.n_slots
        self.spec, self.n, self.A, self.device = tree.spec, tree.n, tree.A, tree.device
        for name in ("leaf_own", "leaf_opp", "leaf_player", "leaf_kind", "enc", "probs", "value"):
            setattr(self, name, getattr(tree, name))

    def set_encoding(self, on: bool) -> None:
        self.tree.set_encoding(on)


class TreeBatch:
    def __init__(self, G: int, n=8, mode: str = "neural", capacity: int = None, n_sim_hint: int = 200,
                 device="cuda", seed: int = 0, tie_random: bool = True, dirichlet_alpha: float = None,
                 dirichlet_epsilon: float = None, enc_dtype=torch.float32, max_depth: int = 128,
                 graph_search: bool = True, graph_first_use: bool = False, leaves: int = None):
        """leaves: None = parity mode, the reference's one leaf per game per simulation step (mcts.py:254-269).
        An integer L >= 1 = throughput mode (neural only; an extension, include/az_b200.h az_tree.n_leaves): up to L
        leaves of a game in flight per step under virtual loss, the evaluator sees G*L leaf slots (`leaf_view`).
        L = 1 is bit-identical to parity mode."""
        if not torch.cuda.is_available():
            raise RuntimeError("TreeBatch needs a CUDA device (alphazero_b200 has no CPU path)")
        if mode not in ("neural", "rollout"):
            raise ValueError(f"unknown mode {mode}")
        # `n`: Othello board size, or a GameSpec (gamespec.OthelloSpec / Connect4Spec)
        self.spec = as_spec(n)
        self.G, self.n, self.mode = int(G), self.spec.n, mode
        self.A = self.spec.action_size
        self.device = torch.device(device)
        if leaves is not None and (mode != "neural" or not 1 <= int(leaves) <= 8):
            raise ValueError("leaves (throughput mode) needs mode='neural' and 1 <= leaves <= 8")
        self.leaves = None if leaves is None else int(leaves)
        self.n_slots = self.G * (self.leaves or 1)       # leaf slots the evaluator sees
        if capacity is None:
            # per search at most (n_sim + 1) child blocks are added and the re-rooted subtree is kept.  Measured on
            # Othello 8x8 self-play (n_sim = 200, tools/pool_use.py): the fullest of 4096 games peaks at 5.4-7.6 k nodes
            # and over several hundred thousand games 48 * (n_sim + 1) = 9.6 k was exceeded once -- the default is twice
            # that (5 GB for 4096 games).  Overflow is detected (status bit) and raised, never silent; SelfPlayEngine
            # answers it by doubling the pools and replaying the batch.
            capacity = max(1024, 96 * (n_sim_hint + 1))
        self.capacity = int(capacity)
        self.max_depth = int(max_depth)
        dev = self.device
        G = self.G
        z = lambda *shape, dtype: torch.zeros(*shape, dtype=dtype, device=dev)
        self.nodes = z(G * 2 * self.capacity * NODE_BYTES, dtype=torch.uint8)
        self.half = z(G, dtype=torch.uint8)
        self.n_alloc = z(G, dtype=torch.int32)
        self.root_own = z(G, dtype=torch.int64)
        self.root_opp = z(G, dtype=torch.int64)
        self.root_player = torch.ones(G, dtype=torch.int8, device=dev)
        self.active = torch.ones(G, dtype=torch.uint8, device=dev)
        self.sims_done = z(G, dtype=torch.int32)
        self.rng_ctr = z(G, dtype=torch.int32)
        self.status = z(G, dtype=torch.int32)
        S = self.n_slots
        self.leaf_node = z(S, dtype=torch.int32)
        self.leaf_kind = z(S, dtype=torch.uint8)
        self.leaf_own = z(S, dtype=torch.int64)
        self.leaf_opp = z(S, dtype=torch.int64)
        self.leaf_player = z(S, dtype=torch.int8)
        self.path = z(S * self.max_depth, dtype=torch.int32)
        self.path_len = z(S, dtype=torch.int32)
        self.enc_dtype = enc_dtype
        self.enc = z(S, self.spec.rows, self.spec.cols, dtype=enc_dtype)
        self.probs = z(S, self.A, dtype=torch.float32)
        self.value = z(S, dtype=torch.float32)
        self.n_rollouts = 0
        self.simulation_time = 0.0
        self.profile = None   # set to {"eval": [], "tree": []} to record (start, end) CUDA events per call
        # a whole search (steps x (leaf evaluation, tree step)) as ONE CUDA graph from its second use on: one launch
        # per move instead of 2 * n_sim, so a busy host cannot starve the device
        self.graph_search = bool(graph_search)
        self.graph_first_use = bool(graph_first_use)   # capture at the first use (batches that exist to be reused)
        self.graph_step_launches = 0  # az_tree_step kernels replayed inside search graphs (not seen by _lib.LAUNCH_CALLS)
        # (n_sim, steps, evaluator uid, evaluator.version, descriptor fields baked into the launches)
        #   -> [CUDAGraph | None (seen once), evaluator (kept alive: the graph replays into its buffers)]
        self._search_graphs = {}
        self._desc = _lib.AzTree()
        d = self._desc
        for name in ("nodes", "half", "n_alloc", "root_own", "root_opp", "root_player", "active", "sims_done",
                     "rng_ctr", "status", "leaf_node", "leaf_kind", "leaf_own", "leaf_opp", "leaf_player", "path",
                     "path_len", "enc", "probs", "value"):
            setattr(d, name, getattr(self, name).data_ptr())
        d.seed = int(seed) & ((1 << 64) - 1)
        d.G, d.n, d.capacity, d.max_depth = self.G, self.n, self.capacity, self.max_depth
        d.game, d.height = self.spec.game_id, self.spec.height
        d.mode = _lib.MODE_NEURAL if mode == "neural" else _lib.MODE_ROLLOUT
        d.tie_random = 1 if tie_random else 0
        d.enc_dtype = {torch.float32: 0, torch.bfloat16: 1}[enc_dtype]
        d.n_leaves = self.leaves or 0
        self.leaf_view = LeafView(self) if (self.leaves or 1) > 1 else self    # what evaluators are built over
        self.set_dirichlet(dirichlet_alpha, dirichlet_epsilon)
        self.reset_roots()

    def __getstate__(self):
        raise TypeError("TreeBatch owns device memory (node pools, CUDA graphs) and cannot be pickled; send the "
                        "position (root_own / root_opp / root_player) and build a TreeBatch in the receiving process. "
                        "MCT / MCTSPlayer / AlphaZeroPlayer pickle without their device tree (mcts.MCT.__getstate__).")

    # ------------------------------------------------------------------ configuration
    def set_encoding(self, on: bool) -> None:
        """Switches the float network encoding written by the select kernel on/off (evaluators that read the leaf
        bitboards directly, e.g. FusedOthelloEvaluator, do not need it)."""
        self._desc.enc = self.enc.data_ptr() if on else None

    def set_dirichlet(self, alpha, epsilon) -> None:
        """Both None (or <= 0) switches the root noise off (mcts.py:235)."""
        self.dirichlet_alpha, self.dirichlet_epsilon = alpha, epsilon
        on = alpha is not None and epsilon is not None
        self._desc.dirichlet_alpha = float(alpha) if on else 0.0
        self._desc.dirichlet_eps = float(epsilon) if on else 0.0

    @property
    def desc(self):
        return ctypes.byref(self._desc)

    # ------------------------------------------------------------------ roots
    def reset_roots(self, mask: torch.Tensor = None) -> None:
        """Start position + fresh trees for the masked games (board.reset() + player.reset())."""
        black, white = self.spec.start_position()
        if mask is None:
            self.root_own.fill_(to_i64(black)); self.root_opp.fill_(to_i64(white)); self.root_player.fill_(1)
            self.active.fill_(1)
        else:
            m = mask.bool()
            self.root_own[m] = to_i64(black); self.root_opp[m] = to_i64(white); self.root_player[m] = 1
            self.active[m] = 1
        self.reset_trees(mask)

    def set_roots(self, own: torch.Tensor, opp: torch.Tensor, player: torch.Tensor, fresh_trees: bool = True) -> None:
        self.root_own.copy_(own); self.root_opp.copy_(opp); self.root_player.copy_(player)
        if fresh_trees:
            self.reset_trees()

    def adopt(self, src: "TreeBatch", sel: torch.Tensor, n_nodes: int = None) -> None:
        """Moves the games `sel` (indices into `src`) into slots 0..k-1 of this batch: root position, kept search tree
        (the live half of the node pool, its first `n_nodes` records -- default the whole half), random counter and
        sticky status.  Slots k.. are switched off.  Used by SelfPlayEngine to carry the last live games of a batch
        over to a smaller batch (smaller leaf batches for the network) when most games have ended."""
        if (src.spec.game_id, src.n, src.spec.height, src.capacity, src.mode, src.leaves) != \
                (self.spec.game_id, self.n, self.spec.height, self.capacity, self.mode, self.leaves):
            raise ValueError("adopt: the two tree batches differ in game, pool capacity, mode or leaves in flight")
        k = int(sel.numel())
        if k > self.G:
            raise ValueError(f"adopt: {k} games do not fit {self.G} slots")
        sel = sel.to(device=src.device, dtype=torch.int64)
        for name in ("root_own", "root_opp", "root_player", "n_alloc", "rng_ctr", "status"):
            getattr(self, name)[:k] = getattr(src, name)[sel]
        words = self.capacity * NODE_BYTES // 8
        m = words if n_nodes is None else min(words, int(n_nodes) * NODE_BYTES // 8)
        sv = src.nodes.view(torch.int64).view(src.G, 2, words)
        dv = self.nodes.view(torch.int64).view(self.G, 2, words)
        dv[:k, 0, :m] = sv[sel, src.half[sel].to(torch.int64), :m]
        self.half[:k] = 0
        self.leaf_kind.zero_()
        self.active.zero_()
        self.active[:k] = 1

    def reset_trees(self, mask: torch.Tensor = None) -> None:
        m = None if mask is None else mask.to(torch.uint8).contiguous()
        _lib.check(_lib.lib().az_tree_reset(self.desc, _ptr(m), _stream()), "az_tree_reset")

    # ------------------------------------------------------------------ search
    def step(self, n_sim: int, finish: bool, select: bool) -> None:
        _lib.check(_lib.lib().az_tree_step(self.desc, int(n_sim), int(finish), int(select), _stream()), "az_tree_step")

    def rollout_leaves(self) -> None:
        _lib.check(_lib.lib().az_tree_rollout(self.desc, _stream()), "az_tree_rollout")

    def search(self, n_sim: int, evaluator=None, extra_steps: int = 1, check: bool = True) -> None:
        """MCT.search(board, n_sim) for every active game (mcts.py:254-269).

        `extra_steps` = 1 covers the root-priming evaluation of fresh roots (mcts.py:231-233), which costs a step but
        no simulation; pass 0 when every root is known to be evaluated already (tree reuse in self-play)."""
        if self.mode == "neural" and evaluator is None:
            raise ValueError("neural mode needs an evaluator")
        start = time.time()
        # an evaluator whose raw output rows the tree step reads itself (az_tree.logits); else probs / value
        binding = getattr(evaluator, "logits_binding", None) if self.mode == "neural" else None
        self._desc.logits, self._desc.logits_ld = (binding[0], int(binding[1])) if binding else (None, 0)
        self.sims_done.zero_()
        self.step(n_sim, finish=False, select=True)
        steps = n_sim + (extra_steps if self.mode == "neural" else 0)
        L = self.leaves or 1
        if L > 1:
            # throughput mode: L simulations per step when no descent is dropped, plus a margin for dropped descents
            # (two descents of one step meeting at the same unevaluated node); whatever is still missing after the
            # planned steps is made up one step at a time below
            per = -(-n_sim // L)
            steps = per + (extra_steps if self.mode == "neural" else 0) + max(2, per // 8)
        self.last_search_steps = steps
        prof = self.profile
        graph = self._search_graph(n_sim, steps, evaluator) if (self.graph_search and prof is None) else None
        if graph is not None:
            graph.replay()
            self.graph_step_launches += steps
            if evaluator is not None and hasattr(evaluator, "n_calls"):
                evaluator.n_calls += steps
                if hasattr(evaluator, "own_launches"):
                    evaluator.own_launches += steps * evaluator.own_kernels_per_enqueue
            steps = 0
        for _ in range(steps):
            if prof is not None:
                e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                e0.record()
            if self.mode == "neural":
                evaluator(self.leaf_view)
            else:
                self.rollout_leaves()
            if prof is not None:
                e1.record()
            self.step(n_sim, finish=True, select=True)
            if prof is not None:
                e2.record()
                prof["eval"].append((e0, e1))
                prof["tree"].append((e1, e2))
                prof["depth_sum"] += self.path_len[self.active.bool()].sum()
                prof["depth_cnt"] += self.active.sum()
        if L > 1:
            self.extra_steps_last_search = 0
            while bool(((self.sims_done < n_sim) & self.active.bool()).any().item()):
                evaluator(self.leaf_view)
                self.step(n_sim, finish=True, select=True)
                self.extra_steps_last_search += 1
        self.n_rollouts = int(n_sim)
        if check:
            self.check_status()
            self.simulation_time = time.time() - start

    def _search_graph(self, n_sim: int, steps: int, evaluator):
        """The CUDA graph of `steps` x (leaf evaluation, az_tree_step) for this evaluator, captured the second time the
        same search is asked for (a one-off search never pays the instantiation); None = run the loop.  Everything a
        step reads that changes between searches (roots, pools, random counters, network outputs) lives in device
        buffers whose addresses are fixed for the life of the tree, so a replay is a fresh search.  Evaluators carry
        a `version` that changes when their buffers do (refresh())."""
        if self.mode == "neural":
            enqueue = getattr(evaluator, "enqueue", None)
            if enqueue is None:
                return None
            ident = (_evaluator_uid(evaluator), getattr(evaluator, "version", 0))
        else:
            enqueue, evaluator = self.rollout_leaves, None
            ident = (0, 0)
        # az_tree is passed to the kernels BY VALUE, so everything in it that a setter can change after the capture
        # (set_dirichlet, set_encoding, the logits binding of search) is part of the key: a replay never runs with
        # stale settings, it is re-captured
        d = self._desc
        baked = (d.enc, d.logits, int(d.logits_ld), float(d.dirichlet_alpha), float(d.dirichlet_eps), int(d.tie_random),
                 int(d.seed), int(d.mode))
        key = (int(n_sim), int(steps)) + ident + (baked,)
        if key not in self._search_graphs:
            for k in [k for k in self._search_graphs if k[:3] == key[:3] and k != key]:
                del self._search_graphs[k]                      # graphs over old weights / old settings
            self._search_graphs[key] = [None, evaluator]
            if not self.graph_first_use:
                return None
        entry = self._search_graphs[key]
        if entry[0] is None:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(steps):
                    enqueue()
                    self.step(n_sim, finish=True, select=True)
            entry[0] = g
        return entry[0]

    def check_status(self) -> None:
        """Raises if any game hit a pool overflow / depth overflow / illegal move (synchronises)."""
        st = int(self.status.max().item())
        if st & _lib.STATUS_OVERFLOW:
            raise RuntimeError("TreeBatch: node pool overflow -- raise `capacity`")
        if st & _lib.STATUS_DEPTH:
            raise RuntimeError("TreeBatch: search path longer than max_depth")
        if st & _lib.STATUS_ILLEGAL:
            raise ValueError("TreeBatch: illegal move passed to advance()")

    # ------------------------------------------------------------------ results
    def root_stats(self):
        """(N int32 [G,65] with -1 = no child, Q float64 [G,65], P float32 [G,65], root_N int32 [G]) on the canvas
        action space (a = r*8+c, 64 = pass) -- what root.children exposes in the reference (mcts.py:88-116)."""
        G = self.G
        N = torch.empty(G, 65, dtype=torch.int32, device=self.device)
        Q = torch.empty(G, 65, dtype=torch.float64, device=self.device)
        P = torch.empty(G, 65, dtype=torch.float32, device=self.device)
        rn = torch.empty(G, dtype=torch.int32, device=self.device)
        _lib.check(_lib.lib().az_tree_root_stats(self.desc, _ptr(N), _ptr(Q), _ptr(P), _ptr(rn), _stream()),
                   "az_tree_root_stats")
        return N, Q, P, rn

    def choose(self, temp: torch.Tensor = None):
        """(action uint8 [G], pi float32 [G, n*n+1]): get_action_probs + sampling (mcts.py:95-116,
        players.py:184-189); temp None = 0 for all games.  Entries of inactive games are left untouched (255 / 0)."""
        action = torch.full((self.G,), 255, dtype=torch.uint8, device=self.device)
        pi = torch.zeros(self.G, self.A, dtype=torch.float32, device=self.device)
        t = None if temp is None else temp.to(device=self.device, dtype=torch.float32).contiguous()
        _lib.check(_lib.lib().az_tree_choose(self.desc, _ptr(t), _ptr(action), _ptr(pi), _stream()), "az_tree_choose")
        return action, pi

    def advance(self, action: torch.Tensor) -> None:
        """board.play_move(action) + MCT.change_root(action) for every active game (255 = leave untouched)."""
        a = action.to(device=self.device, dtype=torch.uint8).contiguous()
        _lib.check(_lib.lib().az_tree_advance(self.desc, _ptr(a), _stream()), "az_tree_advance")

    def nodes_used(self) -> torch.Tensor:
        return self.n_alloc.clone()
