"""Batched self-play: the trainer's sequential episode loop (alphazero/trainer.py:215-286) for G concurrent games.

Every game follows exactly the reference's per-episode procedure -- AlphaZeroPlayer.get_move with the temperature
of the move index, Sample(state, pi, player), play_move, apply_move (tree reuse), outcome back-filled and
normalised to player = +1 -- but all games advance in lock-step on the device: one TreeBatch.search per move for
all live games, one leaf per game per simulation step, the network evaluated on the whole leaf batch.
"""
import dataclasses

import numpy as np
import torch

from . import _lib
from . import boardops as bo
from .boardops import _ptr, _stream
from .engine import TreeBatch
from .fused_net import make_evaluator
from .gamespec import as_spec
from .schedulers import TEMP_SCHEDULERS


class HostStaging:
    """Pinned host buffers for repeated device -> host reads (SampleBatch.cpu(staging)): one buffer per field, reused
    from call to call and re-allocated with 50 % headroom only when a batch outgrows it."""

    def __init__(self):
        self._buf = {}

    def take(self, name: str, like: torch.Tensor) -> torch.Tensor:
        need = like.numel() * like.element_size()
        buf = self._buf.get(name)
        if buf is None or buf.numel() < need:
            buf = torch.empty(max(need + need // 2, 4096), dtype=torch.uint8, pin_memory=True)
            self._buf[name] = buf
        return buf[:need].view(like.dtype).view(like.shape)


@dataclasses.dataclass
class SampleBatch:
    """Normalised training samples (trainer.py:74-78): `own`/`opp` are the bitboards of the side to move / the other
    side, i.e. state*player with own = +1; pi is in network order (pass last); z = winner * player in {-1,0,1}."""
    own: torch.Tensor         # int64 [S]
    opp: torch.Tensor         # int64 [S]
    pi: torch.Tensor          # float32 [S, action_size]
    z: torch.Tensor           # int8 [S]
    episode_idx: torch.Tensor  # int32 [S]
    move_idx: torch.Tensor     # int32 [S]
    n: int                     # columns (Othello: board size; Connect4: width)
    rows: int = None           # rows (None = n, a square Othello board)
    transformation: torch.Tensor = None   # uint8 [S] codes of TRANSFORMATIONS (None = all original samples)
    game: str = "othello"

    # code -> Sample.transformation tag (trainer.py:97,117): twins are appended in this order per sample
    TRANSFORMATIONS = (None, "reflection_horizontal", "rotation_90", "reflection_horizontal+rotation_90", "rotation_180",
                       "reflection_horizontal+rotation_180", "rotation_270", "reflection_horizontal+rotation_270")

    def __len__(self) -> int:
        return int(self.own.numel())

    def cpu(self, staging: "HostStaging | None" = None) -> "SampleBatch":
        """Host copy: asynchronous copies into pinned memory followed by one synchronisation.  Without `staging` every
        call takes fresh pinned tensors from torch's caching host allocator (the first calls at a new size pay for
        cudaHostAlloc: ~30 ms per 70 MB); a caller that reads batch after batch passes ONE HostStaging, whose pinned
        buffers are reused and only ever grow -- the returned batch then aliases them and is valid until the next
        cpu(staging) with the same object."""
        if not self.own.is_cuda:
            return self

        def host(name, x):
            if x is None:
                return None
            dst = torch.empty(x.shape, dtype=x.dtype, pin_memory=True) if staging is None else staging.take(name, x)
            return dst.copy_(x, non_blocking=True)

        out = SampleBatch(host("own", self.own), host("opp", self.opp), host("pi", self.pi), host("z", self.z),
                          host("episode_idx", self.episode_idx), host("move_idx", self.move_idx), self.n, self.rows,
                          host("transformation", self.transformation), self.game)
        torch.cuda.current_stream(self.own.device).synchronize()
        return out

    @property
    def n_rows(self) -> int:
        return self.n if self.rows is None else self.rows

    def augment(self, min_move_idx: int = 2) -> "SampleBatch":
        """The data augmentation of AlphaZeroTrainer.self_play (trainer.py:275-284) on the device: the originals
        followed, for every sample with move_idx >= 2, by its reflection and (Othello) the three rotations of both --
        the order of the reference's memory list.  One az_augment_plan + one az_augment_apply (csrc/sample_kernels.cu)."""
        if not self.own.is_cuda:
            raise RuntimeError("SampleBatch.augment runs on the device (CUDA kernels only, no CPU fallback)")
        if self.transformation is not None:
            raise ValueError("the batch is already augmented")
        L, dev, S = _lib.lib(), self.own.device, len(self)
        game_id = {"othello": _lib.GAME_OTHELLO, "connect4": _lib.GAME_CONNECT4, "tictactoe": _lib.GAME_TICTACTOE}[self.game]
        stream = torch.cuda.current_stream().cuda_stream
        A = self.pi.shape[1]
        scratch = torch.empty(L.az_augment_scratch_bytes(S), dtype=torch.uint8, device=dev)
        total = torch.zeros(1, dtype=torch.int64, device=dev)
        mv = self.move_idx.contiguous()
        _lib.check(L.az_augment_plan(mv.data_ptr(), S, min_move_idx, scratch.data_ptr(), total.data_ptr(), stream),
                   "az_augment_plan")
        n_out = S + int(total.item()) * L.az_augment_twins(game_id)
        out = SampleBatch(torch.empty(n_out, dtype=torch.int64, device=dev), torch.empty(n_out, dtype=torch.int64, device=dev),
                          torch.empty(n_out, A, dtype=torch.float32, device=dev), torch.empty(n_out, dtype=torch.int8, device=dev),
                          torch.empty(n_out, dtype=torch.int32, device=dev), torch.empty(n_out, dtype=torch.int32, device=dev),
                          self.n, self.rows, torch.empty(n_out, dtype=torch.uint8, device=dev), self.game)
        own, opp, pi, z, ep = (t.contiguous() for t in (self.own, self.opp, self.pi, self.z, self.episode_idx))
        _lib.check(L.az_augment_apply(own.data_ptr(), opp.data_ptr(), pi.data_ptr(), z.data_ptr(), ep.data_ptr(),
                                      mv.data_ptr(), S, game_id, self.n_rows, self.n, A, scratch.data_ptr(),
                                      out.own.data_ptr(), out.opp.data_ptr(), out.pi.data_ptr(), out.z.data_ptr(),
                                      out.episode_idx.data_ptr(), out.move_idx.data_ptr(),
                                      out.transformation.data_ptr(), stream), "az_augment_apply")
        return out

    def batch(self, index: torch.Tensor):
        """(input float32 [B, rows, cols], pi float32 [B, A], outcome float32 [B, 1]) of the samples `index` (int64
        device tensor) -- one iteration of AlphaZeroTrainer._batch_generator (trainer.py:309-320)."""
        if not self.own.is_cuda:
            raise RuntimeError("SampleBatch.batch runs on the device (CUDA kernels only, no CPU fallback)")
        dev, B, A = self.own.device, int(index.numel()), self.pi.shape[1]
        index = index.to(device=dev, dtype=torch.int64).contiguous()
        x = torch.empty(B, self.n_rows, self.n, dtype=torch.float32, device=dev)
        pi = torch.empty(B, A, dtype=torch.float32, device=dev)
        z = torch.empty(B, 1, dtype=torch.float32, device=dev)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        _lib.check(_lib.lib().az_samples_batch(self.own.data_ptr(), self.opp.data_ptr(), self.pi.data_ptr(),
                                               self.z.data_ptr(), index.data_ptr(), len(self), B, self.n_rows, self.n, A,
                                               x.data_ptr(), pi.data_ptr(), z.data_ptr(), bad.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream), "az_samples_batch")
        self._bad = bad     # checked lazily by check_batches(): no host sync inside the training loop
        return x, pi, z

    def check_batches(self) -> None:
        bad = getattr(self, "_bad", None)
        if bad is not None and int(bad.item()) != 0:
            raise IndexError("SampleBatch.batch: sample index out of range")

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.own, self.opp, self.pi, self.z, self.episode_idx, self.move_idx))

    def states(self) -> np.ndarray:
        """float64 [S, rows, cols] grids as the reference's Sample.state after normalize()."""
        own, opp = bo.bits_to_numpy(self.own), bo.bits_to_numpy(self.opp)
        rows = self.n if self.rows is None else self.rows
        out = np.zeros((len(own), rows, self.n))
        for r in range(rows):
            for c in range(self.n):
                bit = np.uint64(r * 8 + c)
                out[:, r, c] = ((own >> bit) & np.uint64(1)).astype(np.float64) - ((opp >> bit) & np.uint64(1)).astype(np.float64)
        return out


def default_tail_tiers(G: int) -> tuple:
    """Sizes of the tail tiers of a G-game batch: G/2, G/8, G/64 from 1024 games on, none below."""
    return tuple(sorted(G // d for d in (2, 8, 64))) if G >= 1024 else ()


def normalise_tail_tiers(tail_tiers, G: int) -> tuple:
    """"auto" -> default_tail_tiers(G); else the distinct sizes in (0, G), ascending."""
    if isinstance(tail_tiers, str):
        if tail_tiers != "auto":
            raise ValueError(f"tail_tiers must be 'auto' or a sequence of sizes, not {tail_tiers!r}")
        return default_tail_tiers(G)
    return tuple(sorted({int(x) for x in (tail_tiers or ()) if 0 < int(x) < G}))


def pick_tail_tier(n_live: int, current_size: int, tiers: tuple):
    """The smallest tier that holds `n_live` games and is smaller than the batch they are in, or None (stay)."""
    return next((x for x in tiers if n_live <= x < current_size), None)


class SelfPlayEngine:
    """G concurrent AlphaZero self-play games on one GPU."""

    def __init__(self, nn, G: int, n=8, n_sim: int = 200, dirichlet_alpha: float = 0.03,
                 dirichlet_epsilon: float = 0.25, temp_scheduler_type: str = "linear", temp_max_step: int = 4,
                 temp_min_step: int = 4, seed: int = 0, dtype=torch.bfloat16, capacity: int = None,
                 tie_random: bool = True, device="cuda", max_plies: int = 128, evaluator=None,
                 network_path: str = "auto", tail_tiers="auto", evaluator_factory=None, leaves: int = None):
        # `n`: Othello board size or a GameSpec (gamespec.Connect4Spec(7, 6) for BASELINE config 5)
        self._ctor = dict(nn=nn, G=G, n=n, n_sim=n_sim, dirichlet_alpha=dirichlet_alpha, dirichlet_epsilon=dirichlet_epsilon,
                          temp_scheduler_type=temp_scheduler_type, temp_max_step=temp_max_step, temp_min_step=temp_min_step,
                          seed=seed, dtype=dtype, capacity=capacity, tie_random=tie_random, device=device,
                          max_plies=max_plies, evaluator=evaluator, network_path=network_path, tail_tiers=tail_tiers,
                          evaluator_factory=evaluator_factory, leaves=leaves)
        self.pool_grows = getattr(self, "pool_grows", 0)
        self.spec = spec = as_spec(n)
        self.G, self.n, self.n_sim = G, spec.n, n_sim
        self.max_plies = max_plies
        self.tree = TreeBatch(G, n=spec, mode="neural", capacity=capacity, n_sim_hint=n_sim, device=device, seed=seed,
                              tie_random=tie_random, dirichlet_alpha=dirichlet_alpha,
                              dirichlet_epsilon=dirichlet_epsilon, enc_dtype=dtype, leaves=leaves)
        self.leaves = leaves
        # `evaluator`: a ready evaluator bound to the main tree; `evaluator_factory(tree) -> evaluator`: builds one per
        # tree batch (main and tail tiers), e.g. lambda t: HashEvaluator(salt, tree=t); default: the network's own
        self._evaluator_factory = evaluator_factory
        if evaluator is not None:
            self.evaluator = evaluator
        elif evaluator_factory is not None:
            self.evaluator = evaluator_factory(self.tree.leaf_view)
        else:
            self.evaluator = make_evaluator(nn, self.tree.leaf_view, dtype=dtype, path=network_path)
        # Tail of a batch: games of one batch end within a few plies of each other, and the last plies would run the
        # full-size leaf batch through the network for a handful of live games (Othello 8x8, 4096 games: 1481 / 285 /
        # 29 / 8 / 1 / 1 games live after plies 60..65).  `tail_tiers` = sizes of smaller TreeBatches (own pools,
        # evaluator and search graph) the live games are moved to once they fit:
        # "auto" = G/2, G/8, G/64 for G >= 1024, nothing below; () switches it off.  A caller-supplied evaluator is
        # bound to the main tree, so it switches it off too.
        self.tail_tiers = normalise_tail_tiers(tail_tiers, G) if evaluator is None else ()
        self._tiers = {}             # size -> (TreeBatch, evaluator)
        self._tier_args = dict(nn=nn, dtype=dtype, network_path=network_path, seed=seed, tie_random=tie_random,
                               dirichlet_alpha=dirichlet_alpha, dirichlet_epsilon=dirichlet_epsilon, device=device)
        self.temp = TEMP_SCHEDULERS[temp_scheduler_type](temp_max_step, temp_min_step, spec.max_moves)
        dev = self.tree.device
        T, A = max_plies, spec.action_size
        self.h_own = torch.zeros(T, G, dtype=torch.int64, device=dev)
        self.h_opp = torch.zeros(T, G, dtype=torch.int64, device=dev)
        self.h_player = torch.zeros(T, G, dtype=torch.int8, device=dev)
        self.h_live = torch.zeros(T, G, dtype=torch.bool, device=dev)
        self.h_pi = torch.zeros(T, G, A, dtype=torch.float32, device=dev)
        self.winner = torch.zeros(G, dtype=torch.int8, device=dev)
        # per-ply counters written by az_tree_move: live games after the ply, largest re-rooted tree, simulations, moves,
        # largest tree before the move; mirrored to pinned host memory one small async copy per ply
        self._counters = torch.zeros(T, 8, dtype=torch.int64, device=dev)
        self._host_counters = torch.zeros(T, 8, dtype=torch.int64).pin_memory()
        self.total_sims = 0
        self.total_moves = 0
        self.plies_played = 0
        self.max_nodes_used = 0      # high-water mark of any game's node pool half over the last play()
        for size in self.tail_tiers:      # built (pools, evaluator, search graph) up front: nothing is set up inside play()
            self._tier(size)
        self.search_events = None    # set to [] to record (start event, end event, steps, batch size) of every search of play()

    def set_network(self, nn) -> None:
        """trainer.update_network (trainer.py:383-387): refresh the device snapshot + CUDA graph."""
        self._tier_args["nn"] = nn
        for _, ev in self.trees():
            if hasattr(ev, "refresh"):
                ev.refresh(nn)

    def _tier(self, size: int):
        """(TreeBatch, evaluator) of `size` slots for the tail of a batch, built on first use."""
        if size not in self._tiers:
            a = self._tier_args
            tree = TreeBatch(size, n=self.spec, mode="neural", capacity=self.tree.capacity, n_sim_hint=self.n_sim,
                             device=a["device"], seed=a["seed"] + 0x9E3779B9 * size, tie_random=a["tie_random"],
                             dirichlet_alpha=a["dirichlet_alpha"], dirichlet_epsilon=a["dirichlet_epsilon"],
                             enc_dtype=a["dtype"], max_depth=self.tree.max_depth, graph_first_use=True,
                             leaves=self.leaves)
            if self._evaluator_factory is not None:
                ev = self._evaluator_factory(tree.leaf_view)
            else:
                ev = make_evaluator(a["nn"], tree.leaf_view, dtype=a["dtype"], path=a["network_path"])
            tree.search(self.n_sim, ev, extra_steps=0, check=False)      # captures the tier's search graph now
            self._tiers[size] = (tree, ev)
        return self._tiers[size]

    def trees(self):
        """The main TreeBatch and the tail tiers built so far (with their evaluators)."""
        return [(self.tree, self.evaluator)] + list(self._tiers.values())

    def play(self, start_own: torch.Tensor = None, start_opp: torch.Tensor = None,
             start_player: torch.Tensor = None) -> SampleBatch:
        """Plays G games to the end and returns their samples (device tensors).  A node pool that turns out too small
        (the default is sized from measurements, not from the worst case) is answered by doubling every pool and
        replaying the batch -- never by dropping simulations."""
        while True:
            try:
                return self._play(start_own, start_opp, start_player)
            except RuntimeError as e:
                if "node pool overflow" not in str(e) or self._ctor["evaluator"] is not None or self.pool_grows >= 4:
                    raise
                grows, events, nn = self.pool_grows + 1, self.search_events, self._tier_args["nn"]
                kw = dict(self._ctor, capacity=2 * self.tree.capacity, nn=nn)
                for tr, _ in self.trees():
                    tr.nodes = None                      # release the old pools before the larger ones are allocated
                self._tiers.clear()
                torch.cuda.empty_cache()
                self.pool_grows = grows
                self.__init__(**kw)
                self.search_events = events

    def _play(self, start_own: torch.Tensor = None, start_opp: torch.Tensor = None,
              start_player: torch.Tensor = None) -> SampleBatch:
        """One attempt of play().

        Per ply: one search (a CUDA graph replay) and ONE kernel for everything else (az_tree_move: history record,
        move choice + policy target, re-rooting, terminal test, retiring finished games, counters).  The host only
        needs to know how many games are still live -- to stop and to move the survivors to a smaller batch (tail
        tiers); while nearly all games of a large batch are running it reads that number one ply late from pinned
        memory, so the ply loop never waits for the device."""
        tree, G, spec = self.tree, self.G, self.spec
        L = _lib.lib()
        if start_own is None:
            tree.reset_roots()
        else:
            tree.set_roots(start_own, start_opp, start_player)
            tree.active.fill_(1)
        term, _, win = spec.terminal_score(tree.root_own, tree.root_opp, tree.root_player)
        live = term == 0
        self.winner.copy_(win)
        tree.active.copy_(live.to(torch.uint8))
        self.h_live.zero_()          # a tail tier only writes the rows of its own (live) games
        self._counters.zero_()
        h_live8 = self.h_live.view(torch.uint8)
        t = 0
        # `tree` / `evaluator` = the batch the live games are in (the main one, later a tail tier); `slot_game` = game id
        # of each of its slots (None = identity)
        evaluator, idx, slot_game = self.evaluator, None, None
        used = [tree]
        n_live = known = int(live.sum().item())
        events = []
        while t < self.max_plies and n_live > 0:
            # fresh roots only exist on the first move: afterwards the played child always carries visits
            if self.search_events is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            tree.search(self.n_sim, evaluator, extra_steps=1 if t == 0 else 0, check=False)
            if self.search_events is not None:
                e1.record()
                self.search_events.append((e0, e1, tree.last_search_steps, tree.G))
            _lib.check(L.az_tree_move(tree.desc, float(self.temp[t]), _ptr(slot_game), _ptr(self.h_own[t]), _ptr(self.h_opp[t]),
                                      _ptr(self.h_player[t]), _ptr(h_live8[t]), _ptr(self.h_pi[t]), _ptr(self.winner),
                                      _ptr(self._counters[t]), _stream()), "az_tree_move")
            self._host_counters[t].copy_(self._counters[t], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            events.append(ev)
            t += 1
            # exact count (wait for this ply's counters) in small batches, in the tail tiers and once games have begun
            # to end; one ply late -- no stall -- while more than 90 % of a large batch is still running
            lag = 1 if (tree is self.tree and G >= 1024 and known > 0.9 * G) else 0
            look = t - 1 - lag
            if look < 0:
                continue
            events[look].synchronize()
            known, n_nodes = int(self._host_counters[look, 0]), int(self._host_counters[look, 1])
            if lag:
                continue
            n_live = known
            if n_live == 0:
                break
            size = pick_tail_tier(n_live, tree.G, self.tail_tiers)
            if size is not None:
                sel = torch.nonzero(tree.active).flatten()
                sub, evaluator = self._tier(size)
                sub.adopt(tree, sel, n_nodes)
                idx = sel if idx is None else idx[sel]
                slot_game = torch.zeros(sub.G, dtype=torch.int32, device=tree.device)
                slot_game[:n_live] = idx.to(torch.int32)
                tree = sub
                used.append(tree)
        torch.cuda.current_stream().synchronize()
        for tr in used:
            tr.check_status()
        cnt = self._host_counters[:t]
        if int(cnt[t - 1, 0]) > 0 if t > 0 else n_live > 0:
            raise RuntimeError(f"self-play did not finish within max_plies={self.max_plies}")
        self.plies_played = int((cnt[:, 3] > 0).sum())
        self.total_sims = int(cnt[:, 2].sum())
        self.total_moves = int(cnt[:, 3].sum())
        self.max_nodes_used = int(cnt[:, 4].max()) if t > 0 else 0
        self.plies_recorded = t
        return self._samples(t, self.total_moves)

    def _samples(self, t: int, total: int, fused: bool = True) -> SampleBatch:
        """The history of the first t plies as a SampleBatch ordered by episode, then move (trainer.py:241-265).
        fused: one az_samples_pack launch; else the same through torch indexing (kept as the test's cross-check)."""
        spec, G, dev = self.spec, self.G, self.tree.device
        mask = self.h_live[:t]                                   # [t, G]
        if fused:
            A = spec.action_size
            lengths = mask.sum(dim=0, dtype=torch.int64)
            offset = torch.cumsum(lengths, 0) - lengths
            own = torch.empty(total, dtype=torch.int64, device=dev); opp = torch.empty_like(own)
            pi = torch.empty(total, A, dtype=torch.float32, device=dev)
            z = torch.empty(total, dtype=torch.int8, device=dev)
            ep = torch.empty(total, dtype=torch.int32, device=dev); mv = torch.empty_like(ep)
            _lib.check(_lib.lib().az_samples_pack(_ptr(self.h_own), _ptr(self.h_opp), _ptr(self.h_player),
                                                  _ptr(self.h_live.view(torch.uint8)), _ptr(self.h_pi), _ptr(self.winner),
                                                  _ptr(offset), t, G, A, _ptr(own), _ptr(opp), _ptr(pi), _ptr(z), _ptr(ep),
                                                  _ptr(mv), _stream()), "az_samples_pack")
            return SampleBatch(own, opp, pi, z, ep, mv, spec.cols, spec.rows, None, spec.game)
        move_idx, episode_idx = torch.nonzero(mask.t(), as_tuple=True)[::-1]   # ordered by episode, then move
        own = self.h_own[:t].t()[mask.t()]
        opp = self.h_opp[:t].t()[mask.t()]
        player = self.h_player[:t].t()[mask.t()]
        pi = self.h_pi[:t].transpose(0, 1)[mask.t()]
        z = (self.winner[episode_idx] * player).to(torch.int8)   # outcome * player (trainer.py:77)
        return SampleBatch(own, opp, pi, z, episode_idx.to(torch.int32), move_idx.to(torch.int32), spec.cols, spec.rows,
                           None, spec.game)
