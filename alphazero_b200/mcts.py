"""Drop-in `Node` / `MCT` (alphazero/mcts.py:8-269) on top of the GPU tree kernels.

`MCT` keeps the reference's constructor, attributes and methods, but the tree lives in a one-game `TreeBatch`
(device node pool, az_tree_* kernels).  `root` is a read-only snapshot view: `root.children[move].N / .Q / .P` are
downloaded on access, which is what the reference's callers (get_action_probs, arena display, tests) read.
For throughput use `TreeBatch` / `SelfPlayEngine` directly: this class runs one game per object.
"""
from time import time

import numpy as np
import torch

from . import boardops as bo
from .base import Board, PolicyValueNetwork, TreeEval
from .engine import TreeBatch
from .fused_net import make_evaluator
from .gamespec import OthelloSpec, spec_of
from .utils import fair_max


def action_to_move(a: int, n: int):
    return (n, n) if a == bo.PASS else (a // 8, a % 8)


def move_to_action(move, n: int) -> int:
    if tuple(move) == (n, n):
        return bo.PASS
    r, c = int(move[0]), int(move[1])
    return r * 8 + c if 0 <= r < n and 0 <= c < n else 255


class Node:
    """Snapshot of one tree node as the reference exposes it (mcts.py:8-46): move, N, Q, P, children."""

    def __init__(self, move=None, parent=None, prob: float = None, N: int = 0, Q: float = 0) -> None:
        self.move = move
        self.parent = parent
        self.N = N
        self.Q = Q
        self.P = prob
        self.children = {}
        self.has_dirichlet_noise = False

    def __str__(self) -> str:
        return f"Node: N={self.N} | Q={self.Q} | P={self.P} | n_children={len(self.children)} | move={self.move}"

    def UCT(self, c_exploration: float = np.sqrt(2)) -> float:
        if self.N == 0:
            return float("inf")
        return self.Q + c_exploration * np.sqrt(np.log(self.parent.N) / self.N)

    def PUCT(self, c_puct: float = 1.0) -> float:
        return self.Q + c_puct * self.P * np.sqrt(self.parent.N) / (1 + self.N)


class MCT:
    """Monte-Carlo tree of one game; `eval_method` "rollout" (UCT + random playouts) or "neural" (PUCT + network)."""

    CAPACITY = 1 << 16   # nodes per pool half of the one-game tree
    TIE_RANDOM = True    # fair_max semantics; tests set False for the deterministic lowest-action-index tie-break

    def __init__(self, eval_method: str = None, nn: PolicyValueNetwork = None, dirichlet_alpha: float = None,
                 dirichlet_epsilon: float = None, leaves: int = None) -> None:
        """`leaves` (an extension, not in the reference): None = the reference's one leaf per simulation; L >= 1 =
        throughput mode with up to L leaves in flight per step under virtual loss (engine.TreeBatch)."""
        self.leaves = leaves
        self.n_rollouts = 0
        self.simulation_time = 0
        self.eval_method = TreeEval.to_dict()["rollout" if eval_method is None else eval_method]
        self._nn = nn
        self._dirichlet_alpha = dirichlet_alpha
        self._dirichlet_epsilon = dirichlet_epsilon
        if self._nn is not None and self.eval_method != TreeEval.NEURAL:
            raise ValueError(f"A neural network has been set for the MCT but the evaluation method is {self.eval_method}")
        self._tree = None
        self._evaluator = None
        self._pending = []      # change_root calls made before the first search (no device tree yet)
        self._n = None
        self._spec = None
        self._nn_stamp = None   # weights fingerprint the evaluator's bf16 snapshot was taken at

    # ---- pickling (the reference ships players / MCTs into a multiprocessing.Pool, arena.py:211-234) -------
    def __getstate__(self) -> dict:
        """The device tree (node pools, CUDA graphs) does not travel: a pickled MCT keeps its configuration and its
        network and starts from a fresh root on the other side, like a player after reset() (players.py:151-152).
        Arena workers reset their players before every game (arena.py:56-58), so nothing is lost there."""
        state = dict(self.__dict__)
        state["_tree"] = state["_evaluator"] = state["_nn_stamp"] = None
        state["_pending"] = []
        return state

    def __setstate__(self, state: dict) -> None:
        self.__dict__.update(state)

    @staticmethod
    def _weights_stamp(nn):
        """Changes whenever a parameter / buffer of `nn` is written in place (load_state_dict, optimizer.step()): the
        reference evaluates the live module (base.py:357-367), so the snapshot must follow it."""
        if nn is None:
            return None
        if not hasattr(nn, "parameters"):      # a test double that supplies its own evaluator
            return (id(nn), 0)
        return (id(nn), sum(int(t._version) for t in list(nn.parameters()) + list(nn.buffers())))

    # ---- attributes the reference's callers set directly (players.py:213-215, trainer.py:387) -------------
    @property
    def nn(self):
        return self._nn

    @nn.setter
    def nn(self, nn: PolicyValueNetwork) -> None:
        if self.eval_method != TreeEval.NEURAL:
            raise ValueError(f"Trying to set a neural network for the MCT but the evaluation method is {self.eval_method}")
        self._nn = nn
        self._evaluator = None          # re-snapshot the weights / re-capture the CUDA graph on the next search
        self._nn_stamp = None

    @property
    def dirichlet_alpha(self):
        return self._dirichlet_alpha

    @dirichlet_alpha.setter
    def dirichlet_alpha(self, v):
        self._dirichlet_alpha = v
        if self._tree is not None:
            self._tree.set_dirichlet(self._dirichlet_alpha, self._dirichlet_epsilon)

    @property
    def dirichlet_epsilon(self):
        return self._dirichlet_epsilon

    @dirichlet_epsilon.setter
    def dirichlet_epsilon(self, v):
        self._dirichlet_epsilon = v
        if self._tree is not None:
            self._tree.set_dirichlet(self._dirichlet_alpha, self._dirichlet_epsilon)

    # ---- device tree --------------------------------------------------------------------------------------
    def _ensure_tree(self, board: Board) -> None:
        spec = spec_of(board)    # NotImplementedError for games without device rules
        if self._tree is None or self._spec != spec:
            mode = "neural" if self.eval_method == TreeEval.NEURAL else "rollout"
            self._tree = TreeBatch(1, n=spec, mode=mode, capacity=self.CAPACITY, seed=int(np.random.randint(1 << 31)),
                                   tie_random=self.TIE_RANDOM, dirichlet_alpha=self._dirichlet_alpha,
                                   dirichlet_epsilon=self._dirichlet_epsilon, enc_dtype=torch.bfloat16,
                                   leaves=self.leaves if mode == "neural" else None)
            self._spec, self._n = spec, spec.n
            self._evaluator = None
            self._pending = []
        own, opp = (board.black, board.white) if board.player == 1 else (board.white, board.black)
        t = self._tree
        same = (bo.to_u64(int(t.root_own.item())) == own and bo.to_u64(int(t.root_opp.item())) == opp
                and int(t.root_player.item()) == board.player)
        if not same:   # the caller moved the board without change_root: restart from a fresh root (mcts.py:124-125)
            dev = t.device
            t.set_roots(torch.tensor([bo.to_i64(own)], dtype=torch.int64, device=dev),
                        torch.tensor([bo.to_i64(opp)], dtype=torch.int64, device=dev),
                        torch.tensor([board.player], dtype=torch.int8, device=dev), fresh_trees=True)
        if self.eval_method == TreeEval.NEURAL and self._evaluator is None:
            if self._nn is None:
                raise ValueError("MCT in neural mode needs a network before search")
            if hasattr(self._nn, "make_batched_evaluator"):      # evaluator supplied by the network object itself
                self._evaluator = self._nn.make_batched_evaluator(t.leaf_view)
            else:
                t.set_encoding(True)
                self._evaluator = make_evaluator(self._nn, t.leaf_view)   # switches the encoding off when it reads bitboards
            self._nn_stamp = self._weights_stamp(self._nn)
        elif self.eval_method == TreeEval.NEURAL and self._nn_stamp != self._weights_stamp(self._nn):
            if hasattr(self._evaluator, "refresh"):               # weights changed in place since the snapshot
                self._evaluator.refresh(self._nn)
            self._nn_stamp = self._weights_stamp(self._nn)

    def _adopt(self, old: "MCT") -> None:
        """Re-uses the device pools / CUDA graph of a discarded MCT of the same kind (player.reset() builds a new MCT
        per game, players.py:151-152, 228-234); the tree itself starts fresh."""
        if old is None or old._tree is None or old.eval_method != self.eval_method or old.leaves != self.leaves:
            return
        self._tree, self._n, self._spec = old._tree, old._n, old._spec
        self._tree.set_dirichlet(self._dirichlet_alpha, self._dirichlet_epsilon)
        if old._nn is self._nn:
            self._evaluator, self._nn_stamp = old._evaluator, old._nn_stamp     # (_ensure_tree refreshes a stale one)
        self._tree.reset_trees()
        self._tree.root_own.fill_(-1)      # no position yet: the first search uploads the board

    # ---- reference API --------------------------------------------------------------------------------------
    def get_stats(self):
        return self.n_rollouts, self.simulation_time

    @property
    def root(self) -> Node:
        """Snapshot of the root and its children (insertion order = ascending action index)."""
        node = Node(move=None, parent=None)
        if self._tree is None:
            return node
        N, Q, P, rn = self._tree.root_stats()
        N, Q, P = N[0].cpu().numpy(), Q[0].cpu().numpy(), P[0].cpu().numpy()
        node.N = int(rn.item())
        neural = self.eval_method == TreeEval.NEURAL
        for a in range(65):
            if N[a] >= 0:
                mv = self._spec.action_to_move(a)
                node.children[mv] = Node(move=mv, parent=node, prob=np.float32(P[a]) if neural else None,
                                         N=int(N[a]), Q=float(Q[a]))
        return node

    def get_prior_probs(self) -> dict:
        return {move: child.P for move, child in self.root.children.items()}

    def get_action_probs(self, board: Board, temp: float = 0):
        """mcts.py:95-116: one-hot on the most visited move (random tie-break) for temp == 0, else N^(1/temp)
        normalised; both with the visit counts."""
        children = self.root.children
        if len(children) == 0:
            return {board.pass_move: 1.}
        visit_counts = {move: int(child.N) for move, child in children.items()}
        if temp == 0:
            best_move, _ = fair_max(children.items(), key=lambda x: x[1].N)
            return {best_move: 1}, visit_counts
        values = {move: child.N ** (1. / temp) for move, child in children.items()}
        total = sum(values.values())
        return {move: v / total for move, v in values.items()}, visit_counts

    def change_root(self, move) -> None:
        """mcts.py:118-125: keep the subtree of `move` (played on the device root board as well)."""
        if self._tree is None:
            return                      # nothing searched yet: the next search starts from the board it is given
        a = self._spec.move_to_action(move)
        self._tree.advance(torch.tensor([a], dtype=torch.uint8))
        st = int(self._tree.status.item())
        if st & 4:
            # the move does not belong to the tree's position (caller desynchronised): drop the tree
            self._tree.status.zero_()
            self._tree.root_own.fill_(-1)   # forces a fresh root on the next search

    def search(self, board: Board, n_sim: int = None, compute_time: float = None) -> None:
        """mcts.py:254-269."""
        start = time()
        if n_sim is None and compute_time is None:
            raise ValueError("MCT.search needs to have either n_sim or compute_time specified.")
        self._ensure_tree(board)
        self.n_rollouts = 0
        if n_sim is not None:
            self._tree.search(int(n_sim), self._evaluator)
            self.n_rollouts = int(n_sim)
        else:
            chunk = 32
            while time() - start < compute_time:
                self._tree.search(chunk, self._evaluator)
                self.n_rollouts += chunk
        self.simulation_time = time() - start
