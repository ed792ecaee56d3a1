"""Leaf evaluators for TreeBatch.search (the batched PolicyValueNetwork.evaluate, alphazero/base.py:357-367)."""
import torch

from . import _lib
from .boardops import _ptr, _stream, to_u64


class HashEvaluator:
    """Deterministic test double (see tests/hash_eval.py): priors and values are a hash of the leaf position.
    Used for visit-count identity tests against the reference MCT; runs on the device (az_eval_hash)."""

    def __init__(self, salt: int = 0, tree=None):
        self.salt = to_u64(salt)
        self.n_calls = 0
        self.tree = tree          # bound to one tree: TreeBatch.search may capture it into a whole-search graph
        self.version = 0

    def __call__(self, tree) -> None:
        _lib.check(_lib.lib().az_eval_hash(_ptr(tree.leaf_own), _ptr(tree.leaf_opp), tree.A, tree.G, self.salt,
                                           _ptr(tree.probs), _ptr(tree.value), _stream()), "az_eval_hash")
        self.n_calls += 1

    @property
    def enqueue(self):
        if self.tree is None:     # not bound to one tree: TreeBatch.search runs it from the host loop
            return None
        t = self.tree
        return lambda: _lib.check(_lib.lib().az_eval_hash(_ptr(t.leaf_own), _ptr(t.leaf_opp), t.A, t.G, self.salt,
                                                          _ptr(t.probs), _ptr(t.value), _stream()), "az_eval_hash")


class NetEvaluator:
    """Runs a PolicyValueNetwork on the leaf encodings written by the select kernel, zero-copy:

    tree.enc [G, n, n] is the static input, tree.probs / tree.value the static outputs.  The forward is the
    network's own PyTorch module (weights cast to `dtype`, eval mode) captured once in a CUDA graph and replayed
    every simulation step; exp(log_softmax) is taken in float32 (PolicyValueNetwork.predict, base.py:350-355).
    Re-create (or call `refresh()`) after the weights change (mcts.py:78-83 nn setter)."""

    own_kernels_per_call = 0   # library kernels only

    def __init__(self, nn, tree, dtype=torch.bfloat16, use_cuda_graph: bool = True):
        self.tree = tree
        self.dtype = dtype
        self.use_cuda_graph = use_cuda_graph
        self.graph = None
        self.n_calls = 0
        self.version = 0
        self.refresh(nn)

    def refresh(self, nn) -> None:
        import copy
        self.version += 1            # TreeBatch drops whole-search graphs captured over the old weights
        self.src = nn
        self.net = copy.deepcopy(nn).to(device=self.tree.device, dtype=self.dtype).eval()
        for p in self.net.parameters():
            p.requires_grad_(False)
        if hasattr(self.net, "device"):
            self.net.device = self.tree.device
        self.graph = None
        if self.use_cuda_graph:
            self._capture()

    def _forward(self) -> None:
        t = self.tree
        x = t.enc if t.enc.dtype == self.dtype else t.enc.to(self.dtype)
        with torch.no_grad():
            log_probs, v = self.net.forward(x)
            torch.exp(log_probs.float(), out=t.probs)
            t.value.copy_(v.float().reshape(-1))

    def _capture(self) -> None:
        s = torch.cuda.Stream(device=self.tree.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                self._forward()
        torch.cuda.current_stream().wait_stream(s)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._forward()

    def enqueue(self) -> None:
        """The forward's kernels on the current stream, no graph replay: what TreeBatch captures into a search graph."""
        self._forward()

    def __call__(self, tree=None) -> None:
        if self.graph is not None:
            self.graph.replay()
        else:
            self._forward()
        self.n_calls += 1
