"""Hand-written forward of the convolutional policy-value networks (OthelloNet, Connect4Net) for the leaf batch
(evaluator for TreeBatch.search).

The weights are the reference module's own (any OthelloNet / Connect4Net state_dict, alphazero/games/othello.py:341-359,
connect4.py:369-390); this file
only re-packs them for the kernels:
  * eval-mode BatchNorm folded into the preceding conv / linear layer (y = s*(Wx + b - mu) + beta, s = g/sqrt(v+eps));
  * conv2..conv4: per kernel ROW one 96x32 bf16 tile -- the three taps (dc) of the row side by side on the N side,
    N = dc*32 + c_out -- in the tcgen05 K-major no-swizzle operand layout (csrc/net_kernels.cu);
  * conv1 as one more K=16 bf16 tile (9 taps, zero padded); all conv biases in float32;
  * fc1 columns permuted from the reference's channel-major flatten (c*m*m + pos) to the trunk kernel's
    position-major feature order (pos*32 + c);
  * fc_probs and fc_value concatenated into one [n*n+2, 512] head.
Per step (default, heads="stack"):  az_net_trunk (tcgen05, features in the packed operand layout)  ->  az_net_fcstack
(fc1 + ReLU, fc2 + ReLU, heads as ONE tcgen05 cluster kernel, csrc/fcstack_kernels.cu)  ->  az_net_heads (softmax /
tanh straight into tree.probs / tree.value; inside a whole-search graph the tree step does that itself): no library
kernel on the path.  heads="library" keeps the first version (three cuBLASLt GEMMs) as a cross-check for the tests.
Input is the leaf bitboards (tree.leaf_own / leaf_opp): the float encoding of the board is never materialised.
"""
import os

import numpy as np
import torch

from . import _lib
from .boardops import _ptr, _stream

TAP_BYTES, LAYER_BYTES = 2048, 9 * 2048


def _fold_bn(weight, bias, bn):
    s = bn.weight.detach().double() / torch.sqrt(bn.running_var.detach().double() + bn.eps)
    shape = [-1] + [1] * (weight.dim() - 1)
    w = weight.detach().double() * s.reshape(shape)
    b = (bias.detach().double() - bn.running_mean.detach().double()) * s + bn.bias.detach().double()
    return w, b


def _bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.float().cpu().to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)


def pack_trunk_weights(net) -> torch.Tensor:
    """uint8 blob for az_net_trunk:
    [conv2|conv3|conv4: 3 kernel rows x [ci chunk 4][(dc, co) 96][ci%8] bf16][conv1: [k chunk 2][co 32][8] bf16, taps 0..8 then 0]
    [b1..b4 float32 [4][32]]  -- all with eval-mode BatchNorm folded in -- and, for k_net_trunk_rows, conv2..conv4 once
    more with the roles of kernel row and kernel column exchanged: 3 kernel columns x [ci chunk 4][(dr, co) 96][ci%8]."""
    assert net.n_channels == 32
    n_bytes = int(_lib.lib().az_net_wblob_bytes())
    mma = np.zeros((3, 3, 4, 96, 8), dtype=np.uint16)
    biases = np.zeros((4, 32), dtype=np.float32)
    w1, b1 = _fold_bn(net.conv1.weight, net.conv1.bias, net.bn1)     # [co][1][kh][kw]
    biases[0] = b1.float().cpu().numpy()
    c1 = np.zeros((32, 16), dtype=np.uint16)
    c1[:, :9] = _bf16_bits(w1.reshape(32, 9))
    c1 = c1.reshape(32, 2, 8).transpose(1, 0, 2)                      # [k chunk][co][k % 8]
    for li, (conv, bn) in enumerate(((net.conv2, net.bn2), (net.conv3, net.bn3), (net.conv4, net.bn4))):
        w, b = _fold_bn(conv.weight, conv.bias, bn)                   # [co][ci][kh][kw]
        biases[li + 1] = b.float().cpu().numpy()
        wb = _bf16_bits(w)
        for kh in range(3):
            for kw in range(3):
                mma[li, kh, :, kw * 32:(kw + 1) * 32] = wb[:, :, kh, kw].reshape(32, 4, 8).transpose(1, 0, 2)   # [ci chunk][co][ci % 8]
    rows = np.ascontiguousarray(mma.reshape(3, 3, 4, 3, 32, 8).transpose(0, 3, 2, 1, 4, 5))     # [layer][kw][ci chunk][kh][co][8]
    packed = np.concatenate([mma.reshape(-1).view(np.uint8), np.ascontiguousarray(c1).reshape(-1).view(np.uint8),
                             biases.reshape(-1).view(np.uint8), rows.reshape(-1).view(np.uint8)])
    assert packed.nbytes == n_bytes, (packed.nbytes, n_bytes)
    return torch.from_numpy(packed.copy())


def fcstack_plan(K1: int, N1: int, N2: int, n_out: int, NS: int, RT: int = 1) -> dict:
    """Layer shapes / blob offsets of az_net_fcstack for a K1 -> N1 -> N2 -> n_out stack on clusters of RT row tiles x NS
    column slices (the C side is the single source of truth: az_net_fcstack_plan); raises ValueError when there is no
    mapping.  "variant" names the kernel -- and with it the blob layout -- the C side will run for this shape."""
    import ctypes
    out = (ctypes.c_int64 * 15)()
    if _lib.lib().az_net_fcstack_plan(K1, N1, N2, n_out, RT, NS, out) != 0:
        raise ValueError(f"az_net_fcstack has no mapping for {K1} -> {N1} -> {N2} -> {n_out} with {RT} x {NS} CTAs per cluster")
    v = list(out)
    return {"kch": v[0:3], "Nc": v[3:6], "hpc": v[6], "w_off": v[7:10], "b_off": v[10:13], "bytes": v[13],
            "variant": v[14], "K": [K1, N1, N2], "NS": NS, "RT": RT, "n_out": n_out}


def _kmajor(w, groups: int) -> np.ndarray:
    """[N][K] weights -> bf16 bits [groups][K/64][8][N/groups][8]: per group of output columns and 64-wide K chunk the
    tcgen05 K-major no-swizzle B operand (8-wide K plane, column, K % 8), one contiguous block per (group, chunk)."""
    N, K = w.shape
    return np.ascontiguousarray(_bf16_bits(w).reshape(groups, N // groups, K // 64, 8, 8).transpose(0, 2, 3, 1, 4))


def pack_fcstack(w1, b1, w2, b2, wh, bh, NS: int, RT: int = 1):
    """(uint8 blob, plan) for az_net_fcstack from the three layers' [N][K] weights / [N] biases (BatchNorm folded).
    Variant 1: per layer [NS][K/kch][kch/8][N/NS][8] bf16 -- slice c of the output columns, K chunk, 8-wide K plane,
    column, K % 8 = the tcgen05 K-major no-swizzle B operand, one contiguous block per (slice, chunk) -- then the float32
    biases; the heads are spread over the slices hpc outputs each, zero padded to a multiple of 16 columns.
    Variant 2 (include/az_b200.h): fc1 in 2 NS pass-slices, fc2 in NS slices, the heads in NS K slices of all (padded)
    outputs; biases in natural order.  Variant 3: the same with every column group split once more into the two halves
    the CTAs of a pair stage."""
    n_out, N2 = wh.shape
    plan = fcstack_plan(w1.shape[1], w1.shape[0], w2.shape[0], n_out, NS, RT)
    blob = np.zeros(plan["bytes"], dtype=np.uint8)

    def put(off, arr):
        raw = np.ascontiguousarray(arr).reshape(-1).view(np.uint8)
        blob[off:off + raw.size] = raw

    if plan["variant"] in (2, 3):
        halves = 2 if plan["variant"] == 3 else 1      # variant 3: every column group once more in two halves (CTA pair)
        n3, nc2 = plan["Nc"][2], plan["Nc"][1]
        wh_p = torch.zeros(n3, N2, dtype=torch.float64)
        bh_p = torch.zeros(n3, dtype=torch.float64)
        wh_p[:n_out] = wh.double().cpu()
        bh_p[:n_out] = bh.double().cpu()
        put(plan["w_off"][0], _kmajor(w1, 2 * NS * halves))
        put(plan["w_off"][1], _kmajor(w2, NS * halves))
        # heads: K slice c (and column half r) of every output: [NS][halves][Nc2/8][n3/halves][8]
        put(plan["w_off"][2], _bf16_bits(wh_p).reshape(halves, n3 // halves, NS, nc2 // 8, 8).transpose(2, 0, 3, 1, 4))
        for L, b in enumerate((b1, b2, bh_p)):
            put(plan["b_off"][L], b.float().cpu().numpy().astype(np.float32))
        return torch.from_numpy(blob), plan
    hpc, n3c = plan["hpc"], plan["Nc"][2]
    wh_p = torch.zeros(NS * n3c, N2, dtype=torch.float64)
    bh_p = torch.zeros(NS * n3c, dtype=torch.float64)
    for c in range(NS):
        lo, hi = c * hpc, min(n_out, (c + 1) * hpc)
        if hi > lo:
            wh_p[c * n3c:c * n3c + hi - lo] = wh[lo:hi].double().cpu()
            bh_p[c * n3c:c * n3c + hi - lo] = bh[lo:hi].double().cpu()
    for L, (w, b) in enumerate(((w1, b1), (w2, b2), (wh_p, bh_p))):
        N, K = w.shape
        kch, nc = plan["kch"][L], plan["Nc"][L]
        assert N == NS * nc and K == plan["K"][L]
        put(plan["w_off"][L], _bf16_bits(w).reshape(NS, nc, K // kch, kch // 8, 8).transpose(0, 2, 3, 1, 4))
        put(plan["b_off"][L], b.float().cpu().numpy().astype(np.float32))
    return torch.from_numpy(blob), plan


def pack_rows(x: torch.Tensor, kch: int = 64) -> torch.Tensor:
    """Row-major [G][K] -> the packed A operand layout [ceil(G/128)][K/kch][kch/8][128][8] (rows zero padded)."""
    G, K = x.shape
    T = (G + 127) // 128
    xp = torch.zeros(T * 128, K, dtype=x.dtype, device=x.device)
    xp[:G] = x
    return xp.view(T, 128, K // kch, kch // 8, 8).permute(0, 2, 3, 1, 4).contiguous()


def unpack_rows(packed: torch.Tensor, G: int, K: int, kch: int = 64) -> torch.Tensor:
    """Inverse of pack_rows: the first G rows as a row-major [G][K] tensor."""
    T = packed.numel() // (128 * K)
    return packed.reshape(T, K // kch, kch // 8, 128, 8).permute(0, 3, 1, 2, 4).reshape(T * 128, K)[:G]


def _image_shape(nn):
    """(rows, cols) of the image the network's first convolution sees, or None if the fused trunk has no kernel for it."""
    from .games.connect4 import Connect4Net
    from .games.othello import OthelloNet
    if getattr(nn, "n_channels", None) != 32:
        return None
    if isinstance(nn, OthelloNet) and nn.n in (6, 8):
        return nn.n, nn.n
    if isinstance(nn, Connect4Net) and (nn.width, nn.height) == (7, 6):
        return nn.width, nn.height          # the (height, width) grid VIEWED as (width, height), connect4.py:399
    return None


class FusedNetEvaluator:
    """evaluator(tree): fills tree.probs / tree.value for the leaves in tree.leaf_own / tree.leaf_opp.

    Handles the reference's two convolutional architectures -- OthelloNet (n = 6, 8) and Connect4Net (7 x 6) -- which
    share the trunk (4 x conv3x3/32 + BN + ReLU, pads 1,1,0,0) and differ in the image shape and the fc sizes."""

    @staticmethod
    def supports(nn) -> bool:
        """True for the reference's architectures (layer names / shapes of othello.py:341-359, connect4.py:369-390)."""
        return _image_shape(nn) is not None

    own_kernels_per_call = 2   # k_net_trunk + k_net_heads (replayed inside the CUDA graph)

    def __init__(self, nn, tree, use_cuda_graph: bool = True, heads: str = "stack", heads_in_tree: bool = True,
                 stack_shape=None, trunk_variant: int = 0):
        """heads: "stack" (default) = az_net_fcstack, the hand-written fc1 -> fc2 -> heads kernel; "library" = the
        three cuBLASLt GEMMs of the first version (kept as a cross-check); "fused" = library fc1 / fc2 + az_net_headgemm.
        stack_shape: (RT, NS) cluster shape of az_net_fcstack, default by batch size (env AZ_B200_FCSTACK="RT,NS").
        trunk_variant: 0 = the library's choice (k_net_trunk_rows for 8 x 8 images in batches of 1024 and more, else
        k_net_trunk -- and among the former k_net_trunk_rows2, two CTAs per SM, when there are more groups of 14 boards
        than SMs); 1 = k_net_trunk whatever the batch; 2 = k_net_trunk_rows rather than k_net_trunk_rows2 (include/az_b200.h).
        heads_in_tree: inside a whole-search graph (TreeBatch.search -> enqueue()) the softmax / tanh of the heads is
        left to the tree step, which reads the bf16 logits row itself (az_tree.logits, same arithmetic and bits as
        az_net_heads); a plain call still runs az_net_heads so that tree.probs / tree.value are always valid."""
        if heads not in ("stack", "fused", "library"):
            raise ValueError("heads must be 'stack' (az_net_fcstack), 'library' (cuBLASLt GEMMs + az_net_heads) or "
                             "'fused' (library fc1 / fc2 + az_net_headgemm)")
        self.tree = tree
        self.trunk_variant = int(trunk_variant)
        self.use_cuda_graph = use_cuda_graph
        self.heads = heads
        self.stack_shape = stack_shape
        self.heads_in_tree = bool(heads_in_tree) and heads != "fused"
        self.own_launches = 0
        self.graph = None
        self.n_calls = 0
        self.version = 0
        self.refresh(nn)

    def refresh(self, nn) -> None:
        """(Re)packs the weights; call after the network changed (MCT.nn setter, mcts.py:78-83)."""
        self.version += 1            # TreeBatch drops whole-search graphs captured over the old buffers
        t = self.tree
        dev, G = t.device, t.G
        shape = _image_shape(nn)
        if shape is None:
            raise ValueError(f"the fused trunk has no kernel for {nn}")
        self.rows, self.cols = shape
        self.connect4 = t.spec.game == "connect4"
        # k_net_trunk + k_net_heads (+ k_net_fcstack) (+ az_connect4_image_bits)
        self.own_kernels_per_call = 2 + (1 if self.heads == "stack" else 0) + (1 if self.connect4 else 0)
        if nn.action_size != t.A or (not self.connect4 and nn.n != t.n) or \
                (self.connect4 and (nn.width, nn.height) != (t.spec.cols, t.spec.rows)):
            raise ValueError(f"network {nn} does not match the tree's game {t.spec}")
        self.src = nn
        m = (self.rows - 4) * (self.cols - 4)
        h1, h2 = nn.fc1.out_features, nn.fc2.out_features
        self.wblob = pack_trunk_weights(nn).to(dev)
        w1, b1 = _fold_bn(nn.fc1.weight, nn.fc1.bias, nn.fc_bn1)               # [h1][32*m] channel-major
        w1 = w1.reshape(h1, 32, m).permute(0, 2, 1).reshape(h1, m * 32)
        w2, b2 = _fold_bn(nn.fc2.weight, nn.fc2.bias, nn.fc_bn2)
        bf = lambda x: x.float().to(dev).to(torch.bfloat16).contiguous()
        self.w1t, self.b1 = bf(w1.t()), bf(b1)                                  # [K][N] for addmm
        self.w2t, self.b2 = bf(w2.t()), bf(b2)
        A = self.A = nn.action_size
        self.ld = ((A + 1 + 7) // 8) * 8                                        # head rows padded to a multiple of 8
        wh = torch.zeros(self.ld, h2, dtype=torch.float64)
        bh = torch.zeros(self.ld, dtype=torch.float64)
        wh[:A] = nn.fc_probs.weight.detach().double().cpu(); bh[:A] = nn.fc_probs.bias.detach().double().cpu()
        wh[A] = nn.fc_value.weight.detach().double().cpu()[0]; bh[A] = nn.fc_value.bias.detach().double().cpu()[0]
        self.wht, self.bh = bf(wh.t()), bf(bh)                                  # library-GEMM form (heads="library")
        # fused heads kernel (az_net_headgemm): bf16 [h2/8][NH][8] K-major tiles + float bias[NH]
        self.NH = ((A + 1 + 15) // 16) * 16
        tiles = np.zeros((h2 // 8, self.NH, 8), dtype=np.uint16)
        tiles[:, :A + 1, :] = _bf16_bits(wh[:A + 1].float()).reshape(A + 1, h2 // 8, 8).transpose(1, 0, 2)
        bias = np.zeros(self.NH, dtype=np.float32)
        bias[:A + 1] = bh[:A + 1].float().numpy()
        blob = np.concatenate([tiles.reshape(-1).view(np.uint8), bias.view(np.uint8)])
        assert blob.nbytes == int(_lib.lib().az_net_headgemm_blob_bytes(h2, self.NH))
        self.head_blob = torch.from_numpy(blob.copy()).to(dev)
        self.h2_features = h2
        self.K1, self.N1, self.N2 = m * 32, h1, h2
        if self.heads == "stack":
            self.RT, self.NS = self._stack_shape(G, h1)
            blob, self.plan = pack_fcstack(w1, b1, w2, b2, wh[:A + 1], bh[:A + 1], self.NS, self.RT)
            self.stack_blob = blob.to(dev)
            tiles = -(-((G + 127) // 128) // self.RT) * self.RT          # row tiles, a whole number of clusters
            self.tiles = tiles
            self.features = torch.zeros(tiles * 128 * m * 32, dtype=torch.bfloat16, device=dev)      # packed layout
            self.h1 = torch.zeros(tiles * 128 * h1, dtype=torch.bfloat16, device=dev)
            self.h2 = torch.zeros(tiles * 128 * h2, dtype=torch.bfloat16, device=dev)
            self.logits = torch.zeros(G, self.ld, dtype=torch.bfloat16, device=dev)
        else:
            self.features = torch.zeros(G, m * 32, dtype=torch.bfloat16, device=dev)
            self.h1 = torch.empty(G, h1, dtype=torch.bfloat16, device=dev)
            self.h2 = torch.empty(G, h2, dtype=torch.bfloat16, device=dev)
            self.logits = torch.empty(G, self.ld, dtype=torch.bfloat16, device=dev)
        # (pointer, row pitch) TreeBatch.search puts into az_tree.logits for this evaluator, or None
        self.logits_binding = (self.logits.data_ptr(), self.ld) if self.heads_in_tree else None
        if self.connect4:
            self.img_own = torch.zeros(G, dtype=torch.int64, device=dev)
            self.img_opp = torch.zeros(G, dtype=torch.int64, device=dev)
        self.graph = None
        if self.use_cuda_graph:
            self._capture()

    def _stack_shape(self, G: int, N1: int):
        """(RT, NS) of the az_net_fcstack clusters: NS column slices so that a slice of fc1 fits one accumulator
        (N1/NS <= 256), one row tile per cluster.  Clusters of 4 all fit the GPU at once (32 of them at 4096 positions);
        clusters of 8 (RT = 2, the weights multicast over two row tiles) do not -- the 16th waits for a second wave,
        which doubles the kernel's latency (measured: 37 us against 19.5 us, profiles/r2_fcstack_experiments.txt)."""
        shape = self.stack_shape
        if shape is None and os.environ.get("AZ_B200_FCSTACK"):
            shape = tuple(int(v) for v in os.environ["AZ_B200_FCSTACK"].split(","))
        if shape is not None:
            return int(shape[0]), int(shape[1])
        NS = max(1, N1 // 256)
        RT = 1
        return RT, NS

    def features_rowmajor(self) -> torch.Tensor:
        """bf16 [G][(rows-4)(cols-4)*32] view / copy of the trunk's output whatever layout it was written in."""
        if self.heads == "stack":
            return unpack_rows(self.features, self.tree.G, self.K1)
        return self.features

    def trunk(self, debug=None, debug_layer: int = 0) -> None:
        t = self.tree
        own, opp = t.leaf_own, t.leaf_opp
        if self.connect4:
            _lib.check(_lib.lib().az_connect4_image_bits(_ptr(own), _ptr(opp), t.spec.cols, t.spec.rows, t.G,
                                                         _ptr(self.img_own), _ptr(self.img_opp), _stream()),
                       "az_connect4_image_bits")
            own, opp = self.img_own, self.img_opp
        _lib.check(_lib.lib().az_net_trunk(_ptr(own), _ptr(opp), self.rows, self.cols, t.G, _ptr(self.wblob),
                                           _ptr(self.features),
                                           (1 if self.heads == "stack" else 0) | (2 if self.trunk_variant == 1 else 0) | (4 if self.trunk_variant == 2 else 0), _ptr(debug),
                                           int(debug_layer), _stream()), "az_net_trunk")

    def _forward(self, output_transform: bool = True) -> None:
        t = self.tree
        self.trunk()
        if self.heads == "stack":
            _lib.check(_lib.lib().az_net_fcstack(_ptr(self.features), _ptr(self.stack_blob), _ptr(self.h1), _ptr(self.h2),
                                                 _ptr(self.logits), self.ld, self.A + 1, t.G, self.K1, self.N1, self.N2,
                                                 self.RT, self.NS, _stream()), "az_net_fcstack")
        else:
            with torch.no_grad():
                torch._addmm_activation(self.b1, self.features, self.w1t, out=self.h1)      # fused bias + ReLU epilogue
                torch._addmm_activation(self.b2, self.h1, self.w2t, out=self.h2)
                if self.heads == "library":
                    torch.addmm(self.bh, self.h2, self.wht, out=self.logits)
        if not output_transform:
            return
        if self.heads != "fused":
            _lib.check(_lib.lib().az_net_heads(_ptr(self.logits), self.ld, self.A, t.G, _ptr(t.probs), _ptr(t.value),
                                               _stream()), "az_net_heads")
        else:       # fc_probs / fc_value + softmax / tanh in one tcgen05 kernel
            _lib.check(_lib.lib().az_net_headgemm(_ptr(self.h2), self.h2_features, _ptr(self.head_blob), self.NH, self.A,
                                                  t.G, _ptr(t.probs), _ptr(t.value), _stream()), "az_net_headgemm")

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

    @property
    def own_kernels_per_enqueue(self) -> int:
        return self.own_kernels_per_call - (1 if self.logits_binding is not None else 0)

    def enqueue(self) -> None:
        """The forward's kernels on the current stream, no graph replay: what TreeBatch captures into a search graph
        (without az_net_heads when the tree step reads the logits itself)."""
        self._forward(output_transform=self.logits_binding is None)

    def __call__(self, tree=None) -> None:
        if self.graph is not None:
            self.graph.replay()
        else:
            self._forward()
        self.n_calls += 1
        self.own_launches += self.own_kernels_per_call


FusedOthelloEvaluator = FusedNetEvaluator   # name used before Connect4Net was covered


class TicTacToeNetEvaluator:
    """evaluator(tree) for the reference's TicTacToeNet (tictactoe.py:262-316): the whole forward -- 9 -> 9 -> 9 ->
    (9 + 1) units with the eval-mode BatchNorm1d layers folded in, softmax, tanh -- is one fp32 kernel reading the leaf
    bitboards (az_tictactoe_net), one thread per position."""

    own_kernels_per_call = 1

    @staticmethod
    def supports(nn) -> bool:
        from .games.tictactoe import TicTacToeNet
        return isinstance(nn, TicTacToeNet)

    def __init__(self, nn, tree, use_cuda_graph: bool = False):
        if tree.spec.game != "tictactoe":
            raise ValueError(f"network {nn} does not match the tree's game {tree.spec}")
        self.tree = tree
        self.n_calls = 0
        self.version = 0
        self.refresh(nn)

    def refresh(self, nn) -> None:
        self.version += 1
        self.src = nn
        w1, b1 = _fold_bn(nn.fc1.weight, nn.fc1.bias, nn.bn1)
        w2, b2 = _fold_bn(nn.fc2.weight, nn.fc2.bias, nn.bn2)
        parts = [w1, b1, w2, b2, nn.fc_probs.weight.detach().double(), nn.fc_probs.bias.detach().double(),
                 nn.fc_value.weight.detach().double(), nn.fc_value.bias.detach().double()]
        flat = torch.cat([p.reshape(-1).cpu() for p in parts]).float()
        assert flat.numel() == int(_lib.lib().az_tictactoe_net_floats())
        self.weights = flat.to(self.tree.device).contiguous()

    def enqueue(self) -> None:
        t = self.tree
        _lib.check(_lib.lib().az_tictactoe_net(_ptr(t.leaf_own), _ptr(t.leaf_opp), t.G, _ptr(self.weights),
                                               _ptr(t.probs), _ptr(t.value), _stream()), "az_tictactoe_net")

    def __call__(self, tree=None) -> None:
        self.enqueue()
        self.n_calls += 1


def make_evaluator(nn, tree, dtype=torch.bfloat16, path: str = "auto"):
    """The leaf evaluator for `nn` on `tree`: the hand-written kernels when there is one for the architecture
    (OthelloNet 6/8, Connect4Net 7x6: tcgen05 trunk; TicTacToeNet: fused MLP) -- these read the leaf bitboards, so the
    select kernel's encoding is switched off -- else the module's own forward under a CUDA graph.  path: "auto" |
    "fused" (raise if there is no kernel) | "torch"."""
    from .evaluators import NetEvaluator
    if path not in ("auto", "fused", "torch"):
        raise ValueError(f"unknown network path {path!r}")
    if path != "torch":
        for cls in (FusedNetEvaluator, TicTacToeNetEvaluator):
            if cls.supports(nn):
                # AZ_B200_HEADS=library: the first version's three cuBLASLt GEMMs instead of az_net_fcstack (A/B measurements)
                kw = {"heads": os.environ["AZ_B200_HEADS"]} if cls is FusedNetEvaluator and os.environ.get("AZ_B200_HEADS") else {}
                ev = cls(nn, tree, **kw)
                tree.set_encoding(False)
                return ev
        if path == "fused":
            raise ValueError(f"there is no hand-written forward kernel for {nn}")
    return NetEvaluator(nn, tree, dtype=dtype)
