"""GPU: the hand-written network forward (tcgen05 conv trunk + library GEMMs + fused heads) against the fp32
PyTorch forward of the same module: every conv layer (debug dumps) and the final policy/value within 1e-2."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from test_gpu_selfplay import _random_positions

pytestmark = pytest.mark.gpu


def _net(n, seed=0):
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(seed)
    net = OthelloNet(n=n)
    net.train()
    with torch.no_grad():
        for _ in range(5):   # non-trivial BatchNorm running statistics, as a trained net has
            net(torch.randint(-1, 2, (256, n, n)).float())
    net.eval()
    return net


def _setup(n, G_target=300, seed=1):
    from alphazero_b200 import boardops as bo
    from alphazero_b200.engine import TreeBatch
    black, white, player = _random_positions(n, G_target, seed)
    G = len(black)
    own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
    tree = TreeBatch(G, n=n, mode="neural")
    tree.leaf_own.copy_(bo.bits_from_numpy(own)); tree.leaf_opp.copy_(bo.bits_from_numpy(opp))
    x = bo.encode(tree.leaf_own, tree.leaf_opp, n).cpu()
    return tree, x, G


@pytest.mark.parametrize("n", [8, 6])
def test_trunk_layers_match_fp32(n):
    from alphazero_b200.fused_net import FusedOthelloEvaluator
    net = _net(n)
    tree, x, G = _setup(n)
    ev = FusedOthelloEvaluator(net, tree, use_cuda_graph=False)
    with torch.no_grad():
        a1 = F.relu(net.bn1(net.conv1(x.view(-1, 1, n, n))))
        a2 = F.relu(net.bn2(net.conv2(a1)))
        a3 = F.relu(net.bn3(net.conv3(a2)))
        a4 = F.relu(net.bn4(net.conv4(a3)))
    rpb = (n + 1) * 8                      # rows per board in the kernel's row space: ring row + n rows of 8 cells
    from alphazero_b200 import _lib
    NB = _lib.lib().az_net_trunk_boards_per_group(n, G)
    assert NB in ((512 - 8) // rpb, (256 - 8) // rpb)
    n_groups = (G + NB - 1) // NB
    for layer, ref, off in ((1, a1, 0), (2, a2, 0), (3, a3, 1)):
        dbg = torch.zeros(n_groups, 768, 32, dtype=torch.bfloat16, device="cuda")
        ev.trunk(debug=dbg, debug_layer=layer)
        torch.cuda.synchronize()
        d = dbg.float().cpu()
        m = ref.shape[-1]
        got = torch.zeros_like(ref)
        for g in range(G):
            grp, b = divmod(g, NB)
            rows = d[grp, b * rpb + 8:(b + 1) * rpb].reshape(n, 8, 32)
            got[g] = rows[off:off + m, off:off + m].permute(2, 0, 1)
        scale = ref.abs().max().item()
        err = (got - ref).abs().max().item()
        assert err < 0.02 * scale + 0.02, (n, layer, err, scale)
        if layer == 1:   # the zero ring rows / pad cells that pad conv2
            blk = d[0, :2 * rpb].reshape(2, n + 1, 8, 32)
            assert blk[:, 0].abs().max() == 0 and blk[:, :, n:].abs().sum() == 0
    ev.trunk()
    torch.cuda.synchronize()
    m = n - 4
    feat = ev.features_rowmajor().float().cpu().reshape(G, m * m, 32).permute(0, 2, 1).reshape(G, 32, m, m)
    assert (feat - a4).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02


@pytest.mark.parametrize("variant,G", [(0, 1030), (2, 2200), (0, 2200), (0, 4300)])
def test_trunk_rows_layers_match_fp32(variant, G):
    """k_net_trunk_rows / k_net_trunk_rows2 (8 x 8 boards, batches of 1024 and more: tile = one board row of 14 boards, one
    accumulator per output tile): every conv layer through its debug dump, the zero pad lanes, the features, and the
    policy / value against fp32; and against k_net_trunk on the same batch."""
    from alphazero_b200 import _lib
    from alphazero_b200.fused_net import FusedOthelloEvaluator
    n = 8
    net = _net(n, seed=11)
    tree, x, G0 = _setup(n, 400, seed=13)
    # 1030 boards: 73 full groups of 14 and one of 8 (k_net_trunk_rows); 2200: more groups than SMs = k_net_trunk_rows2
    # (two CTAs per SM; variant 2 keeps k_net_trunk_rows), the last group of 2; 4300: CTAs with more than one group.
    # The positions repeat, which also checks that a board's result does not depend on where it sits in a group.
    from alphazero_b200.engine import TreeBatch
    idx = torch.arange(G) % G0
    big = TreeBatch(G, n=n, mode="neural", capacity=64)
    big.leaf_own.copy_(tree.leaf_own[idx.cuda()]); big.leaf_opp.copy_(tree.leaf_opp[idx.cuda()])
    ev = FusedOthelloEvaluator(net, big, use_cuda_graph=False, trunk_variant=variant)
    assert _lib.lib().az_net_trunk_boards_per_group(8, G) == 14
    with torch.no_grad():
        a1 = F.relu(net.bn1(net.conv1(x.view(-1, 1, n, n))))
        a2 = F.relu(net.bn2(net.conv2(a1)))
        a3 = F.relu(net.bn3(net.conv3(a2)))
        a4 = F.relu(net.bn4(net.conv4(a3)))
    NB, n_groups = 14, (G + 13) // 14
    lanes = (torch.arange(14)[:, None] * 9 + torch.arange(8)[None, :]).reshape(-1)          # lane of (board, column)
    for layer, ref, off in ((1, a1, 0), (2, a2, 0), (3, a3, 1)):
        dbg = torch.zeros(n_groups, 1024, 32, dtype=torch.bfloat16, device="cuda")
        ev.trunk(debug=dbg, debug_layer=layer)
        torch.cuda.synchronize()
        d = dbg.float().cpu().reshape(n_groups, 8, 128, 32)
        cells = d[:, :, lanes].reshape(n_groups, 8, 14, 8, 32).permute(0, 2, 4, 1, 3).reshape(n_groups * 14, 32, 8, 8)[:G]
        m = ref.shape[-1]
        got = cells[:, :, off:off + m, off:off + m]
        scale = ref.abs().max().item()
        err = (got - ref[idx]).abs().max().item()
        assert err < 0.02 * scale + 0.02, (layer, err, scale)
        if layer == 1:    # conv2's zero padding: the pad column, the two unused lanes and the boards a partial group lacks
            pad = torch.ones(128, dtype=torch.bool)
            pad[lanes] = False
            assert d[:, :, pad].abs().max() == 0
            assert d[-1, :, (G - (n_groups - 1) * 14) * 9:].abs().max() == 0
    ev.trunk()
    torch.cuda.synchronize()
    feat = ev.features_rowmajor().float().cpu().reshape(G, 16, 32).permute(0, 2, 1).reshape(G, 32, 4, 4)
    assert (feat - a4[idx]).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02
    assert (feat[G0:2 * G0] == feat[:G0]).all()
    # the first kernel on the same batch accumulates the same products in another order: both within the layer tolerance,
    # and the whole forward within the north star's 1e-2 of fp32
    ev1 = FusedOthelloEvaluator(net, big, use_cuda_graph=False, trunk_variant=1)
    ev1.trunk()
    torch.cuda.synchronize()
    feat1 = ev1.features_rowmajor().float().cpu().reshape(G, 16, 32).permute(0, 2, 1).reshape(G, 32, 4, 4)
    assert (feat - feat1).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02
    if variant == 0 and G > 148 * 14:
        # k_net_trunk_rows2 feeds every accumulator the same MMAs in the same order as k_net_trunk_rows: the same bits
        ev2 = FusedOthelloEvaluator(net, big, use_cuda_graph=False, trunk_variant=2)
        ev2.trunk()
        torch.cuda.synchronize()
        assert torch.equal(ev2.features_rowmajor(), ev.features_rowmajor())
    ev(big)
    torch.cuda.synchronize()
    ref_p, ref_v = net.predict(x)
    assert (big.probs.cpu() - ref_p[idx]).abs().max().item() < 1e-2
    assert (big.value.cpu() - ref_v.reshape(-1)[idx]).abs().max().item() < 1e-2


@pytest.mark.parametrize("G", [2073, 2087, 3001, 4097, 8200])
def test_trunk_rows2_bit_identical_to_trunk_rows(G):
    """Batches with more groups of 14 boards than SMs go to k_net_trunk_rows2 (two CTAs per SM; from 297 groups on a CTA
    runs several groups back to back): the same features, bit for bit, as k_net_trunk_rows on every size -- a last
    group of 1, 5, 9 boards, one CTA more than SMs, CTAs with two groups."""
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.fused_net import FusedOthelloEvaluator
    net = _net(8, seed=21)
    tree, x, G0 = _setup(8, 500, seed=23)
    idx = (torch.arange(G) * 7) % G0
    big = TreeBatch(G, n=8, mode="neural", capacity=64)
    big.leaf_own.copy_(tree.leaf_own[idx.cuda()]); big.leaf_opp.copy_(tree.leaf_opp[idx.cuda()])
    feats = []
    for variant in (0, 2, 0):               # twice through the default: the second launch finds used shared memory / TMEM
        ev = FusedOthelloEvaluator(net, big, use_cuda_graph=False, trunk_variant=variant)
        ev.features.fill_(-3.0)
        ev.trunk()
        torch.cuda.synchronize()
        feats.append(ev.features_rowmajor().clone())
    assert torch.equal(feats[0], feats[1]) and torch.equal(feats[0], feats[2])
    with torch.no_grad():
        a = x.view(-1, 1, 8, 8)
        for conv, bn in ((net.conv1, net.bn1), (net.conv2, net.bn2), (net.conv3, net.bn3), (net.conv4, net.bn4)):
            a = F.relu(bn(conv(a)))
    got = feats[0].float().cpu().reshape(G, 16, 32).permute(0, 2, 1).reshape(G, 32, 4, 4)
    assert (got - a[idx]).abs().max().item() < 0.02 * a.abs().max().item() + 0.02


@pytest.mark.parametrize("n,graph", [(8, True), (8, False), (6, True)])
def test_fused_forward_within_1e2_of_fp32(n, graph):
    from alphazero_b200.fused_net import FusedOthelloEvaluator
    net = _net(n, seed=3)
    tree, x, G = _setup(n, 700, seed=5)
    ev = FusedOthelloEvaluator(net, tree, use_cuda_graph=graph)
    ev(tree)
    ev(tree)
    torch.cuda.synchronize()
    ref_p, ref_v = net.predict(x)
    assert (tree.probs.cpu() - ref_p).abs().max().item() < 1e-2
    assert (tree.value.cpu() - ref_v.reshape(-1)).abs().max().item() < 1e-2
    assert torch.allclose(tree.probs.sum(dim=1).cpu(), torch.ones(G), atol=1e-4)


def test_fused_forward_large_batch_and_tail_groups():
    """4096 + 3 positions: more groups than SMs (persistent loop) and a partially filled last group."""
    from alphazero_b200 import boardops as bo
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.fused_net import FusedOthelloEvaluator
    n = 8
    net = _net(n, seed=7)
    black, white, player = _random_positions(n, 600, 9)
    own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
    G = 4099
    idx = np.arange(G) % len(own)
    tree = TreeBatch(G, n=n, mode="neural", capacity=64)
    tree.leaf_own.copy_(bo.bits_from_numpy(own[idx])); tree.leaf_opp.copy_(bo.bits_from_numpy(opp[idx]))
    ev = FusedOthelloEvaluator(net, tree)
    ev(tree)
    torch.cuda.synchronize()
    x = bo.encode(tree.leaf_own[:len(own)], tree.leaf_opp[:len(own)], n).cpu()
    ref_p, ref_v = net.predict(x)
    p, v = tree.probs.cpu(), tree.value.cpu()
    assert (p[:len(own)] - ref_p).abs().max().item() < 1e-2
    # identical positions give identical outputs wherever they sit in the batch / group / tile
    assert (p[idx == 5] == p[5]).all() and (v[idx == 5] == v[5]).all()
    assert (p[-3:] == p[(G - 3) % len(own):(G - 3) % len(own) + 3]).all()


@pytest.mark.parametrize("n,G", [(8, 4096), (8, 300), (6, 1000)])
def test_fused_heads_kernel_matches_library_heads(n, G):
    """az_net_headgemm (tcgen05 GEMM + softmax / tanh epilogue, fp32 logits) against the cuBLASLt GEMM + az_net_heads
    path it replaces (bf16 logits): same inputs, outputs within the bf16 rounding of the logits; ragged last tile."""
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.fused_net import FusedNetEvaluator
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(n + G)
    net = OthelloNet(n=n)
    with torch.no_grad():
        net.fc_probs.weight.mul_(4.0)          # spread the policy so that the comparison is not between uniform vectors
    tree = TreeBatch(G, n=n, mode="neural", capacity=64)
    bmask = sum(1 << (r * 8 + c) for r in range(n) for c in range(n))
    a = torch.randint(0, 2**62, (G,), device="cuda") & bmask
    b = torch.randint(0, 2**62, (G,), device="cuda") & bmask & ~a
    tree.leaf_own.copy_(a); tree.leaf_opp.copy_(b)
    outs = {}
    for heads in ("library", "fused", "stack"):
        ev = FusedNetEvaluator(net, tree, heads=heads)
        tree.probs.zero_(); tree.value.zero_()
        ev()
        torch.cuda.synchronize()
        outs[heads] = (tree.probs.clone(), tree.value.clone())
    assert (outs["fused"][0].sum(1) - 1).abs().max() < 1e-5
    assert (outs["fused"][0] - outs["library"][0]).abs().max() < 3e-3
    assert (outs["fused"][1] - outs["library"][1]).abs().max() < 5e-3
    # the hand-written fc stack against the three library GEMMs: both round h1 / h2 / logits to bf16
    assert (outs["stack"][0] - outs["library"][0]).abs().max() < 3e-3
    assert (outs["stack"][1] - outs["library"][1]).abs().max() < 8e-3
    # against the module itself in fp32: the north star's 1e-2
    net32 = net.to("cuda"); net32.device = torch.device("cuda")
    from alphazero_b200.gamespec import OthelloSpec
    p32, v32 = net32.predict(OthelloSpec(n).encode(a, b))
    assert (outs["fused"][0] - p32).abs().max() < 1e-2 and (outs["fused"][1] - v32.reshape(-1)).abs().max() < 1e-2
    assert (outs["stack"][0] - p32).abs().max() < 1e-2 and (outs["stack"][1] - v32.reshape(-1)).abs().max() < 1e-2


@pytest.mark.parametrize("K1,N1,N2,n_out,RT,NS,G", [
    (512, 1024, 512, 66, 1, 4, 4096),    # OthelloNet 8x8 at the bench's batch: 32 clusters of 4, k_net_fcstack2
    (128, 1024, 512, 38, 1, 4, 700),     # OthelloNet 6x6 on k_net_fcstack2, ragged last tile
    (192, 512, 256, 80, 1, 4, 333),      # k_net_fcstack2: odd number of input chunks, one block per pass, 64-column fc2 slices
    (512, 1024, 512, 96, 1, 4, 200),     # more outputs than k_net_fcstack2's exchange buffer holds: the first kernel
    (512, 1024, 512, 66, 2, 4, 4096),    # k_net_fcstack3: 16 groups of 4 CTA pairs (cta_group::2)
    (512, 1024, 512, 96, 2, 4, 300),     # the first kernel on clusters of 8, ragged
    (512, 1024, 512, 66, 2, 4, 300),     # ragged: three row tiles, the second cluster half empty
    (512, 1024, 512, 66, 1, 4, 129),
    (512, 1024, 512, 66, 1, 8, 1000),
    (128, 1024, 512, 38, 2, 4, 700),     # OthelloNet 6x6
    (192, 64, 32, 8, 1, 1, 2048),        # Connect4Net: one CTA per row tile, 32-wide last chunk
    (192, 64, 32, 8, 2, 1, 333),
])
def test_fcstack_raw_shapes(K1, N1, N2, n_out, RT, NS, G):
    """az_net_fcstack through the C ABI on random operands against float64 torch (h1 / h2 rounded to bf16 as the
    kernel's epilogues do): every cluster shape the evaluators use, ragged last tiles, pad columns untouched."""
    from alphazero_b200 import _lib
    from alphazero_b200.fused_net import pack_fcstack, pack_rows
    torch.manual_seed(K1 + G)
    w1 = torch.randn(N1, K1, dtype=torch.float64) / K1 ** 0.5
    w2 = torch.randn(N2, N1, dtype=torch.float64) / N1 ** 0.5
    wh = torch.randn(n_out, N2, dtype=torch.float64) / N2 ** 0.5
    b1, b2, bh = (torch.randn(n, dtype=torch.float64) * 0.1 for n in (N1, N2, n_out))
    x = torch.randn(G, K1).to(torch.bfloat16)
    bf = lambda t: t.float().to(torch.bfloat16)
    blob, plan = pack_fcstack(w1, b1, w2, b2, wh, bh, NS, RT)
    blob = blob.cuda()
    tiles = -(-((G + 127) // 128) // RT) * RT
    xp = torch.zeros(tiles * 128, K1, dtype=torch.bfloat16)
    xp[:G] = x
    xp = pack_rows(xp, plan["kch"][0]).cuda()
    s1 = torch.zeros(tiles * 128 * N1, dtype=torch.bfloat16, device="cuda")
    s2 = torch.zeros(tiles * 128 * N2, dtype=torch.bfloat16, device="cuda")
    ld = ((n_out + 7) // 8) * 8
    logits = torch.full((G, ld), -77.0, dtype=torch.bfloat16, device="cuda")
    for _ in range(2):                   # twice: the scratch buffers are reused
        _lib.check(_lib.lib().az_net_fcstack(xp.data_ptr(), blob.data_ptr(), s1.data_ptr(), s2.data_ptr(),
                                             logits.data_ptr(), ld, n_out, G, K1, N1, N2, RT, NS,
                                             torch.cuda.current_stream().cuda_stream), "az_net_fcstack")
    torch.cuda.synchronize()
    xd = x.cuda()
    h1 = bf(torch.relu(xd.double() @ bf(w1).cuda().double().t() + b1.float().double().cuda()))
    h2 = bf(torch.relu(h1.double() @ bf(w2).cuda().double().t() + b2.float().double().cuda()))
    want = h2.double() @ bf(wh).cuda().double().t() + bh.float().double().cuda()
    got = logits[:, :n_out].double()
    # one bf16 ulp of the logits (+ the rare h1 / h2 element that rounds the other way under fp32 accumulation)
    assert (got - want).abs().max().item() <= 2.0 ** -7 * max(1.0, want.abs().max().item())
    assert (logits[:, n_out:] == -77.0).all()
    # argument checks: shapes without a mapping are refused, not mis-launched
    assert _lib.lib().az_net_fcstack(xp.data_ptr(), blob.data_ptr(), s1.data_ptr(), s2.data_ptr(), logits.data_ptr(),
                                     ld, n_out, G, K1, N1, N2, 4, 4, 0) != 0
    assert _lib.lib().az_net_fcstack(xp.data_ptr(), blob.data_ptr(), s1.data_ptr(), s2.data_ptr(), logits.data_ptr(),
                                     n_out - 1, n_out, G, K1, N1, N2, RT, NS, 0) != 0


@pytest.mark.parametrize("K,A,G", [(32, 65, 200), (64, 7, 129), (512, 37, 1000), (256, 95, 64)])
def test_headgemm_raw_shapes(K, A, G):
    """az_net_headgemm through the C ABI on random operands, over the feature / action sizes it accepts (including
    feature counts whose operand tile is smaller than the probs staging area), against float64 torch."""
    from alphazero_b200 import _lib
    torch.manual_seed(K + A)
    NH = ((A + 1 + 15) // 16) * 16
    h2 = torch.randn(G, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(A + 1, K) / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn(A + 1)
    tiles = np.zeros((K // 8, NH, 8), dtype=np.uint16)
    tiles[:, :A + 1, :] = w.view(torch.int16).numpy().view(np.uint16).reshape(A + 1, K // 8, 8).transpose(1, 0, 2)
    b = np.zeros(NH, dtype=np.float32); b[:A + 1] = bias.numpy()
    blob = torch.from_numpy(np.concatenate([tiles.reshape(-1).view(np.uint8), b.view(np.uint8)]).copy()).cuda()
    assert blob.numel() == _lib.lib().az_net_headgemm_blob_bytes(K, NH)
    probs = torch.full((G, A), -1.0, device="cuda"); value = torch.full((G,), -9.0, device="cuda")
    _lib.check(_lib.lib().az_net_headgemm(h2.data_ptr(), K, blob.data_ptr(), NH, A, G, probs.data_ptr(), value.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "az_net_headgemm")
    torch.cuda.synchronize()
    logits = h2.double().cpu() @ w.double().t() + bias.double()
    assert (probs.cpu().double() - torch.softmax(logits[:, :A], 1)).abs().max() < 1e-5
    assert (value.cpu().double() - torch.tanh(logits[:, A])).abs().max() < 1e-5
    # argument checks: unsupported sizes are refused, not mis-launched
    assert _lib.lib().az_net_headgemm(h2.data_ptr(), K + 8, blob.data_ptr(), NH, A, G, probs.data_ptr(), value.data_ptr(), 0) != 0
    assert _lib.lib().az_net_headgemm(h2.data_ptr(), K, blob.data_ptr(), NH, NH, G, probs.data_ptr(), value.data_ptr(), 0) != 0


@pytest.mark.parametrize("game", ["othello8", "othello6", "connect4"])
def test_tree_step_reading_logits_gives_identical_search(game):
    """az_tree.logits: the tree step takes softmax / tanh of the evaluator's raw bf16 rows itself (inside the
    whole-search graph az_net_heads is not launched at all) -- node statistics, priors and chosen moves must be
    bit-identical to the search that reads az_net_heads' probs / value."""
    from alphazero_b200.engine import TreeBatch
    from alphazero_b200.fused_net import FusedNetEvaluator
    from alphazero_b200.gamespec import Connect4Spec, OthelloSpec
    from alphazero_b200.games.connect4 import Connect4Net
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(5)
    if game == "connect4":
        spec, net = Connect4Spec(7, 6), Connect4Net(board_width=7, board_height=6)
    else:
        n = int(game[-1])
        spec, net = OthelloSpec(n), OthelloNet(n=n)
    with torch.no_grad():
        net.fc_probs.weight.mul_(4.0)
    G, n_sim = 300, 60
    stats = {}
    for in_tree in (False, True):
        tree = TreeBatch(G, n=spec, mode="neural", n_sim_hint=n_sim, seed=11, tie_random=False)
        tree.reset_roots()
        ev = FusedNetEvaluator(net, tree, heads_in_tree=in_tree)
        assert (ev.logits_binding is not None) == in_tree
        rec = []
        for move in range(4):
            tree.search(n_sim, ev)
            N, Q, P, rn = tree.root_stats()
            action, pi = tree.choose(None)
            rec.append((N.clone(), Q.clone(), P.clone(), rn.clone(), action.clone()))
            tree.advance(action)
            tree.check_status()
        assert any(e[0] is not None for e in tree._search_graphs.values())      # moves 1.. replayed the search graph
        # first search from the host loop (trunk + az_net_heads per step), three from the graph
        per_call = 4 if game == "connect4" else 3   # trunk, fcstack, heads (+ az_connect4_image_bits)
        assert ev.own_launches == (n_sim + 1) * per_call + 3 * (n_sim + 1) * (per_call - (1 if in_tree else 0))
        stats[in_tree] = rec
    for a, b in zip(stats[False], stats[True]):
        for x, y in zip(a, b):
            assert torch.equal(x, y)
