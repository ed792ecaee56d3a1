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
    feat = ev.features.float().cpu().reshape(G, m * m, 32).permute(0, 2, 1).reshape(G, 32, m, m)
    assert (feat - a4).abs().max().item() < 0.02 * a4.abs().max().item() + 0.02


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
