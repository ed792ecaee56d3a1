"""CPU: the weight blob of az_net_trunk (fused_net.pack_trunk_weights) -- its size is the library's, both sections hold
the folded conv2..conv4 weights at the addresses the two kernels read (include/az_b200.h: az_net_wblob_bytes)."""
import numpy as np
import torch


def test_trunk_blob_sections_hold_the_folded_weights():
    from alphazero_b200 import _lib
    from alphazero_b200.fused_net import _bf16_bits, _fold_bn, pack_trunk_weights
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(5)
    net = OthelloNet(n=8)
    net.train()
    with torch.no_grad():
        for _ in range(3):               # non-trivial BatchNorm statistics to fold
            net(torch.randint(-1, 2, (64, 8, 8)).float())
    net.eval()
    blob = pack_trunk_weights(net).numpy()
    assert blob.nbytes == _lib.lib().az_net_wblob_bytes() == 2 * 3 * 18432 + 1024 + 512
    first = blob[:3 * 18432].view(np.uint16).reshape(3, 3, 4, 3, 32, 8)       # [layer][kh][ci chunk][kw][co][ci % 8]
    rows = blob[3 * 18432 + 1536:].view(np.uint16).reshape(3, 3, 4, 3, 32, 8)  # [layer][kw][ci chunk][kh][co][ci % 8]
    for li, (conv, bn) in enumerate(((net.conv2, net.bn2), (net.conv3, net.bn3), (net.conv4, net.bn4))):
        w, b = _fold_bn(conv.weight, conv.bias, bn)                           # [co][ci][kh][kw]
        bits = _bf16_bits(w).reshape(32, 4, 8, 3, 3)                          # [co][ci chunk][ci % 8][kh][kw]
        assert np.array_equal(first[li], bits.transpose(3, 1, 4, 0, 2))
        assert np.array_equal(rows[li], bits.transpose(4, 1, 3, 0, 2))
        bias = blob[3 * 18432 + 1024:3 * 18432 + 1536].view(np.float32).reshape(4, 32)
        assert np.array_equal(bias[li + 1], b.float().cpu().numpy())
    c1 = blob[3 * 18432:3 * 18432 + 1024].view(np.uint16).reshape(2, 32, 8)   # conv1: [k chunk][co][k % 8], taps 0..8 then 0
    w1, _ = _fold_bn(net.conv1.weight, net.conv1.bias, net.bn1)
    taps = c1.transpose(1, 0, 2).reshape(32, 16)
    assert np.array_equal(taps[:, :9], _bf16_bits(w1.reshape(32, 9))) and not taps[:, 9:].any()
