"""CPU: the host side of az_net_fcstack -- plan, weight packing, operand packing -- against a NumPy walk-through of the
kernel's own address arithmetic (tests/fcstack_emulator.py) and the plain matrix products it must equal."""
import numpy as np
import pytest
import torch

from fcstack_emulator import emulate, emulate2


def _stack(K1, N1, N2, n_out, seed):
    g = torch.Generator().manual_seed(seed)
    w1 = torch.randn(N1, K1, generator=g, dtype=torch.float64) / K1 ** 0.5
    w2 = torch.randn(N2, N1, generator=g, dtype=torch.float64) / N1 ** 0.5
    wh = torch.randn(n_out, N2, generator=g, dtype=torch.float64) / N2 ** 0.5
    b1, b2, bh = (torch.randn(n, generator=g, dtype=torch.float64) * 0.1 for n in (N1, N2, n_out))
    return w1, b1, w2, b2, wh, bh


def _bf(x):
    return x.float().to(torch.bfloat16).float()


@pytest.mark.parametrize("K1,N1,N2,n_out,RT,NS,G", [
    (512, 1024, 512, 66, 2, 4, 300),     # OthelloNet 8x8 on CTA pairs (variant 3): three row tiles -> two groups
    (128, 1024, 512, 38, 2, 4, 200),     # OthelloNet 6x6 on CTA pairs
    (512, 1024, 512, 96, 2, 4, 300),     # the first kernel on clusters of 8, the last one half empty
    (512, 1024, 512, 66, 1, 4, 300),     # the evaluator's shape: k_net_fcstack2 (variant 2 of the plan)
    (128, 1024, 512, 38, 1, 4, 128),     # OthelloNet 6x6 (variant 2)
    (192, 512, 256, 80, 1, 4, 130),      # variant 2 with one block per pass, an odd number of input chunks, 64-column fc2 slices
    (512, 1024, 512, 96, 1, 4, 100),     # too many outputs for variant 2's exchange buffer: the first kernel
    (128, 1024, 512, 38, 1, 8, 70),      # eight column slices
    (192, 64, 32, 8, 1, 1, 200),         # Connect4Net: no cluster
    (192, 64, 32, 8, 2, 1, 500),         # Connect4Net, the weights multicast over two row tiles
])
def test_packing_matches_the_kernels_addressing(K1, N1, N2, n_out, RT, NS, G):
    from alphazero_b200.fused_net import pack_fcstack, pack_rows
    w1, b1, w2, b2, wh, bh = _stack(K1, N1, N2, n_out, seed=K1 + NS)
    blob, plan = pack_fcstack(w1, b1, w2, b2, wh, bh, NS, RT)
    assert blob.numel() == plan["bytes"]
    assert plan["variant"] == ({1: 2, 2: 3}[RT] if NS == 4 and n_out <= 80 else 1)
    x = _bf(torch.randn(G, K1, generator=torch.Generator().manual_seed(1)))
    tiles = -(-((G + 127) // 128) // RT) * RT
    xp = torch.zeros(tiles * 128, K1)
    xp[:G] = x
    packed = pack_rows(xp.to(torch.bfloat16), plan["kch"][0]).reshape(-1).view(torch.uint8).numpy()
    ld = ((n_out + 7) // 8) * 8
    if plan["variant"] >= 2:
        got = emulate2(packed, blob.numpy(), plan, G, ld)
    else:
        got = emulate(packed, blob.numpy(), plan, G, ld, RT, NS)
    h1 = _bf(torch.relu(x.double() @ _bf(w1).double().t() + b1.float().double()))
    h2 = _bf(torch.relu(h1.double() @ _bf(w2).double().t() + b2.float().double()))
    want = _bf(h2.double() @ _bf(wh).double().t() + bh.float().double()).numpy()
    assert np.isnan(got[:, n_out:]).all()            # the pad columns of a logits row are never written
    assert np.isfinite(got[:, :n_out]).all()
    # fp32 accumulation order differs from the float64 product: allow a bf16 ulp of the logits
    assert np.abs(got[:, :n_out] - want).max() <= 2.0 ** -7 * max(1.0, np.abs(want).max())


def test_plan_refuses_shapes_without_a_mapping():
    from alphazero_b200.fused_net import fcstack_plan
    with pytest.raises(ValueError):
        fcstack_plan(512, 1024, 512, 66, 2)          # a slice of fc1 would not fit its 256 accumulator columns
    with pytest.raises(ValueError):
        fcstack_plan(500, 1024, 512, 66, 4)          # K1 not a whole number of chunks
    p = fcstack_plan(512, 1024, 512, 66, 4, RT=2)
    assert p["variant"] == 3 and p["Nc"] == [128, 128, 80]
    assert fcstack_plan(192, 512, 256, 80, 4, RT=2)["variant"] == 1      # partial sums would not fit the h2 scratch
    p = fcstack_plan(512, 1024, 512, 96, 4, RT=2)
    assert p["variant"] == 1
    assert p["kch"] == [64, 64, 64] and p["Nc"] == [256, 128, 32] and p["hpc"] == 24      # even: bf16 pairs
    assert p["bytes"] == (1024 * 512 + 512 * 1024 + 128 * 512) * 2 + (1024 + 512 + 128) * 4
    p = fcstack_plan(512, 1024, 512, 66, 4)
    assert p["variant"] == 2 and p["Nc"] == [128, 128, 80]       # two passes of 128 columns, all heads padded to 80
    assert p["bytes"] == (1024 * 512 + 512 * 1024 + 80 * 512) * 2 + (1024 + 512 + 80) * 4


def test_pack_rows_round_trip():
    from alphazero_b200.fused_net import pack_rows, unpack_rows
    x = torch.arange(300 * 128, dtype=torch.float32).reshape(300, 128).to(torch.bfloat16)
    p = pack_rows(x)
    assert p.shape == (3, 2, 8, 128, 8)
    assert torch.equal(unpack_rows(p, 300, 128), x)
    # element (g, k) sits at tile g/128, chunk k/64, plane (k%64)/8, row g%128, k%8 -- the trunk's store address
    g, k = 133, 77
    assert p[g // 128, k // 64, (k % 64) // 8, g % 128, k % 8] == x[g, k]
