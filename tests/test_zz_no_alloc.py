"""B200: az_net_fcstack_static -- the fully connected stack with its h1 / h2 scratch in a static device array inside
the library, for callers that cannot allocate device memory -- against az_net_fcstack (caller-allocated scratch) on the
same operands, through the C ABI.  No counterpart in the reference.
Named so that pytest collects it last."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 2048 * 1536 * 2          # AZ_FCSTACK_STATIC_BYTES of csrc/fcstack_kernels.cu


def _operands(K1, N1, N2, n_out, RT, NS, G, seed):
    from alphazero_b200.fused_net import pack_fcstack, pack_rows
    g = torch.Generator().manual_seed(seed)
    w1 = torch.randn(N1, K1, generator=g, dtype=torch.float64) / K1 ** 0.5
    w2 = torch.randn(N2, N1, generator=g, dtype=torch.float64) / N1 ** 0.5
    wh = torch.randn(n_out, N2, generator=g, dtype=torch.float64) / N2 ** 0.5
    b1, b2, bh = (torch.randn(n, generator=g, dtype=torch.float64) * 0.1 for n in (N1, N2, n_out))
    blob, plan = pack_fcstack(w1, b1, w2, b2, wh, bh, NS, RT)
    tiles = -(-((G + 127) // 128) // RT) * RT
    xs = []
    for i in range(2):                   # two different inputs: the second call must not see the first one's h1 / h2
        x = torch.zeros(tiles * 128, K1, dtype=torch.bfloat16)
        x[:G] = (torch.randn(G, K1, generator=g) * (1.0 + i)).to(torch.bfloat16)
        xs.append(pack_rows(x, plan["kch"][0]).cuda())
    return blob.cuda(), xs, tiles


@pytest.mark.parametrize("K1,N1,N2,n_out,RT,NS,G", [
    (512, 1024, 512, 66, 1, 4, 2048),    # OthelloNet 8x8 (k_net_fcstack2) at the static array's limit
    (192, 512, 256, 60, 1, 2, 1100),     # the first kernel: leaves h1 / h2 activations where the next case's counters lie
    (512, 1024, 512, 66, 2, 4, 300),     # CTA pairs (k_net_fcstack3): its counters in the h2 region must be cleared first
    (192, 64, 32, 8, 1, 1, 1000),        # Connect4Net
])
def test_static_scratch_matches_the_allocating_entry_point(K1, N1, N2, n_out, RT, NS, G):
    from alphazero_b200 import _lib
    L = _lib.lib()
    blob, xs, tiles = _operands(K1, N1, N2, n_out, RT, NS, G, seed=K1 + G)
    ld = ((n_out + 7) // 8) * 8
    stream = torch.cuda.current_stream().cuda_stream
    # every buffer the test needs exists before the calls under test: they must not allocate
    h1 = torch.zeros(tiles * 128 * N1, dtype=torch.bfloat16, device="cuda")
    h2 = torch.zeros(tiles * 128 * N2, dtype=torch.bfloat16, device="cuda")
    want = [torch.full((G, ld), -77.0, dtype=torch.bfloat16, device="cuda") for _ in xs]
    got = [torch.full((G, ld), -77.0, dtype=torch.bfloat16, device="cuda") for _ in xs]
    for x, out in zip(xs, want):
        _lib.check(L.az_net_fcstack(x.data_ptr(), blob.data_ptr(), h1.data_ptr(), h2.data_ptr(), out.data_ptr(), ld,
                                    n_out, G, K1, N1, N2, RT, NS, stream), "az_net_fcstack")
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    for x, out in zip(xs, got):          # two calls in a row on the shared static array
        _lib.check(L.az_net_fcstack_static(x.data_ptr(), blob.data_ptr(), out.data_ptr(), ld, n_out, G, K1, N1, N2,
                                           RT, NS, stream), "az_net_fcstack_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
    for w, g in zip(want, got):
        scale = max(1.0, w[:, :n_out].float().abs().max().item())
        # the tolerance of test_fcstack_raw_shapes (one bf16 ulp of the logits); the kernels are deterministic, so in
        # fact the two entry points agree bit for bit
        assert (g[:, :n_out].float() - w[:, :n_out].float()).abs().max().item() <= 2.0 ** -7 * scale
        assert torch.equal(g, w)
        assert (g[:, n_out:] == -77.0).all()
    assert not torch.equal(got[0], got[1])


def test_static_scratch_rejects_what_does_not_fit():
    from alphazero_b200 import _lib
    L = _lib.lib()
    K1, N1, N2, n_out, RT, NS = 512, 1024, 512, 66, 1, 4
    G = 2049                             # 17 row tiles: 17 * 128 * 1536 * 2 bytes > the array
    assert -(-G // 128) * 128 * (N1 + N2) * 2 > STATIC_BYTES
    dummy = torch.zeros(64, dtype=torch.uint8, device="cuda")        # never touched: the call fails before any launch
    rc = L.az_net_fcstack_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 72, n_out, G, K1, N1, N2, RT, NS,
                                 torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.AZ_ERR_CAPACITY
    msg = L.az_last_cuda_error()
    msg = msg.decode() if isinstance(msg, bytes) else ctypes.cast(msg, ctypes.c_char_p).value.decode()
    assert "az_net_fcstack_static" in msg and "2049 positions" in msg and str(STATIC_BYTES) in msg
    torch.cuda.synchronize()             # nothing was launched, the context is healthy
    # shapes the kernels have no mapping for are refused like in az_net_fcstack
    assert L.az_net_fcstack_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 72, n_out, 128, 500, N1, N2,
                                   RT, NS, 0) == _lib.AZ_ERR_ARG
    assert L.az_net_fcstack_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 72, n_out, 0, K1, N1, N2,
                                   RT, NS, 0) == _lib.AZ_OK
