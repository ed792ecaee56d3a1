"""CPU, world_size 2, gloo: the one exchange of the multi-GPU path -- all-gather of variable-length sample batches --
and the game sharding arithmetic."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _make_batch(rank, count, A=65):
    from alphazero_b200.selfplay import SampleBatch
    g = torch.Generator().manual_seed(100 + rank)
    return SampleBatch(
        own=torch.randint(0, 2**62, (count,), generator=g), opp=torch.randint(0, 2**62, (count,), generator=g),
        pi=torch.rand(count, A, generator=g), z=torch.randint(-1, 2, (count,), generator=g).to(torch.int8),
        episode_idx=torch.arange(count, dtype=torch.int32) // 3, move_idx=torch.arange(count, dtype=torch.int32) % 3, n=8)


def _worker(rank, world, port, counts, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from alphazero_b200.distributed import gather_samples
    mine = _make_batch(rank, counts[rank])
    full = gather_samples(mine, episode_offset=1000 * rank)
    torch.save({k: getattr(full, k) for k in ("own", "opp", "pi", "z", "episode_idx", "move_idx")},
               os.path.join(out_dir, f"rank{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("counts", [(7, 12), (5, 0), (9, 9)])
def test_gather_samples_world2(tmp_path, counts):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), counts, str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]
    parts = [_make_batch(r, counts[r]) for r in range(world)]
    want = {k: torch.cat([getattr(p, k) + (1000 * r if k == "episode_idx" else 0) for r, p in enumerate(parts)])
            for k in ("own", "opp", "pi", "z", "episode_idx", "move_idx")}
    for r in range(world):      # every rank holds the concatenation of the per-rank dumps, bit for bit
        for k, v in want.items():
            assert torch.equal(got[r][k], v.to(got[r][k].dtype)), (r, k)
    # a checksum of checksums: order-independent fingerprint equals the sum of the per-rank fingerprints
    fp = lambda b: int((b["own"] ^ b["opp"]).sum().item())
    assert fp(got[0]) == sum(int((p.own ^ p.opp).sum().item()) for p in parts)


def test_shard_range_partitions_games():
    from alphazero_b200.distributed import shard_range
    for total in (0, 1, 7, 4096, 32768, 16385):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _bcast_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from alphazero_b200.distributed import broadcast_network, rank_world
    from alphazero_b200.games.connect4 import Connect4Config, Connect4Net
    assert rank_world() == (rank, world)
    torch.manual_seed(1000 + rank)                      # every rank starts from different weights and BN statistics
    net = Connect4Net(config=Connect4Config())
    net.train()
    net(torch.randn(8, 6, 7))                           # moves the BatchNorm running statistics (buffers)
    broadcast_network(net, src=0)
    torch.save(net.state_dict(), os.path.join(out_dir, f"net{rank}.pt"))
    dist.destroy_process_group()


def test_broadcast_network_world2(tmp_path):
    """The weight hand-off of the sharded training loop: after the broadcast every rank holds rank 0's parameters AND
    buffers (BatchNorm running statistics, num_batches_tracked), bit for bit."""
    from alphazero_b200.games.connect4 import Connect4Config, Connect4Net
    world = 2
    mp.spawn(_bcast_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    sd = [torch.load(os.path.join(tmp_path, f"net{r}.pt")) for r in range(world)]
    torch.manual_seed(1000)
    want = Connect4Net(config=Connect4Config())
    want.train()
    want(torch.randn(8, 6, 7))
    for k, v in want.state_dict().items():
        assert torch.equal(sd[0][k], v), k
        assert torch.equal(sd[1][k], v), k
    assert int(sd[1]["bn1.num_batches_tracked"]) == 1
