"""Multi-GPU plumbing: games shard across ranks with no data-path collective; the single exchange is the
all-gather of the self-play samples to the training rank.  One process per GPU, torch.distributed
(NCCL on GPUs, gloo in the CPU tests)."""
import datetime

import torch
import torch.distributed as dist

from .selfplay import SampleBatch


def rank_world(group=None):
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


_HOST_GROUPS = {}
HOST_WAIT = datetime.timedelta(hours=24)


def host_barrier(group=None) -> None:
    """Host-side rendezvous (gloo, 24 h timeout) entered before every collective of the training loop.

    Rank 0 optimises, evaluates and saves alone while the other ranks already stand in the next collective
    (AlphaZeroTrainer.train): at the real configurations that takes longer than NCCL's watchdog allows an enqueued
    collective to wait (10 minutes by default), which would abort the job.  Waiting HERE -- on the host, through a
    side process group with a long timeout -- keeps every NCCL collective short-lived: it is only enqueued once all
    ranks have arrived.  The side group is created collectively at the first call (every rank reaches the same call
    sites in the same order)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    key = id(group) if group is not None else 0
    if key not in _HOST_GROUPS:
        ranks = None if group is None else dist.get_process_group_ranks(group)
        _HOST_GROUPS[key] = dist.new_group(ranks=ranks, backend="gloo", timeout=HOST_WAIT)
    dist.barrier(group=_HOST_GROUPS[key])


def broadcast_network(nn: torch.nn.Module, src: int = 0, group=None) -> None:
    """Rank `src`'s parameters and buffers (BatchNorm statistics included) replace everybody's, as ONE flat
    broadcast per dtype -- the weight hand-off after AlphaZeroTrainer.update_network (trainer.py:383-387) when
    the self-play games are sharded over ranks.  OthelloNet: 1.1 M parameters = 4.5 MB."""
    host_barrier(group)
    tensors = [t for t in list(nn.parameters()) + list(nn.buffers())]
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault(t.dtype, []).append(t)
    for dtype, ts in by_dtype.items():
        flat = torch.cat([t.detach().reshape(-1) for t in ts])
        if dist.get_backend(group) == "nccl" and not flat.is_cuda:      # a network kept on the CPU (config.device)
            flat = flat.cuda()
        dist.broadcast(flat, src=src, group=group)
        off = 0
        with torch.no_grad():
            for t in ts:
                t.copy_(flat[off: off + t.numel()].reshape(t.shape))
                off += t.numel()


def shard_range(total: int, rank: int, world: int):
    """Games [lo, hi) owned by `rank`: contiguous, sizes differ by at most one."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_samples(batch: SampleBatch, episode_offset: int = 0, group=None) -> SampleBatch:
    """All-gathers variable-length sample batches: counts first, then payloads padded to the longest rank.
    Every rank gets the concatenation in rank order (episode indices shifted by each rank's `episode_offset`)."""
    host_barrier(group)
    world = dist.get_world_size(group)
    dev = batch.own.device
    count = torch.tensor([len(batch)], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(count) for _ in range(world)]
    dist.all_gather(counts, count, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(counts)

    def pad(t):
        out = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        out[: t.shape[0]] = t
        return out

    def gather(t):
        mine = pad(t)
        bufs = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(bufs, mine, group=group)
        return torch.cat([b[:c] for b, c in zip(bufs, counts)], dim=0)

    ep = batch.episode_idx + int(episode_offset)
    return SampleBatch(gather(batch.own), gather(batch.opp), gather(batch.pi), gather(batch.z), gather(ep),
                       gather(batch.move_idx), batch.n, batch.rows, None, batch.game)
