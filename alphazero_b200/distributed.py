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


_REC = 32      # bytes of a sample's fixed-size wire record: own 8 | opp 8 | episode 4 | move 4 | z 1 | action 1 | one-hot 1 | pad


def _pack_records(batch: SampleBatch, episode_offset: int):
    """(uint8 [S, 32] records, float32 [M, A] dense policies) -- the wire form of a batch.  From move 5 on the policy
    target is one-hot (temperature 0: trainer.py:241-246, schedulers.py:33-40), so it travels as one action byte; only
    the M samples whose pi is a real distribution send their A floats (about 8 % of an Othello game)."""
    S = len(batch)
    dev = batch.own.device
    pi = batch.pi
    if S == 0:
        return torch.zeros(0, _REC, dtype=torch.uint8, device=dev), pi.reshape(0, pi.shape[1] if pi.dim() == 2 else 0)
    top, act = pi.max(dim=1)
    onehot = (top == 1.0) & (pi.sum(dim=1) == 1.0)
    rec = torch.zeros(S, _REC, dtype=torch.uint8, device=dev)
    rec[:, 0:8] = batch.own.contiguous().view(torch.uint8).view(S, 8)
    rec[:, 8:16] = batch.opp.contiguous().view(torch.uint8).view(S, 8)
    rec[:, 16:20] = (batch.episode_idx + int(episode_offset)).to(torch.int32).contiguous().view(torch.uint8).view(S, 4)
    rec[:, 20:24] = batch.move_idx.to(torch.int32).contiguous().view(torch.uint8).view(S, 4)
    rec[:, 24] = batch.z.contiguous().view(torch.uint8)
    rec[:, 25] = act.to(torch.uint8)
    rec[:, 26] = onehot.to(torch.uint8)
    return rec, pi[~onehot].contiguous()


def _unpack_records(rec: torch.Tensor, dense: torch.Tensor, A: int, like: SampleBatch) -> SampleBatch:
    S = rec.shape[0]
    rec = rec.contiguous()
    col = lambda a, b, dt: rec[:, a:b].contiguous().view(dt).view(S)
    own, opp = col(0, 8, torch.int64), col(8, 16, torch.int64)
    ep, mv = col(16, 20, torch.int32), col(20, 24, torch.int32)
    z = rec[:, 24].contiguous().view(torch.int8)
    onehot = rec[:, 26].bool()
    pi = torch.zeros(S, A, dtype=torch.float32, device=rec.device)
    rows = torch.nonzero(onehot).flatten()
    pi[rows, rec[rows, 25].long()] = 1.0
    pi[~onehot] = dense
    return SampleBatch(own, opp, pi, z, ep, mv, like.n, like.rows, None, like.game)


def gather_samples(batch: SampleBatch, episode_offset: int = 0, group=None) -> SampleBatch:
    """All-gathers variable-length sample batches to every rank, in rank order (episode indices shifted by each rank's
    `episode_offset`): one all-gather of the counts, one of the 32-byte sample records, one of the dense policies --
    each a single padded buffer.  At 8 ranks x 4096 Othello games this is ~100 MB on the wire instead of the 564 MB of
    seven per-field all-gathers of float32 policies."""
    host_barrier(group)
    world = dist.get_world_size(group)
    dev = batch.own.device
    A = batch.pi.shape[1]
    rec, dense = _pack_records(batch, episode_offset)
    count = torch.tensor([rec.shape[0], dense.shape[0]], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(count) for _ in range(world)]
    dist.all_gather(counts, count, group=group)
    counts = [c.tolist() for c in counts]

    def gather(t, sizes):
        m = max(max(sizes), 1)
        mine = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        mine[: t.shape[0]] = t
        bufs = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(bufs, mine, group=group)
        return [b[:c] for b, c in zip(bufs, sizes)]

    recs = gather(rec, [c[0] for c in counts])
    denses = gather(dense, [c[1] for c in counts])
    return _unpack_records(torch.cat(recs, dim=0), torch.cat(denses, dim=0), A, batch)
