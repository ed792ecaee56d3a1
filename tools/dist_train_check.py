"""Run under torchrun (one process per GPU, NCCL): two iterations of AlphaZeroTrainer.train with the episodes sharded
over the ranks; checks that every rank ends with the same memory size, bit-identical weights and a refreshed evaluator.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/dist_train_check.py [game]
"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist


def main():
    game = sys.argv[1] if len(sys.argv) > 1 else "othello"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    from alphazero_b200.games.registers import CONFIGS_REGISTER
    from alphazero_b200.trainer import AlphaZeroTrainer
    kw = dict(simulations=8, iterations=2, episodes=13, epochs=1, batch_size=32, eval_episodes=4, eval_opponent="random",
              do_eval=True, save=True, save_checkpoints=False, push=False, push_checkpoints=False)
    if game == "othello":
        kw["board_size"] = 6
    cfg = CONFIGS_REGISTER[game](**kw)
    torch.manual_seed(100 + rank)            # different initial weights per rank: train() must broadcast rank 0's
    np.random.seed(100 + rank)
    out = tempfile.mkdtemp() if rank == 0 else "/nonexistent-on-purpose"
    tr = AlphaZeroTrainer()
    tr.train(config=cfg, experiment_name="dist", models_path=out)
    n = torch.tensor([len(tr.memory_batch)], device="cuda")
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    assert len({int(s.item()) for s in sizes}) == 1, sizes
    flat = torch.cat([t.detach().float().reshape(-1).cuda() for t in list(tr.nn.parameters()) + list(tr.nn.buffers())])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    assert torch.equal(flat, ref), "weights differ between ranks"
    eps = tr.memory_batch.episode_idx
    assert int(eps.max().item()) == cfg.episodes - 1 and int(eps.min().item()) == 0      # global episode numbering
    if rank == 0:
        assert os.path.isfile(os.path.join(out, "dist", "dist.pt")) and os.path.isfile(os.path.join(out, "dist", "loss.json"))
        print(f"dist_train_check ok: game={game} world={world} samples={int(n.item())} params={flat.numel()}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
