"""CPU: host-side logic of the batched self-play engine that needs no device (tail-tier selection, constructor guards)."""
import pytest
import torch

from alphazero_b200.selfplay import SampleBatch, default_tail_tiers, normalise_tail_tiers, pick_tail_tier


def test_default_tail_tiers():
    assert default_tail_tiers(4096) == (64, 512, 2048)
    assert default_tail_tiers(2048) == (32, 256, 1024)      # BASELINE config 5's per-GPU share
    assert default_tail_tiers(1024) == (16, 128, 512)
    assert default_tail_tiers(1023) == () and default_tail_tiers(48) == ()


def test_normalise_tail_tiers():
    assert normalise_tail_tiers("auto", 4096) == (64, 512, 2048)
    assert normalise_tail_tiers((), 4096) == () and normalise_tail_tiers(None, 4096) == ()
    assert normalise_tail_tiers((24, 6, 2, 24, 0, 48, 100), 48) == (2, 6, 24)    # distinct, inside (0, G), ascending
    with pytest.raises(ValueError):
        normalise_tail_tiers("always", 4096)


def test_pick_tail_tier_follows_the_live_counts_of_a_batch():
    """Live games after plies 59..65 of a 4096-game Othello batch (profiles/r1_move_phases_othello.txt)."""
    tiers = default_tail_tiers(4096)
    size, seen = 4096, []
    for n_live in (4075, 1481, 285, 29, 8, 1, 1):
        nxt = pick_tail_tier(n_live, size, tiers)
        if nxt is not None:
            size = nxt
        seen.append(size)
    assert seen == [4096, 2048, 512, 64, 64, 64, 64]
    assert pick_tail_tier(2048, 4096, tiers) == 2048        # fits exactly
    assert pick_tail_tier(10, 64, tiers) is None             # already in the smallest tier
    assert pick_tail_tier(10, 4096, ()) is None               # tiers switched off
    assert pick_tail_tier(30, 4096, tiers) == 64              # skips tiers it does not need


def test_engine_refuses_to_run_without_a_device():
    if torch.cuda.is_available():
        pytest.skip("checks the CPU-only failure mode")
    from alphazero_b200.selfplay import SelfPlayEngine
    with pytest.raises(RuntimeError):
        SelfPlayEngine(None, 8, n=6, n_sim=2)                 # no CPU fallback: TreeBatch raises


def test_sample_batch_cpu_is_identity_on_host_tensors():
    z = torch.zeros(3, dtype=torch.int8)
    b = SampleBatch(torch.zeros(3, dtype=torch.int64), torch.zeros(3, dtype=torch.int64), torch.zeros(3, 37), z,
                    torch.zeros(3, dtype=torch.int32), torch.zeros(3, dtype=torch.int32), 6)
    assert b.cpu() is b and len(b) == 3


def test_trainer_memory_setter_rebuilds_the_batch():
    """trainer.memory = [...] (the reference lets callers assign it, trainer.py:219) must train on THOSE samples: the
    device batch is rebuilt from the list and round-trips."""
    import numpy as np
    from alphazero_b200.trainer import AlphaZeroTrainer, Sample, batch_from_samples, samples_from_batch
    rng = np.random.RandomState(0)
    samples = []
    for i in range(20):
        st = rng.choice([0.0, 1.0, -1.0], size=(6, 6))
        pi = rng.rand(37); pi /= pi.sum()
        player = 1 if i % 2 == 0 else -1
        samples.append(Sample(st * player, pi, player, outcome=int(rng.choice([-1, 0, 1])), episode_idx=i // 5, move_idx=i % 5))
    batch = batch_from_samples(samples, "othello", device="cpu")
    back = samples_from_batch(batch)
    for a, b in zip(samples, back):
        assert np.array_equal(a.state * a.player, b.state) and np.allclose(a.pi, b.pi, atol=1e-7)
        assert b.outcome == a.outcome * a.player and (b.episode_idx, b.move_idx) == (a.episode_idx, a.move_idx)
    tr = AlphaZeroTrainer()
    tr.game = "othello"
    tr.memory = samples
    assert tr.memory_batch is not None and len(tr.memory_batch) == 20 and tr.memory is samples
    tr.memory = None
    assert tr.memory_batch is None
