"""CPU: the parts of bench.py that read committed evidence (ncu summaries, measured peaks) -- no device needed."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("az_bench", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_roofline_traffic_comes_from_the_committed_ncu_summaries():
    b = _bench()
    trunk = b.ncu_traffic("k_net_trunk")
    tree_shallow = b.ncu_traffic("k_tree_step")
    tree_mid = b.ncu_traffic("k_tree_step", ("r1_ncu_tree_midgame.json", "r1_ncu_full_summary.json"))
    # the trunk reads 16 B per position + the weight blob once per CTA; activations never reach DRAM
    assert trunk is not None and 4096 * 16 <= trunk < 1e6
    # a tree step moves megabytes, more on the deeper trees of the middle game
    assert tree_shallow is not None and tree_mid is not None and 1e6 < tree_shallow <= tree_mid < 5e7
    assert b.ncu_traffic("no_such_kernel") is None
    assert b.ncu_traffic("k_tree_step", ("missing.json",)) is None


def test_measured_peaks_have_the_roofline_denominators():
    p = _bench().measured_peaks()
    assert p["source"] in ("measured", "fallback")
    assert 3000 < p["hbm_gbs"] < 9000
    assert 500 < p["bf16_tflops_sustained"] <= p["bf16_tflops"] < 2500


def test_default_invocation_is_the_headline_config():
    b = _bench()
    import sys
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        a = b.parse_args()
    finally:
        sys.argv = argv
    assert (a.gpus, a.games, a.sims, a.game, a.impl, a.net) == (1, 4096, 200, "othello", "ours", "fused")
    assert a.warmup >= 3 and a.steps >= 1 and not a.no_tail_tiers
