"""CPU: the drop-in surface mirrors the reference's API (names, signatures, defaults, helper semantics).
The comparisons against the imported reference run only where /root/reference exists (the build container)."""
import dataclasses
import inspect

import numpy as np
import pytest
import torch

from ref_import import import_reference, reference_available

needs_ref = pytest.mark.skipif(not reference_available(), reason="reference checkout not present")


def _public_methods(cls):
    return {n: m for n, m in inspect.getmembers(cls, predicate=inspect.isfunction) if not n.startswith("_")}


@needs_ref
def test_class_surfaces_match_reference():
    import_reference()
    import alphazero.arena as r_arena
    import alphazero.base as r_base
    import alphazero.mcts as r_mcts
    import alphazero.players as r_players
    import alphazero.games.othello as r_oth
    import alphazero.games.connect4 as r_c4
    import alphazero.trainer as r_trainer
    import alphazero_b200.games.connect4 as c4
    import alphazero_b200.arena as arena
    import alphazero_b200.base as base
    import alphazero_b200.mcts as mcts
    import alphazero_b200.players as players
    import alphazero_b200.games.othello as oth
    import alphazero_b200.trainer as trainer
    pairs = [(r_base.Board, base.Board), (r_base.Player, base.Player), (r_base.PolicyValueNetwork, base.PolicyValueNetwork),
             (r_base.TemperatureScheduler, base.TemperatureScheduler), (r_mcts.MCT, mcts.MCT), (r_mcts.Node, mcts.Node),
             (r_players.RandomPlayer, players.RandomPlayer), (r_players.GreedyPlayer, players.GreedyPlayer),
             (r_players.MCTSPlayer, players.MCTSPlayer), (r_players.AlphaZeroPlayer, players.AlphaZeroPlayer),
             (r_oth.OthelloBoard, oth.OthelloBoard), (r_oth.OthelloNet, oth.OthelloNet), (r_arena.Arena, arena.Arena),
             (r_c4.Connect4Board, c4.Connect4Board), (r_c4.Connect4Net, c4.Connect4Net),
             (r_trainer.Sample, trainer.Sample), (r_trainer.AlphaZeroTrainer, trainer.AlphaZeroTrainer)]
    # estimate_training_duration: timers are out of scope
    skip = {"human_display", "pixel_display", "add_child", "select_node", "rollout", "nn_evaluation", "back_propagate",
            "estimate_training_duration"}
    for ref_cls, cls in pairs:
        for name, fn in _public_methods(ref_cls).items():
            if name in skip:
                continue
            assert hasattr(cls, name), f"{cls.__name__}.{name} missing"
            ref_params = [p for p in inspect.signature(fn).parameters]
            params = [p for p in inspect.signature(getattr(cls, name)).parameters]
            if params[-2:] == ["a", "k"] or "a" in params:   # abstract placeholders
                continue
            assert params[:len(ref_params)] == ref_params, f"{cls.__name__}.{name}: {params} vs {ref_params}"
        ref_init = list(inspect.signature(ref_cls.__init__).parameters)
        init = list(inspect.signature(cls.__init__).parameters)
        assert init[:len(ref_init)] == ref_init, f"{cls.__name__}.__init__: {init} vs {ref_init}"
    assert set(players.PLAYERS_REGISTER) == set(r_players.PLAYERS_REGISTER) - {"human"}


@needs_ref
def test_config_defaults_and_enums_match_reference():
    import_reference()
    import alphazero.base as r_base
    import alphazero.games.othello as r_oth
    import alphazero_b200.base as base
    import alphazero_b200.games.othello as oth
    assert dataclasses.asdict(oth.OthelloConfig()) == dataclasses.asdict(r_oth.OthelloConfig())
    assert dataclasses.asdict(base.Config()) == dataclasses.asdict(r_base.Config())
    for name in ("MoveFormat", "TreeEval", "DisplayMode", "DataTransf"):
        assert {m.name: m.value for m in getattr(base, name)} == {m.name: m.value for m in getattr(r_base, name)}
    assert base.TreeEval.to_dict().keys() == r_base.TreeEval.to_dict().keys()


@needs_ref
def test_schedulers_match_reference():
    import_reference()
    from alphazero.schedulers import TEMP_SCHEDULERS as R
    from alphazero_b200.schedulers import TEMP_SCHEDULERS as O
    for kind, args in (("linear", (4, 4, 60)), ("linear", (3, 9, 60)), ("constant", (5, 5, 32))):
        assert list(R[kind](*args)) == list(O[kind](*args))
    with pytest.raises(ValueError):
        O["linear"](5, 4, 10)
    with pytest.raises(ValueError):
        O["constant"](5, 4, 10)


@needs_ref
@pytest.mark.parametrize("n", [6, 8])
def test_othello_net_matches_reference(n):
    import_reference()
    from alphazero.games.othello import OthelloNet as RefNet
    from alphazero_b200.games.othello import OthelloNet
    torch.manual_seed(0)
    net = OthelloNet(n=n)
    ref = RefNet(n=n)
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    ref.load_state_dict(net.state_dict())
    assert net.get_parameters_count() == ref.get_parameters_count() == {6: 707782, 8: 1115362}[n]
    x = torch.randint(-1, 2, (7, n, n)).float()
    p, v = net.predict(x)
    rp, rv = ref.predict(x)
    assert torch.equal(p, rp) and torch.equal(v, rv)
    rng = np.random.default_rng(0)
    probs = rng.random(n * n + 1).astype(np.float32)
    legal = [(1, 2), (3, 3), (n, n)]
    a, b = net.get_normalized_probs(probs, legal), ref.get_normalized_probs(probs, legal)
    assert a.keys() == b.keys() and all(a[k] == b[k] and type(a[k]) is type(b[k]) for k in a)
    tiny = np.zeros(n * n + 1, np.float32)
    assert net.get_normalized_probs(tiny, legal) == ref.get_normalized_probs(tiny, legal)
    mp = {(0, 1): 0.25, (n, n): 0.75}
    assert (net.to_neural_output(mp) == ref.to_neural_output(mp)).all()
    out = rng.random(n * n + 1)
    for axis in (0, 1):
        assert (net.reflect_neural_output(out, axis) == ref.reflect_neural_output(out, axis)).all()
    for angle in (90, 180, 270):
        assert (net.rotate_neural_output(out, angle) == ref.rotate_neural_output(out, angle)).all()
    with pytest.raises(ValueError):
        net.rotate_neural_output(out[:-1], 90)


@needs_ref
def test_sample_augmentation_matches_reference():
    import_reference()
    from alphazero.base import DataTransf as RT
    from alphazero.trainer import Sample as RefSample
    from alphazero_b200.base import DataTransf
    from alphazero_b200.games.othello import OthelloNet
    from alphazero_b200.trainer import Sample
    net = OthelloNet(n=6)
    rng = np.random.default_rng(1)
    state = rng.integers(-1, 2, (6, 6)).astype(float)
    pi = rng.random(37)
    ours = Sample(state.copy(), pi.copy(), player=-1, outcome=1, episode_idx=3, move_idx=5)
    ref = RefSample(state.copy(), pi.copy(), player=-1, outcome=1, episode_idx=3, move_idx=5)
    ours.normalize(); ref.normalize()
    assert (ours.state == ref.state).all() and ours.outcome == ref.outcome and ours.player == ref.player == 1
    a = ours.create_reflection_twin(net.reflect_neural_output, DataTransf.REFLECT_H)
    b = ref.create_reflection_twin(net.reflect_neural_output, RT.REFLECT_H)
    assert (a.state == b.state).all() and (a.pi == b.pi).all() and a.transformation == b.transformation
    for mo, mr in ((DataTransf.ROTATE_90, RT.ROTATE_90), (DataTransf.ROTATE_180, RT.ROTATE_180), (DataTransf.ROTATE_270, RT.ROTATE_270)):
        x = a.create_rotation_twin(net.rotate_neural_output, mo)
        y = b.create_rotation_twin(net.rotate_neural_output, mr)
        assert (x.state == y.state).all() and (x.pi == y.pi).all() and x.transformation == y.transformation
    with pytest.raises(ValueError):
        ours.create_reflection_twin(net.reflect_neural_output, DataTransf.ROTATE_90)


def test_constructor_error_behaviour():
    from alphazero_b200.mcts import MCT
    from alphazero_b200.players import AlphaZeroPlayer, MCTSPlayer
    from alphazero_b200.arena import Arena
    from alphazero_b200.games.othello import OthelloNet
    with pytest.raises(ValueError):
        MCTSPlayer()
    with pytest.raises(ValueError):
        MCTSPlayer(n_sim=3, compute_time=0.1)
    with pytest.raises(ValueError):
        MCT(eval_method="rollout", nn=OthelloNet(n=6))
    m = MCT()
    with pytest.raises(ValueError):
        m.nn = OthelloNet(n=6)
    p = AlphaZeroPlayer(n_sim=5, nn=None, dirichlet_alpha=0.3, dirichlet_epsilon=0.25)
    assert p.mct.dirichlet_alpha == 0.3 and p.mct.nn is None
    with pytest.raises(ValueError):
        Arena._check_inputs(3)
    with pytest.raises(ValueError):
        Arena._check_inputs(12, n_process=3)
    with pytest.raises(ValueError):
        Arena._check_inputs(10, n_process=4)
    with pytest.raises(ValueError):
        Arena._check_inputs(4, start_player=3)
    with pytest.raises(ValueError):
        OthelloNet()


def test_fair_max_semantics():
    from alphazero_b200.utils import fair_max
    x = np.random.randn(100)
    assert fair_max(x) == x.max()
    np.random.seed(0)
    state = np.random.get_state()[1].copy()
    assert fair_max([1, 5, 2]) == 5                        # a unique maximum consumes no randomness
    assert (np.random.get_state()[1] == state).all()
    picks = {fair_max([("a", 1), ("b", 1), ("c", 0)], key=lambda t: t[1])[0] for _ in range(60)}
    assert picks == {"a", "b"}


@needs_ref
@pytest.mark.parametrize("w,h", [(7, 6), (5, 8)])
def test_connect4_net_and_config_match_reference(w, h):
    import_reference()
    import alphazero.games.connect4 as r_c4
    import alphazero.games.registers as r_reg
    import alphazero_b200.games.connect4 as c4
    import alphazero_b200.games.registers as reg
    assert dataclasses.asdict(c4.Connect4Config()) == dataclasses.asdict(r_c4.Connect4Config())
    for name in ("CONFIGS_REGISTER", "BOARDS_REGISTER", "NETWORKS_REGISTER", "MOVE_FORMATS_REGISTER", "DATA_AUGMENT_STRATEGIES"):
        assert set(getattr(reg, name)) == set(getattr(r_reg, name))
    assert reg.GAMES_SET == r_reg.GAMES_SET
    assert reg.MOVE_FORMATS_REGISTER["connect4"].value == r_reg.MOVE_FORMATS_REGISTER["connect4"].value
    assert [t.value for t in reg.DATA_AUGMENT_STRATEGIES["connect4"].rotations] == \
        [t.value for t in r_reg.DATA_AUGMENT_STRATEGIES["connect4"].rotations] == []
    torch.manual_seed(0)
    net = c4.Connect4Net(board_width=w, board_height=h)
    ref = r_c4.Connect4Net(board_width=w, board_height=h)
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    ref.load_state_dict(net.state_dict())
    assert net.get_parameters_count() == ref.get_parameters_count()
    if (w, h) == (7, 6):
        assert net.get_parameters_count() == 43208
    x = torch.randint(-1, 2, (9, h, w)).float()          # (height, width) grids, viewed as (width, height) inside
    p, v = net.predict(x)
    rp, rv = ref.predict(x)
    assert torch.equal(p, rp) and torch.equal(v, rv)
    rng = np.random.default_rng(0)
    probs = rng.random(w).astype(np.float32)
    legal = [np.int64(0), np.int64(2), np.int64(w - 1)]
    a, b = net.get_normalized_probs(probs, legal), ref.get_normalized_probs(probs, legal)
    assert a.keys() == b.keys() and all(a[k] == b[k] and type(a[k]) is type(b[k]) for k in a)
    tiny = np.zeros(w, np.float32)
    assert net.get_normalized_probs(tiny, legal) == ref.get_normalized_probs(tiny, legal)
    mp = {np.int64(1): 0.25, np.int64(3): 0.75}
    assert (net.to_neural_output(mp) == ref.to_neural_output(mp)).all()
    out = rng.random(w)
    assert (net.reflect_neural_output(out, 1) == ref.reflect_neural_output(out, 1)).all()
    with pytest.raises(ValueError):
        net.reflect_neural_output(out[:-1], 1)
    with pytest.raises(ValueError):
        c4.Connect4Net(board_width=3, board_height=6)


def test_connect4_golden_net_loads(golden_dir=None):
    """The committed reference state_dict + outputs (tests/golden/connect4_net_7x6.npz) reproduce on CPU."""
    import os
    from alphazero_b200.games.connect4 import Connect4Net
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "connect4_net_7x6.npz")
    d = np.load(here)
    net = Connect4Net(board_width=7, board_height=6)
    net.load_state_dict({k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith("sd.")})
    p, v = net.predict(torch.from_numpy(d["input"]))
    assert np.abs(p.numpy() - d["probs"]).max() < 1e-6 and np.abs(v.numpy() - d["value"]).max() < 1e-6


@needs_ref
def test_tictactoe_classes_match_reference():
    """TicTacToeConfig defaults, class surface, network parameter names, outputs and the policy-vector helpers."""
    import_reference()
    import alphazero.games.tictactoe as r_ttt
    import alphazero.games.registers as r_reg
    import alphazero_b200.games.tictactoe as ttt
    import alphazero_b200.games.registers as reg
    assert dataclasses.asdict(ttt.TicTacToeConfig()) == dataclasses.asdict(r_ttt.TicTacToeConfig())
    assert [t.value for t in reg.DATA_AUGMENT_STRATEGIES["tictactoe"].rotations] == \
        [t.value for t in r_reg.DATA_AUGMENT_STRATEGIES["tictactoe"].rotations]
    for ref_cls, cls in ((r_ttt.TicTacToeBoard, ttt.TicTacToeBoard), (r_ttt.TicTacToeNet, ttt.TicTacToeNet)):
        for name, fn in _public_methods(ref_cls).items():
            if name in ("human_display", "pixel_display", "get_alignments_sums", "get_nb_free_cells_on_alignments"):
                continue      # rendering / NumPy helpers of the reference's rule code
            assert hasattr(cls, name), name
            ref_params = list(inspect.signature(fn).parameters)
            assert list(inspect.signature(getattr(cls, name)).parameters)[:len(ref_params)] == ref_params, name
        assert list(inspect.signature(cls.__init__).parameters) == list(inspect.signature(ref_cls.__init__).parameters)
    torch.manual_seed(0)
    net, ref = ttt.TicTacToeNet(), r_ttt.TicTacToeNet()
    assert list(net.state_dict().keys()) == list(ref.state_dict().keys())
    ref.load_state_dict(net.state_dict())
    assert net.get_parameters_count() == ref.get_parameters_count()
    x = torch.randint(-1, 2, (11, 3, 3)).float()
    p, v = net.predict(x)
    rp, rv = ref.predict(x)
    assert torch.equal(p, rp) and torch.equal(v, rv)
    p1, _ = net.predict(x[0])                       # a single 3x3 input gets a batch dimension (tictactoe.py:305-306)
    assert p1.shape == (1, 9)
    probs = np.random.RandomState(0).rand(9).astype(np.float32)
    moves = [(0, 1), (2, 2), (1, 0)]
    assert net.get_normalized_probs(probs, moves) == ref.get_normalized_probs(probs, moves)
    assert net.get_normalized_probs(np.zeros(9, np.float32), moves) == {m: 1 / 3 for m in moves}
    mp = {(0, 1): 0.25, (2, 2): 0.75}
    assert (net.to_neural_output(mp) == ref.to_neural_output(mp)).all()
    for axis in (0, 1):
        assert (net.reflect_neural_output(probs, axis) == ref.reflect_neural_output(probs, axis)).all()
    for angle in (90, 180, 270):
        assert (net.rotate_neural_output(probs, angle) == ref.rotate_neural_output(probs, angle)).all()
    with pytest.raises(ValueError):
        net.reflect_neural_output(np.zeros(10), 1)
