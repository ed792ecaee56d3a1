"""ctypes binding of libaz_b200.so (C ABI declared in include/az_b200.h).

The library is built in-tree by `build_library()` (nvcc, sm_100a only).  Loading fails loudly: there is no CPU or
PyTorch fallback for any of the kernels.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libaz_b200.so")
CSRC = os.path.join(_HERE, "csrc")
SOURCES = ["board_kernels.cu", "connect4_kernels.cu", "tictactoe_kernels.cu", "tree_kernels.cu", "net_kernels.cu", "trunk_rows_kernels.cu", "trunk_rows2_kernels.cu", "fcstack_kernels.cu", "fcstack2_kernels.cu", "fcstack3_kernels.cu", "sample_kernels.cu", "player_kernels.cu"]

AZ_OK, AZ_ERR_ARG, AZ_ERR_CUDA, AZ_ERR_CAPACITY = 0, -1, -2, -3
AZ_PASS = 64
LEAF_IDLE, LEAF_NN, LEAF_PRIME, LEAF_ROLLOUT, LEAF_DRAW, LEAF_BLACK, LEAF_WHITE = range(7)
STATUS_OVERFLOW, STATUS_DEPTH, STATUS_ILLEGAL = 1, 2, 4
MODE_ROLLOUT, MODE_NEURAL = 0, 1
GAME_OTHELLO, GAME_CONNECT4, GAME_TICTACTOE = 0, 1, 2
NODE_EVAL, NODE_EXPANDED, NODE_NOISED = 1, 2, 4

c_void_p, c_int, c_i64, c_u64, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_size_t


class AzTree(ctypes.Structure):
    """Mirror of `struct az_tree` (include/az_b200.h); field order and types must match exactly."""
    _fields_ = [(name, c_void_p) for name in (
        "nodes", "half", "n_alloc", "root_own", "root_opp", "root_player", "active", "sims_done", "rng_ctr",
        "status", "leaf_node", "leaf_kind", "leaf_own", "leaf_opp", "leaf_player", "path", "path_len", "enc",
        "probs", "value")] + [
        ("seed", ctypes.c_uint64),
        ("G", ctypes.c_int32), ("n", ctypes.c_int32), ("capacity", ctypes.c_int32), ("max_depth", ctypes.c_int32),
        ("mode", ctypes.c_int32), ("tie_random", ctypes.c_int32), ("enc_dtype", ctypes.c_int32),
        ("dirichlet_alpha", ctypes.c_float), ("dirichlet_eps", ctypes.c_float), ("game", ctypes.c_int32),
        ("height", ctypes.c_int32), ("logits_ld", ctypes.c_int32), ("logits", c_void_p),
        ("n_leaves", ctypes.c_int32), ("reserved", ctypes.c_int32),
    ]


# symbol -> (restype, argtypes); every function declared in include/az_b200.h must appear here
SIGNATURES = {
    "az_abi_version": (c_int, []),
    "az_last_cuda_error": (ctypes.c_char_p, []),
    "az_othello_legal": (c_int, [c_void_p, c_void_p, c_int, c_i64, c_void_p, c_void_p]),
    "az_othello_flips": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_i64, c_void_p, c_void_p]),
    "az_othello_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_i64, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p]),
    "az_othello_terminal_score": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_i64, c_void_p, c_void_p,
                                          c_void_p, c_void_p]),
    "az_othello_perft": (c_int, [c_u64, c_u64, c_int, c_int, c_int, c_void_p, c_size_t,
                                 ctypes.POINTER(ctypes.c_uint64), c_void_p]),
    "az_othello_playouts": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_i64, c_i64, c_u64, c_u64, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "az_othello_encode": (c_int, [c_void_p, c_void_p, c_int, c_i64, c_void_p, c_int, c_void_p]),
    "az_connect4_legal": (c_int, [c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p]),
    "az_connect4_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p]),
    "az_connect4_terminal_score": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p,
                                           c_void_p, c_void_p]),
    "az_connect4_playouts": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_i64, c_u64, c_u64, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p]),
    "az_connect4_encode": (c_int, [c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_int, c_void_p]),
    "az_connect4_perft": (c_int, [c_u64, c_u64, c_int, c_int, c_int, c_int, c_void_p, c_size_t,
                                  ctypes.POINTER(ctypes.c_uint64), c_void_p]),
    "az_tictactoe_legal": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "az_tictactoe_apply": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "az_tictactoe_terminal_score": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p,
                                            c_void_p]),
    "az_tictactoe_encode": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_int, c_void_p]),
    "az_tictactoe_playouts": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_i64, c_u64, c_u64, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p]),
    "az_tictactoe_net_floats": (c_int, []),
    "az_tictactoe_net": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "az_tree_reset": (c_int, [ctypes.POINTER(AzTree), c_void_p, c_void_p]),
    "az_tree_step": (c_int, [ctypes.POINTER(AzTree), c_int, c_int, c_int, c_void_p]),
    "az_tree_rollout": (c_int, [ctypes.POINTER(AzTree), c_void_p]),
    "az_tree_root_stats": (c_int, [ctypes.POINTER(AzTree), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "az_tree_choose": (c_int, [ctypes.POINTER(AzTree), c_void_p, c_void_p, c_void_p, c_void_p]),
    "az_tree_advance": (c_int, [ctypes.POINTER(AzTree), c_void_p, c_void_p]),
    "az_tree_move": (c_int, [ctypes.POINTER(AzTree), ctypes.c_float] + [c_void_p] * 9),
    "az_eval_hash": (c_int, [c_void_p, c_void_p, c_int, c_i64, c_u64, c_void_p, c_void_p, c_void_p]),
    "az_greedy_move": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_u64, ctypes.c_uint32, c_int,
                                c_void_p, c_void_p, c_void_p]),
    "az_net_wblob_bytes": (c_int, []),
    "az_net_trunk_boards_per_group": (c_int, [c_int, c_i64]),
    "az_net_trunk": (c_int, [c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "az_net_fcstack_plan": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "az_net_fcstack_debug": (c_int, [c_void_p]),
    "az_net_fcstack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "az_net_fcstack_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_i64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "az_connect4_image_bits": (c_int, [c_void_p, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p, c_void_p]),
    "az_net_heads": (c_int, [c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p, c_void_p]),
    "az_net_headgemm_blob_bytes": (c_i64, [c_int, c_int]),
    "az_net_headgemm": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_i64, c_void_p, c_void_p, c_void_p]),
    "az_augment_scratch_bytes": (c_i64, [c_i64]),
    "az_augment_twins": (c_int, [c_int]),
    "az_augment_plan": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p]),
    "az_augment_apply": (c_int, [c_void_p] * 6 + [c_i64, c_int, c_int, c_int, c_int] + [c_void_p] * 9),
    "az_samples_pack": (c_int, [c_void_p] * 7 + [c_int, c_i64, c_int] + [c_void_p] * 7),
    "az_samples_batch": (c_int, [c_void_p] * 5 + [c_i64, c_i64, c_int, c_int, c_int] + [c_void_p] * 5),
}

_lib = None


def build_library(verbose: bool = False) -> str:
    """Compiles csrc/*.cu into libaz_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-o", LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    return LIB_PATH


def library_is_stale() -> bool:
    if not os.path.isfile(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(_ROOT, "include", "az_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.isfile(d))


def lib() -> ctypes.CDLL:
    """Returns the loaded library; raises RuntimeError if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "alphazero_b200 has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


LAUNCH_CALLS = 0  # number of kernel-launching C-ABI calls made through check() (bench.py's gpu_launches)


def check(rc: int, what: str = "") -> None:
    global LAUNCH_CALLS
    LAUNCH_CALLS += 1
    if rc == AZ_OK:
        return
    if rc == AZ_ERR_ARG:
        raise ValueError(f"{what}: invalid argument")
    if rc == AZ_ERR_CAPACITY:
        raise RuntimeError(f"{what}: scratch capacity exceeded")
    msg = lib().az_last_cuda_error().decode()
    raise RuntimeError(f"{what}: CUDA error: {msg}")
