"""CPU: the device headers (othello_bits.cuh / philox.cuh are __host__ __device__) compiled for the host and
checked against the oracle -- catches bit-logic bugs before a GPU is involved.  Not a product path."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hb():
    so = os.path.join(ROOT, "tests", "_host_bits.so")
    subprocess.run(["g++", "-O2", "-x", "c++", "-shared", "-fPIC", "-o", so,
                    os.path.join(ROOT, "tests", "host_bits_harness.cpp")], check=True)
    L = ctypes.CDLL(so)
    L.hb_nth.argtypes = [ctypes.c_uint64, ctypes.c_int]
    return L


@pytest.mark.parametrize("n", [4, 6, 8])
def test_bitboard_logic_vs_oracle(hb, n):
    P = orc.P
    rng = np.random.default_rng(n)
    G = 4000
    b0, w0 = orc.start(n)
    black, white = np.full(G, b0, np.uint64), np.full(G, w0, np.uint64)
    player = np.ones(G, np.int8)
    checked = 0
    for ply in range(n * n + 4):
        legal = orc.legal(black, white, player, n)
        own, opp = np.where(player > 0, black, white), np.where(player > 0, white, black)
        l2 = np.zeros(G, np.uint64)
        hb.hb_legal(P(own), P(opp), n, ctypes.c_int64(G), P(l2))
        assert (legal == l2).all()
        term, _, _ = orc.terminal(black, white, player, n)
        t2 = np.zeros(G, np.uint8)
        hb.hb_terminal(P(own), P(opp), n, ctypes.c_int64(G), P(t2))
        assert (term == t2).all()
        act = np.full(G, 64, np.uint8)
        for g in range(G):
            m = int(legal[g])
            if m:
                bits = [i for i in range(64) if (m >> i) & 1]
                act[g] = bits[rng.integers(len(bits))]
        ob, ow, op, fl, st = orc.apply(black, white, player, act, n)
        live = term == 0
        assert (st[live] == 0).all()
        f2 = np.zeros(G, np.uint64)
        hb.hb_flips(P(own), P(opp), P(act), ctypes.c_int64(G), P(f2))
        assert (fl[live] == f2[live]).all()
        o2, p2 = own.copy(), opp.copy()
        hb.hb_apply(P(o2), P(p2), P(act), ctypes.c_int64(G))
        nb, nw = np.where(op > 0, o2, p2), np.where(op > 0, p2, o2)
        assert (nb[live] == ob[live]).all() and (nw[live] == ow[live]).all()
        black, white, player = np.where(live, ob, black), np.where(live, ow, white), np.where(live, op, player)
        checked += int(live.sum())
    assert checked > G * (n * n - 8)


def test_nth_set_bit_and_philox(hb):
    rng = np.random.default_rng(0)
    for _ in range(3000):
        m = int(rng.integers(1, 2**63)) | (int(rng.integers(0, 2)) << 63)
        bits = [i for i in range(64) if (m >> i) & 1]
        k = int(rng.integers(len(bits)))
        assert hb.hb_nth(m, k) == bits[k]
    for _ in range(100):
        c = (ctypes.c_uint32 * 4)(*[int(x) for x in rng.integers(0, 2**32, 4)])
        k = (ctypes.c_uint32 * 2)(*[int(x) for x in rng.integers(0, 2**32, 2)])
        o1, o2 = (ctypes.c_uint32 * 4)(), (ctypes.c_uint32 * 4)()
        hb.hb_philox(c, k, o1)
        orc.lib().orc_philox4x32_10(c, k, o2)
        assert list(o1) == list(o2)
