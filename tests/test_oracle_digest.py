"""CPU: the oracle's boards against the digests of ~1 M positions per game computed by the unmodified reference
(tests/golden/reference_digests.json, made by tests/golden/make_golden_digest.py): legal-move sets, game-over flags,
scores, winners and -- through the walk itself -- every flip of every played move."""
import json
import os

import numpy as np
import pytest

import oracle_lib as orc
from digest_walk import walk_digest

BACKENDS = {
    "othello8": (orc.start(8), lambda b, w, p: orc.legal(b, w, p, 8), lambda b, w, p, a: orc.apply(b, w, p, a, 8)[:3],
                 lambda b, w, p: orc.terminal(b, w, p, 8)),
    "othello6": (orc.start(6), lambda b, w, p: orc.legal(b, w, p, 6), lambda b, w, p, a: orc.apply(b, w, p, a, 6)[:3],
                 lambda b, w, p: orc.terminal(b, w, p, 6)),
    "connect4_7x6": ((0, 0), lambda b, w, p: orc.c4_legal(b, w, p, 7, 6), lambda b, w, p, a: orc.c4_apply(b, w, p, a, 7, 6)[:3],
                     lambda b, w, p: orc.c4_terminal(b, w, p, 7, 6)),
}


@pytest.mark.parametrize("key", sorted(BACKENDS))
def test_oracle_reproduces_the_reference_digests(golden_dir, key):
    d = json.load(open(os.path.join(golden_dir, "reference_digests.json")))["games"][key]
    start, legal_fn, apply_fn, terminal_fn = BACKENDS[key]
    total = 0
    for ch in d["chunks"]:
        sha, n = walk_digest(ch["seed"], ch["games"], start, legal_fn, apply_fn, terminal_fn)
        assert n == ch["positions"] and sha == ch["sha256"], (key, ch["seed"])
        total += n
    assert total == d["positions"] and total >= 1_000_000
