"""Differential test: the C oracle against the compiled, unmodified reference (oracle/_ref)
on fuzzed targets -- normalizeGaps/trimAln strings, the graph before and after mergeNodes
(every adjacency list in order, every count/visited flag) and the consensus fragments."""
import random

import pytest

from helpers import (Oracle, Ref, have_ref, rand_target, MODE_PIPELINE, MODE_DUMP_PRE, MODE_DUMP_POST,
                     MODE_NORM, MODE_MERGE, MODE_SINGLE)

pytestmark = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (reference tree absent)")


def test_normalize_trim_fuzz():
    o, r = Oracle(), Ref()
    rng = random.Random(1)
    for it in range(3000):
        n = rng.randint(1, 120)
        alpha = "ACGT-" if it % 3 else "ACGTN-."
        q = "".join(rng.choice(alpha) for _ in range(n)).encode()
        t = "".join(rng.choice(alpha) for _ in range(n)).encode()
        trim = rng.choice([0, 1, 2, 5, 10, 50, 500])
        start = rng.randint(1, 1000)
        assert o.normalize_trim(q, t, start, trim) == r.normalize_trim(q, t, start, trim), (q, t, trim)


@pytest.mark.parametrize("seed", range(6))
def test_graph_and_consensus_fuzz(seed):
    o, r = Oracle(), Ref()
    rng = random.Random(100 + seed)
    for it in range(150):
        full = it % 4 == 0
        L, bb, alns = rand_target(rng, err=rng.choice([0.05, 0.15, 0.3]), full=full,
                                  alphabet="ACGT" if it % 5 else "ACGTN", dots=(it % 7 == 0))
        trim = rng.choice([0, 3, 10])
        minw = rng.choice([0, 1, 2, 4])
        minlen = rng.choice([0, 5, 20])
        use_bb = bb if it % 2 else None
        mode = MODE_PIPELINE | MODE_DUMP_PRE | MODE_DUMP_POST
        a = o.consensus(L, alns, backbone=use_bb, mode=mode, trim=trim, min_weight=minw, min_len=minlen)
        b = r.consensus(L, alns, backbone=use_bb, mode=mode, trim=trim, min_weight=minw, min_len=minlen)
        assert a["status"] == 0
        assert a["dump"] == b["dump"], f"graph differs seed={seed} it={it}"
        assert a["frags"] == b["frags"]
        m2 = MODE_NORM | MODE_MERGE | MODE_SINGLE
        assert o.consensus(L, alns, backbone=use_bb, mode=m2, min_weight=minw)["single"] == \
            r.consensus(L, alns, backbone=use_bb, mode=m2, min_weight=minw)["single"]
