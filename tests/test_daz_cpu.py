"""The dazcon front on the host (no GPU): LAS reader, pile / hit bookkeeping, hit scoring and the trace-point
realignment of pbdagcon_b200/host/daz_provider.cpp against (1) the reference's own unit vectors
(reference test/cpp/TargetHitTest.cpp:4-77, test/cpp/TargetTest.cpp:4-22), (2) the committed goldens made from the
unmodified reference DazAlnProvider (tests/golden/make_daz_golden.py) and (3) that provider itself, live, where
oracle/_ref was built."""
import os
import subprocess

import pytest

import daz_helpers as dh

pytestmark = pytest.mark.skipif(not os.path.exists(dh.DAZCON), reason="dazcon front not built (dazzler sources absent at build time)")


def test_target_hit_score_kats():
    # reference test/cpp/TargetHitTest.cpp: single_overlap_perfect, single_overlap_inaccurate, multi_overlap_inaccurate
    checkers = [dh.host()] + ([dh.ref()] if dh.have_dazref() else [])
    for c in checkers:
        assert c.ovl_score([(27, 7013, 1231, 8217, 0)]) == pytest.approx(6986, rel=1e-6)
        assert c.ovl_score([(10, 5000, 2000, 6000, 230)]) == pytest.approx(3770, rel=1e-6)
        assert c.ovl_score([(10, 5000, 2000, 6000, 230), (5005, 7005, 6001, 7995, 53)]) == pytest.approx(4721, rel=1e-6)
    if len(checkers) == 2:   # bit-identical floats on a sweep of chains, incl. overlapping records that replace / drop
        import random
        rng = random.Random(4)
        chains = []
        for _ in range(300):
            chain, a = [], rng.randint(0, 50)
            for _ in range(rng.randint(1, 5)):
                la = rng.randint(100, 4000)
                chain.append((a, a + la, rng.randint(0, 5000), rng.randint(5000, 9000), rng.randint(0, la // 3)))
                a = max(a + rng.randint(-300, la + 200), 0)
            chains.append(chain)
        assert checkers[0].ovl_scores(chains) == checkers[1].ovl_scores(chains)


def test_target_add_record_grouping():
    # reference test/cpp/TargetTest.cpp:4-22 (listed as broken in the reference's makefile; the grouping rule itself holds)
    import ctypes as C
    h = dh.host()
    recs = [1, 3, 0, 1, 3, 0, 2, 3, 1]
    assert h.lib.dgh_daz_group_hits((C.c_int * 9)(*recs), 3) == 2


def test_provider_dump_matches_reference_goldens():
    """Every (targSeq, alignments) the provider yields, for five option sets (min-cov, max-hits, -x, -o, target list):
    md5 equal to the reference provider's dump recorded in golden_dazcon.json."""
    h = dh.host()
    for d in dh.gold()["dumps"]:
        s = h.dump(str(dh.DB), str(dh.LAS), d["min_cov"], d["max_hits"], d["sort_cov"], d["proper"], d["targets"])
        assert len(s) == d["bytes"] and dh.md5(s) == d["md5"], d
        assert len(dh.parse_dump(s)) == d["n_targets"]


@pytest.mark.skipif(not dh.have_dazref(), reason="oracle/_ref/libdaz_ref.so not built")
def test_provider_dump_matches_reference_live():
    h, r = dh.host(), dh.ref()
    for kw in (dict(), dict(min_cov=2, max_hits=5, sort_cov=1), dict(proper=1, min_cov=1), dict(targets=(1, 2, 77, 137, 138), min_cov=3)):
        a, b = h.dump(str(dh.DB), str(dh.LAS), **kw), r.dump(str(dh.DB), str(dh.LAS), **kw)
        assert a == b, kw
    t = dh.parse_dump(h.dump(str(dh.DB), str(dh.LAS)))
    assert len(t) > 100 and all(len(seq) == tlen for _, tlen, seq, _ in t)
    assert all(len(q) == len(tt) for _, _, _, alns in t for _, _, q, tt in alns)


def test_missing_inputs_fail_loudly(tmp_path):
    r = subprocess.run([dh.DAZCON, "-a", str(dh.LAS)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "seq-file" in r.stderr
    r = subprocess.run([dh.DAZCON, "-s", str(dh.DB)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "align-file" in r.stderr
    h = dh.host()
    assert h.dump(str(dh.DB), str(tmp_path / "nope.las")).startswith(b"ERROR Open failed")
