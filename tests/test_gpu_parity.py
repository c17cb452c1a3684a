"""GPU parity: libdagcon_b200.so (through its C ABI) against the oracle, stage by stage and end to end.
Bit-exact is the bar: strings, vertex/edge lists in order, counts, flags, fragment ranges and sequences."""
import os
import random

import numpy as np
import pytest

from helpers import (Oracle, Ref, have_ref, checker, rand_target, MODE_PIPELINE, MODE_DUMP_PRE, MODE_DUMP_POST, MODE_NORM,
                     MODE_TRIM, MODE_MERGE, MODE_SINGLE)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import pbdagcon_b200.api as api
    return api


@pytest.fixture(scope="module")
def synth():
    import pbdagcon_b200.synth as synth
    return synth


def _batch(api, synth, targets, with_backbone=False):
    return api.PackedBatch.from_synth(synth.from_python(targets, with_backbone=with_backbone))


def test_normalize_trim_kats(api, synth):
    # reference test/cpp/AlignmentTest.cpp:19-47 and :80-143
    eng = api.Consensus(min_cov=0, min_len=0, trim=0)
    alns = [(b"CAC", b"CGC", 1), (b"-C--CGT", b"CCGAC-T", 1), (b"ATAT-AGCCGGC", b"ATATTA---GGC", 1)]
    got = eng.debug_normalize(_batch(api, synth, [(20, None, alns)]))
    assert got[0][:2] == (b"C-AC", b"CG-C")
    assert got[1][:2] == (b"CCG--T", b"CCGACT")
    assert got[2][:2] == (b"ATAT-AGCCGGC", b"ATATTAG--G-C")
    eng.close()
    t, q = b"ACG-TCA-GCA", b"AC-C-C-T---"
    for k, exp in [(0, (q, t, 1)), (3, (b"C-C-T", b"-TCA-", 4)), (4, (b"C", b"C", 5)), (5, (b"", b"", 6))]:
        eng = api.Consensus(min_cov=0, min_len=0, trim=k, flags=api.F_NO_NORMALIZE)
        assert eng.debug_normalize(_batch(api, synth, [(20, None, [(q, t, 1)])]))[0] == exp
        eng.close()
    eng = api.Consensus(min_cov=0, min_len=0, trim=500, flags=api.F_NO_NORMALIZE)
    assert eng.debug_normalize(_batch(api, synth, [(20, None, [(q, t, 1)])]))[0][:2] == (b"", b"")
    eng.close()


@pytest.mark.parametrize("trim", [0, 1, 2, 5, 10, 50, 500])
def test_normalize_trim_fuzz(api, synth, trim):
    o = checker()
    rng = random.Random(1000 + trim)
    alns = []
    for it in range(4000):
        n = rng.randint(1, 150) if it % 10 else rng.randint(150, 2500)
        alpha = "ACGT-" if it % 3 else "ACGTN-."
        if it % 7 == 0:
            alpha = "AC-"          # long gap runs, many pulls
        q = "".join(rng.choice(alpha) for _ in range(n)).encode()
        t = "".join(rng.choice(alpha) for _ in range(n)).encode()
        alns.append((q, t, rng.randint(1, 1000)))
    eng = api.Consensus(min_cov=0, min_len=0, trim=trim)
    got = eng.debug_normalize(_batch(api, synth, [(100000, None, alns)]))
    eng.close()
    for (q, t, s), g in zip(alns, got):
        assert g == o.normalize_trim(q, t, s, trim), (q, t, trim)


def test_normalize_min_len_filter(api, synth):
    eng = api.Consensus(min_cov=0, min_len=10, trim=0)
    got = eng.debug_normalize(_batch(api, synth, [(50, None, [(b"ACGT", b"ACGT", 1), (b"ACGTACGTACGT", b"ACGTACGTACGT", 1)])]))
    eng.close()
    assert got[0] is None and got[1] == (b"ACGTACGTACGT", b"ACGTACGTACGT", 1)


def _fuzz_targets(seed, n, **kw):
    rng = random.Random(seed)
    out = []
    for it in range(n):
        full = it % 4 == 0
        L, bb, alns = rand_target(rng, err=rng.choice([0.05, 0.15, 0.3]), full=full,
                                  alphabet="ACGT" if it % 5 else "ACGTN", dots=(it % 7 == 0), **kw)
        out.append((L, bb, alns))
    return out


@pytest.mark.parametrize("seed", range(4))
def test_graph_dumps_fuzz(api, synth, seed):
    """Every vertex, both adjacency lists in order, counts and visited flags, before and after mergeNodes."""
    o = checker()
    targets = _fuzz_targets(200 + seed, 40)
    trim = [0, 3, 10, 0][seed]
    use_bb = seed % 2 == 1
    eng = api.Consensus(min_cov=0, min_len=0, trim=trim)
    b = _batch(api, synth, targets, with_backbone=use_bb)
    for ti, (L, bb, alns) in enumerate(targets):
        exp = o.consensus(L, alns, backbone=bb if use_bb else None, mode=MODE_PIPELINE | MODE_DUMP_PRE | MODE_DUMP_POST, trim=trim)
        lines = exp["dump"].split("\n")
        nl = (len(lines) - 1) // 2
        pre, post = "\n".join(lines[:nl]) + "\n", "\n".join(lines[nl:2 * nl]) + "\n"
        assert eng.debug_graph(b, ti, 0) == pre, f"pre-merge graph differs target {ti}"
        assert eng.debug_graph(b, ti, 1) == post, f"post-merge graph differs target {ti}"
    eng.close()


@pytest.mark.parametrize("seed", range(8))
def test_consensus_fuzz(api, synth, seed):
    o = Oracle()
    r = Ref() if have_ref() else None
    rng = random.Random(300 + seed)
    targets = _fuzz_targets(300 + seed, 150)
    trim = rng.choice([0, 3, 10])
    minw = rng.choice([0, 1, 2, 4])
    minlen = rng.choice([0, 5, 20])
    use_bb = seed % 2 == 0
    eng = api.Consensus(min_cov=minw, min_len=minlen, trim=trim)
    res = eng(_batch(api, synth, targets, with_backbone=use_bb))
    eng.close()
    for ti, (L, bb, alns) in enumerate(targets):
        if len(alns) < minw:
            assert res.status[ti] == api.T_SKIPPED and res.frags(ti) == []
            continue
        exp = o.consensus(L, alns, backbone=bb if use_bb else None, mode=MODE_PIPELINE, trim=trim, min_weight=minw, min_len=minlen)
        assert res.status[ti] == api.T_OK
        assert res.frags(ti) == exp["frags"], f"target {ti}"
        if r is not None:
            assert res.frags(ti) == r.consensus(L, alns, backbone=bb if use_bb else None, mode=MODE_PIPELINE, trim=trim,
                                                min_weight=minw, min_len=minlen)["frags"]


def test_raw_consensus_kat(api, synth):
    # reference test/cpp/AlnGraphBoostTest.cpp:11-47
    alns = [
        (b"ATAT-AGCCGGC", b"ATATTA---GGC", 1),
        (b"ATAT-ACGGC", b"ATATTA-GGC", 1),
        (b"ATCAT--CCGGC", b"AT-ATTA--GGC", 1),
        (b"ATAT-ACCGAG-", b"ATATTA--G-GC", 1),
        (b"ATAT-AGCCGGC", b"ATATTA---GGC", 1),
    ]
    eng = api.Consensus(min_cov=0, min_len=0, trim=0, flags=api.F_NO_NORMALIZE | api.F_NO_TRIM | api.F_SINGLE)
    res = eng(_batch(api, synth, [(9, b"ATATTAGGC", alns)], with_backbone=True))
    eng.close()
    assert [f[2] for f in res.frags(0)] == [b"ATATAGCCGGC"]


@pytest.mark.parametrize("flags_mode", [(1, MODE_TRIM | MODE_MERGE), (2, MODE_NORM | MODE_MERGE), (4, MODE_NORM | MODE_TRIM),
                                        (8, MODE_PIPELINE | MODE_SINGLE)])
def test_mode_flags(api, synth, flags_mode):
    flags, mode = flags_mode
    o = checker()
    targets = _fuzz_targets(77, 60)
    if flags & 1:   # feed already-normalised strings so that un-normalised addAln stays well defined
        targets = [(L, bb, [o.normalize_trim(q, t, s, 0) for q, t, s in alns]) for L, bb, alns in targets]
    eng = api.Consensus(min_cov=2, min_len=0, trim=3, flags=flags)
    res = eng(_batch(api, synth, targets))
    eng.close()
    for ti, (L, bb, alns) in enumerate(targets):
        if len(alns) < 2:
            continue
        exp = o.consensus(L, alns, mode=mode, trim=3, min_weight=2, min_len=0)
        if flags & 8:
            assert [f[2] for f in res.frags(ti)] == [exp["single"]], f"target {ti}"
        else:
            assert res.frags(ti) == exp["frags"], f"target {ti}"


def test_edge_cases(api, synth):
    o = Oracle()
    eng = api.Consensus(min_cov=1, min_len=0, trim=0)
    # empty batch
    res = eng(api.PackedBatch(np.zeros(0, np.uint32), np.zeros(1, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.uint32),
                              np.zeros(0, np.uint64), np.zeros(0, np.uint64), np.zeros(16, np.uint8)))
    assert res.n_targets == 0
    targets = [
        (10, None, []),                                              # no alignments -> skipped
        (10, None, [(b"ACGTACGTAC", b"ACGTACGTAC", 1)]),             # perfect full-length match
        (10, None, [(b"ACGTACGTACG", b"ACGTACGTACG", 1)]),           # runs one base off the backbone -> rejected
        (10, None, [(b"ACG", b"ACG", 0)]),                           # start 0 -> rejected
        (10, None, [(b"----", b"ACGT", 3)]),                         # deletions only
        (10, None, [(b"ACGT", b"----", 5)]),                         # insertions only
        (10, None, [(b"AC", b"AC", 9), (b"AC", b"AC", 1)]),          # touches both backbone ends
        (1, None, [(b"A", b"A", 1), (b"A", b"A", 1), (b"C", b"-", 1), (b"C", b"-", 2)]),
        (5, None, [(b"-", b"-", 1), (b".", b".", 2)]),               # all-gap columns normalise to empty alignments
    ]
    res = eng(_batch(api, synth, targets))
    eng.close()
    assert res.status[0] == api.T_SKIPPED
    assert res.status[2] == api.T_BAD_ALIGNMENT and res.frags(2) == []
    assert res.status[3] == api.T_BAD_ALIGNMENT
    for ti in (1, 4, 5, 6, 7, 8):
        L, _, alns = targets[ti]
        exp = o.consensus(L, alns, mode=MODE_PIPELINE, trim=0, min_weight=1, min_len=0)
        assert exp["status"] == 0
        assert res.status[ti] == api.T_OK
        assert res.frags(ti) == exp["frags"], f"target {ti}"


def test_min_len_and_min_cov_filters(api, synth):
    """main.cpp:66-72,118 (targets with < minCov alignments) and :132 (raw columns < minLen are skipped but still
    count toward minCov); fragments shorter than minLen are dropped (AlnGraphBoost.cpp:359,371)."""
    o = checker()
    targets = _fuzz_targets(55, 80, min_span=5)
    eng = api.Consensus(min_cov=3, min_len=60, trim=2)
    res = eng(_batch(api, synth, targets))
    eng.close()
    seen_skip = seen_frag = False
    for ti, (L, bb, alns) in enumerate(targets):
        if len(alns) < 3:
            assert res.status[ti] == api.T_SKIPPED
            seen_skip = True
            continue
        exp = o.consensus(L, alns, mode=MODE_PIPELINE, trim=2, min_weight=3, min_len=60)
        assert res.frags(ti) == exp["frags"]
        seen_frag |= bool(exp["frags"])
    assert seen_skip and seen_frag


@pytest.mark.parametrize("shape", [(2000, 10, 6), (8000, 30, 3), (20000, 50, 32), (40000, 100, 32)])
def test_synthetic_config_shapes(api, synth, shape):
    """PacBio-like synthetic targets of the BASELINE config shapes, pbdagcon flags -c 8 -m 500 -t 50."""
    L, cov, nt = shape
    o = checker()
    sb = synth.gen_batch(seed=2, L=L, cov=cov, t0=0, nt=nt)
    eng = api.Consensus(min_cov=8, min_len=500, trim=50)
    res = eng(api.PackedBatch.from_synth(sb))
    eng.close()
    # the checker runs on every host thread (ctypes releases the GIL): 32 targets of 40 kb x 100x are a minute of CPU otherwise
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 1)) as ex:
        exps = list(ex.map(lambda ti: o.consensus(L, sb.target_alns(ti), mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500), range(nt)))
    for ti in range(nt):
        assert res.status[ti] == api.T_OK
        assert res.frags(ti) == exps[ti]["frags"], f"target {ti}"
        assert sum(len(f[2]) for f in res.frags(ti)) > 0.4 * L


def test_chunking_does_not_change_results(api, synth):
    sb = synth.gen_batch(seed=9, L=1500, cov=12, t0=0, nt=96)
    pb = api.PackedBatch.from_synth(sb)
    base = None
    for scratch in [0, 3 << 20, 1 << 20]:
        eng = api.Consensus(min_cov=4, min_len=100, trim=20, scratch_bytes=scratch)
        res = eng(pb)
        tm = eng.timings()
        eng.close()
        got = [res.frags(t) for t in range(res.n_targets)]
        if scratch:
            assert tm["chunks"] > 1
        if base is None:
            base = got
            o = checker()
            for t in (0, 50, 95):
                assert got[t] == o.consensus(1500, sb.target_alns(t), mode=MODE_PIPELINE, trim=20, min_weight=4, min_len=100)["frags"]
        assert got == base


def test_async_submit_collect_order(api, synth):
    """dgc_submit/dgc_collect: two batches in flight, results come back in submit order and equal the synchronous call."""
    eng = api.Consensus(min_cov=4, min_len=100, trim=20)
    pbs = [api.PackedBatch.from_synth(synth.gen_batch(seed=20 + i, L=1200, cov=10, t0=0, nt=8)) for i in range(5)]
    sync = [[eng(pb).frags(t) for t in range(8)] for pb in pbs]
    got = []
    for i in range(4):
        eng.submit(pbs[i])
    with pytest.raises(api.DagconError):
        eng.submit(pbs[4])          # DGC_MAX_INFLIGHT reached
    r = eng.collect()
    got.append([r.frags(t) for t in range(8)])
    eng.submit(pbs[4])
    for i in range(4):
        r = eng.collect()
        got.append([r.frags(t) for t in range(8)])
    with pytest.raises(api.DagconError):
        eng.collect()
    eng.close()
    assert got == sync


def test_staged_upload_matches_resident_run(api, synth, monkeypatch):
    """dgc_submit stages a large batch: the arena travels in pieces while earlier pieces are normalised and built, and
    chunks close when the arena estimate or the scratch budget runs out.  Results must equal the resident dgc_run."""
    sb = synth.gen_batch(seed=31, L=1500, cov=12, t0=0, nt=120)
    pb = api.PackedBatch.from_synth(sb)
    eng = api.Consensus(min_cov=4, min_len=100, trim=20)
    base = eng(pb)
    want = [base.frags(t) for t in range(base.n_targets)]
    want_status = list(base.status)
    eng.close()
    monkeypatch.setenv("DGC_STAGE_MIN_BYTES", "1")
    for piece_bytes, scratch in [(1 << 19, 0), (1 << 18, 6 << 20), (1 << 21, 0)]:
        monkeypatch.setenv("DGC_STAGE_PIECE_BYTES", str(piece_bytes))
        eng = api.Consensus(min_cov=4, min_len=100, trim=20, scratch_bytes=scratch)
        for rep in range(3):            # the second and third pass reuse (and re-size) the slot's arenas
            eng.submit(pb)
            r = eng.collect()
            assert list(r.status) == want_status
            assert [r.frags(t) for t in range(r.n_targets)] == want, (piece_bytes, scratch, rep)
        tm = eng.timings()
        if scratch:
            assert tm["chunks"] > 1
        # two staged batches in flight at once
        eng.submit(pb); eng.submit(pb)
        for _ in range(2):
            r = eng.collect()
            assert [r.frags(t) for t in range(r.n_targets)] == want
        eng.close()


def test_pool_exhaustion_rebuilds_with_worst_case_pool(api, synth, monkeypatch):
    """A target whose list pool runs dry (high coverage: backbone in-lists beyond the 12 inline slots, merge survivors
    beyond 6) is rebuilt with a worst-case pool; DGC_TEST_TINY_POOL makes every first-pass pool too small."""
    o = checker()
    sb = synth.gen_batch(seed=41, L=1200, cov=80, t0=0, nt=12)
    pb = api.PackedBatch.from_synth(sb)
    want = [o.consensus(1200, sb.target_alns(t), mode=MODE_PIPELINE, trim=20, min_weight=8, min_len=100)["frags"] for t in range(12)]
    eng = api.Consensus(min_cov=8, min_len=100, trim=20)
    res = eng(pb)
    eng.close()
    assert list(res.status) == [api.T_OK] * 12 and [res.frags(t) for t in range(12)] == want
    monkeypatch.setenv("DGC_TEST_TINY_POOL", "1")
    monkeypatch.setenv("DGC_STAGE_MIN_BYTES", "1")
    monkeypatch.setenv("DGC_STAGE_PIECE_BYTES", str(1 << 20))
    eng = api.Consensus(min_cov=8, min_len=100, trim=20)
    res = eng.run(eng.upload(pb))                  # simple path
    assert list(res.status) == [api.T_OK] * 12 and [res.frags(t) for t in range(12)] == want
    assert eng.timings()["chunks"] >= 2            # the rebuild ran as a chunk of its own
    eng.submit(pb)                                 # staged path
    res = eng.collect()
    assert list(res.status) == [api.T_OK] * 12 and [res.frags(t) for t in range(12)] == want
    eng.close()
    monkeypatch.setenv("DGC_TEST_TINY_POOL", "0")
    eng = api.Consensus(min_cov=8, min_len=100, trim=20)   # the flag is process-wide: switch it off again
    eng.close()


def test_config2_shape_invariants_at_scale(api, synth):
    """BASELINE configs[1] shape (8 kb x 30x, -c 8 -m 500 -t 50) at 1500 targets: the resident run, the staged submit and
    a run squeezed into many chunks agree bit for bit, a second run reproduces the first, every target yields
    consensus, and a sample matches the oracle."""
    import hashlib
    nt, L = 1500, 8000
    pieces = [synth.gen_batch(seed=2, L=L, cov=30, t0=t0, nt=125) for t0 in range(0, nt, 125)]
    pb = api.PackedBatch.from_synth(synth.concat(pieces))

    def digest(res):
        h = hashlib.md5()
        for t in range(res.n_targets):
            h.update(bytes([int(res.status[t])]))
            for a, b, s in res.frags(t):
                h.update(b"%d_%d:" % (a, b) + s)
        return h.hexdigest()

    eng = api.Consensus(min_cov=8, min_len=500, trim=50)
    db = eng.upload(pb)
    r1 = eng.run(db)
    d1 = digest(r1)
    assert digest(eng.run(db)) == d1, "a second run over the same resident batch differs"
    eng.free(db)
    eng.submit(pb)
    assert digest(eng.collect()) == d1, "staged submit differs from the resident run"
    eng.close()
    eng = api.Consensus(min_cov=8, min_len=500, trim=50, scratch_bytes=2 << 30)
    r3 = eng(pb)
    assert eng.timings()["chunks"] >= 4 and digest(r3) == d1, "chunked run differs"
    eng.close()
    assert all(int(s) == api.T_OK for s in r1.status)
    assert all(sum(len(f[2]) for f in r1.frags(t)) > 0.5 * L for t in range(nt))
    o = checker()
    for t in range(0, nt, 97):
        alns = pieces[t // 125].target_alns(t % 125)
        assert r1.frags(t) == o.consensus(L, alns, mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500)["frags"], f"target {t}"


def test_skip_edge_and_enter_edge_notes(api, synth):
    """The build remembers a position's first skip edge (deletion length <= 6, slot <= 31) and the slot of the enter edge
    that starts at a vertex in spare accumulator bytes (graph_rec.cuh b2_edge).  Everything around those codes: deletions of
    lengths 1..9 leaving the same vertex in every order (the remembered one, others, lengths that do not fit the code),
    the same deletion after the vertex's out-list has outgrown its record (30 insertions first), alignments that start on
    the same vertex many times, start with a deletion, or consist of deletions only."""
    o = checker()
    rng = random.Random(4242)
    L = 60
    bb = bytes(rng.choice(b"ACGT") for _ in range(L))

    def aln(start, ops):
        """ops over target positions from `start` (1-based): 'm' match, ('d', k) delete k bases, ('i', base) insert."""
        q, t, p = b"", b"", start - 1
        for op in ops:
            if op == "m":
                q += bb[p:p + 1]; t += bb[p:p + 1]; p += 1
            elif op[0] == "d":
                q += b"-" * op[1]; t += bb[p:p + op[1]]; p += op[1]
            else:
                q += bytes([op[1]]); t += b"-"
        return (q, t, start)

    alns = []
    other = lambda p: rng.choice([c for c in b"ACGT" if c != bb[p]])
    # deletions of every length after vertex 10 (target positions 11..), in shuffled order, several times each
    lens = [1, 2, 3, 6, 7, 9, 1, 2, 7, 6, 3, 1, 9, 2, 1]
    rng.shuffle(lens)
    for k in lens:
        alns.append(aln(6, ["m"] * 5 + [("d", k)] + ["m"] * 8))
    # vertex 30: thirty insertions first (its out-list leaves the record), then the same two deletions again and again
    for _ in range(30):
        alns.append(aln(26, ["m"] * 5 + [("i", other(30))] + ["m"] * 6))
    for k in [2, 2, 5, 2, 5, 8, 2]:
        alns.append(aln(26, ["m"] * 5 + [("d", k)] + ["m"] * 6))
    # starts: the same start vertex many times, a start that is a deletion, deletions only, a start on vertex 1
    for _ in range(9):
        alns.append(aln(40, ["m"] * 10))
    alns.append(aln(40, [("d", 3)] + ["m"] * 6))
    alns.append(aln(41, [("d", 4)]))
    alns.append(aln(1, ["m"] * 12))
    alns.append(aln(1, ["m"] * 3 + [("d", 2)] + ["m"] * 4))
    rng.shuffle(alns)
    eng = api.Consensus(min_cov=1, min_len=1, trim=0, flags=api.F_NO_NORMALIZE)
    b = _batch(api, synth, [(L, None, alns)])
    exp = o.consensus(L, alns, mode=MODE_DUMP_PRE | MODE_DUMP_POST | MODE_MERGE, trim=0, min_weight=1, min_len=1)
    lines = exp["dump"].split("\n")
    nl = (len(lines) - 1) // 2
    pre, post = "\n".join(lines[:nl]) + "\n", "\n".join(lines[nl:2 * nl]) + "\n"
    assert eng.debug_graph(b, 0, 0) == pre
    assert eng.debug_graph(b, 0, 1) == post
    res = eng(b)
    eng.close()
    assert res.status[0] == api.T_OK and res.frags(0) == exp["frags"]


@pytest.mark.parametrize("A", [251, 252, 253, 254, 300])
def test_backbone_vertex_degree_edges(api, synth, A):
    """A backbone vertex whose out-list (and its successor's in-list) grows to A + 1 entries: every alignment matches
    position 8, inserts one base of its own and rejoins at position 9.  252 alignments are the most the byte-wide
    per-position list lengths of the build hold (254 entries); beyond that the dispatcher launches the 16-bit variant.
    The reference has no such boundary: every size must come out identical to it."""
    o = checker()
    rng = random.Random(A)
    L = 24
    bb = bytes(rng.choice(b"ACGT") for _ in range(L))
    alns = []
    for i in range(A):
        ins = rng.choice([c for c in b"ACGT" if c != bb[8]])      # never equal to the next target base: stays an insertion
        q = bb[4:8] + bytes([ins]) + bb[8:14]
        t = bb[4:8] + b"-" + bb[8:14]
        alns.append((q, t, 5))
    eng = api.Consensus(min_cov=1, min_len=1, trim=0)
    b = _batch(api, synth, [(L, None, alns)])
    exp = o.consensus(L, alns, mode=MODE_PIPELINE | MODE_DUMP_PRE | MODE_DUMP_POST, trim=0, min_weight=1, min_len=1)
    lines = exp["dump"].split("\n")
    nl = (len(lines) - 1) // 2
    pre, post = "\n".join(lines[:nl]) + "\n", "\n".join(lines[nl:2 * nl]) + "\n"
    assert f"8 {chr(bb[7])}" in pre and len(pre.split("\n")[8].split("|")[1].split()) == A + 1   # vertex 8 has A + 1 out-edges
    assert eng.debug_graph(b, 0, 0) == pre
    assert eng.debug_graph(b, 0, 1) == post
    res = eng(b)
    eng.close()
    assert res.status[0] == api.T_OK and res.frags(0) == exp["frags"]


def test_alignment_count_limit_surfaces(api, synth, tmp_path):
    """65 000 alignments per target is the documented limit of the 16-bit degrees (DGC_T_TOO_MANY_ALNS beyond): the
    largest accepted target equals the reference, one more is rejected with a status, and the CLI says so on stderr."""
    import subprocess
    o = checker()
    L = 60
    rng = random.Random(65)
    bb = bytes(rng.choice(b"ACGT") for _ in range(L))

    def alns(n):
        return [(bb[s:s + 6], bb[s:s + 6], s + 1) for s in ((7 * i) % (L - 6) for i in range(n))]
    eng = api.Consensus(min_cov=4, min_len=3, trim=1)
    res = eng(_batch(api, synth, [(L, None, alns(65000)), (L, None, alns(65001)), (L, None, alns(40))]))
    eng.close()
    assert list(res.status) == [api.T_OK, api.T_TOO_MANY_ALNS, api.T_OK]
    assert res.frags(1) == []
    for ti, n in ((0, 65000), (2, 40)):
        assert res.frags(ti) == o.consensus(L, alns(n), mode=MODE_PIPELINE, trim=1, min_weight=4, min_len=3)["frags"]
    # the same through the CLI: the rejected target is named on stderr, the others are printed
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pbdagcon_b200", "host", "pbdagcon")
    path = tmp_path / "many.m5"
    with open(path, "w") as f:
        for tid, n in (("big", 65001), ("ok", 40)):
            for q, t, s in alns(n):
                f.write(f"q 0 0 0 + {tid} {L} {s - 1} 0 + 0 0 0 0 0 0 {q.decode()} | {t.decode()}\n")
    r = subprocess.run([exe, "-c", "4", "-m", "3", "-t", "1", str(path)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0
    assert "target big rejected (status 3: too many alignments)" in r.stderr
    assert r.stdout.startswith(">ok/")


def test_staged_output_buffers_grow_across_chunks(api, synth, monkeypatch):
    """The staged path sizes the compacted-output buffers from its first piece.  Here the first pieces hold skipped and
    tiny targets, so the estimate falls far short, the scratch budget closes several chunks, and out_cons / out_frags
    must grow at a later close WITHOUT losing the results of the chunks already compacted (the copy has to be ordered
    behind the compaction kernels on the slot's stream)."""
    rng = random.Random(5)
    small = []
    for i in range(80):
        L, bb, alns = rand_target(rng, L=200, n_aln=(2 if i % 2 else 5), err=0.1)
        small.append((L, None, alns))
    head = synth.from_python(small)
    tail = synth.gen_batch(seed=33, L=1500, cov=12, t0=0, nt=90)
    pb = api.PackedBatch.from_synth(synth.concat([head, tail]))
    eng = api.Consensus(min_cov=4, min_len=50, trim=5)
    base = eng(pb)                                          # simple path: everything copied first, one plan
    want = [base.frags(t) for t in range(base.n_targets)]
    want_status = list(base.status)
    eng.close()
    assert want_status.count(api.T_SKIPPED) == 40 and sum(len(w) for w in want[80:]) >= 90
    monkeypatch.setenv("DGC_STAGE_MIN_BYTES", "1")
    monkeypatch.setenv("DGC_STAGE_PIECE_BYTES", str(1 << 17))
    for scratch in (2 << 20, 5 << 20):
        eng = api.Consensus(min_cov=4, min_len=50, trim=5, scratch_bytes=scratch)
        for rep in range(2):
            eng.submit(pb)
            r = eng.collect()
            assert eng.timings()["chunks"] >= 3
            assert list(r.status) == want_status
            assert [r.frags(t) for t in range(r.n_targets)] == want, (scratch, rep)
        eng.close()


def test_packed4_wire_format_equals_ascii(api, synth):
    """DGC_BATCH_PACKED4: the strings as 4-bit codes, two columns per byte.  Same normalised strings and the same
    consensus as the ASCII batch, on fuzzed targets with '.', 'N', odd lengths and strings at every byte phase; through
    the warp normaliser, the serial one (F_NO_NORMALIZE) and the staged upload."""
    targets = _fuzz_targets(901, 120) + [(L, bb, alns) for L, bb, alns in _fuzz_targets(902, 40, min_span=5) if alns]
    sb = synth.from_python(targets)
    asc = api.PackedBatch.from_synth(sb)
    pk = api.PackedBatch.from_synth_packed4(sb)
    assert pk is not None and pk.packed4 and pk.arena.size < 0.6 * asc.arena.size
    # the same packed strings at byte phases 1..3 (the device realigns to words itself)
    shifted = []
    for ph in (1, 2, 3):
        arena = np.concatenate([np.zeros(ph, np.uint8), pk.arena])
        shifted.append(api.PackedBatch(pk.tgt_tlen, pk.tgt_aln_begin, pk.aln_start, pk.aln_len, pk.aln_q_off + np.uint64(ph),
                                       pk.aln_t_off + np.uint64(ph), arena, packed4=True))
    for flags in (0, api.F_NO_NORMALIZE):
        eng = api.Consensus(min_cov=2, min_len=10, trim=3, flags=flags)
        want_norm = eng.debug_normalize(asc)
        want = eng(asc)
        for b in [pk] + shifted:
            assert eng.debug_normalize(b) == want_norm
            got = eng(b)
            assert list(got.status) == list(want.status)
            assert [got.frags(t) for t in range(got.n_targets)] == [want.frags(t) for t in range(want.n_targets)]
        eng.close()
    # a character outside "-ACGTN." cannot be packed: the packer says so (and sends ASCII)
    assert api.PackedBatch.from_synth_packed4(synth.from_python([(10, None, [(b"ACgT", b"ACGT", 1)])])) is None


def test_packed4_large_alignments_and_staged_upload(api, synth, monkeypatch):
    sb = synth.gen_batch(seed=61, L=6000, cov=14, t0=0, nt=40)      # several 1024-column tiles per alignment, look-ahead margins
    asc, pk = api.PackedBatch.from_synth(sb), api.PackedBatch.from_synth_packed4(sb)
    eng = api.Consensus(min_cov=4, min_len=200, trim=20)
    want = eng(asc)
    wf = [want.frags(t) for t in range(want.n_targets)]
    assert [eng(pk).frags(t) for t in range(want.n_targets)] == wf
    eng.close()
    monkeypatch.setenv("DGC_STAGE_MIN_BYTES", "1")
    monkeypatch.setenv("DGC_STAGE_PIECE_BYTES", str(1 << 18))
    eng = api.Consensus(min_cov=4, min_len=200, trim=20, scratch_bytes=64 << 20)
    eng.submit(pk)
    r = eng.collect()
    assert eng.timings()["chunks"] >= 2 and [r.frags(t) for t in range(r.n_targets)] == wf
    eng.close()
