"""The `-a` step on the host (pbdagcon_b200/host/simple_align.cpp) against the reference's one known-answer vector
(reference test/cpp/SimpleAlignerTest.cpp:8-21) and against properties every alignment must have.  The arithmetic
the reference uses here lives in blasr_libcpp, whose source is absent: parity beyond the vector is unpinned."""
import ctypes as C
import os
import random

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_SO = os.path.join(ROOT, "pbdagcon_b200", "host", "libdagcon_host.so")

COMP = bytes.maketrans(b"ACGT", b"TGCA")


def revcomp(s):
    return s.translate(COMP)[::-1]


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOST_SO):
        import __graft_entry__ as ge
        ge.build_host()
    h = C.CDLL(HOST_SO)
    h.dgh_simple_align.restype = C.c_int
    h.dgh_simple_align.argtypes = [C.c_char_p, C.c_char_p, C.c_char, C.c_uint32, C.POINTER(C.c_uint32),
                                   C.POINTER(C.c_uint32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    h.dgh_free.argtypes = [C.c_void_p]
    return h


def align(host, q, t, strand, tlen, start, end):
    s, e = C.c_uint32(start), C.c_uint32(end)
    qo, to = C.c_void_p(), C.c_void_p()
    if not host.dgh_simple_align(q, t, strand, tlen, C.byref(s), C.byref(e), C.byref(qo), C.byref(to)):
        return None
    try:
        return C.string_at(qo), C.string_at(to), s.value, e.value
    finally:
        host.dgh_free(qo)
        host.dgh_free(to)


def test_reference_known_answer(host):
    # SimpleAlignerTest.cpp:8-21, verbatim inputs and expectations
    t = b"ACAGAGATGCAAGGTAAAGTACAATTGAAAAACTAACCTCTTCCAGCGAGACTTATAGCGA"
    q = b"ACAGAAGATGAAGGTAAATACAATGAAAAAACTACCTCGGTTCCAGCGAGAACTATAGCGA"
    qs, ts, start, end = align(host, q, t, b"-", 2092, 765, 826)
    assert ts == b"TCGCTATAAGT-CTCGCTGGAA--GAGGTTAGTTTTT-CAATTGTACTTTACCTTGCATCT-CTGT"
    assert start == 1267
    assert len(qs) == len(ts)
    assert qs.replace(b"-", b"") == revcomp(q)


def mutate(rng, seg, err):
    out = []
    for b in seg:
        u = rng.random()
        if u < err * 0.73:
            out.append(rng.choice(b"ACGT")); out.append(b)      # insertion before the base
        elif u < err * 0.93:
            continue                                            # deletion
        elif u < err:
            out.append(rng.choice([c for c in b"ACGT" if c != b]))
        else:
            out.append(b)
    return bytes(out)


@pytest.mark.parametrize("n,err", [(40, 0.0), (200, 0.1), (1500, 0.15), (6000, 0.15), (3000, 0.3)])
def test_alignment_properties(host, n, err):
    rng = random.Random(1000 + n)
    for strand in (b"+", b"-"):
        t = bytes(rng.choice(b"ACGT") for _ in range(n))
        q = mutate(rng, t, err)
        tlen, start0 = n + 500, 100
        qs, ts, start, end = align(host, q, t, strand, tlen, start0, start0 + n)
        assert len(qs) == len(ts) and len(qs) > 0
        assert all(a != ord("-") or b != ord("-") for a, b in zip(qs, ts))
        assert qs[0] == ts[0] and qs[-1] == ts[-1]            # local: starts and ends on a match
        fq, ft = (qs, ts) if strand == b"+" else (revcomp(qs), revcomp(ts))
        tq, tt = fq.replace(b"-", b""), ft.replace(b"-", b"")
        tb = t.find(tt)
        assert tq in q and tb >= 0
        te = tb + len(tt)
        # SimpleAligner.cpp:49-61
        if strand == b"+":
            assert start == start0 + tb + 1 and end == start0 + tb + te
        else:
            assert end == start0 + tb + te and start == tlen - end + 1
        # a sane alignment of a 15 % error read: most columns match
        same = sum(a == b for a, b in zip(qs, ts))
        assert same >= (1.0 - 1.6 * err) * len(tt) - 5
        if err == 0.0:
            assert qs == ts == (t if strand == b"+" else revcomp(t))


def test_band_widens_for_long_indels(host):
    rng = random.Random(7)
    a = bytes(rng.choice(b"ACGT") for _ in range(1200))
    ins = bytes(rng.choice(b"ACGT") for _ in range(400))
    b = bytes(rng.choice(b"ACGT") for _ in range(300))
    q, t = a + ins + b, a + b                                  # one 400-base insertion: off any 64-wide band
    qs, ts, start, end = align(host, q, t, b"+", 5000, 10, 10 + len(t))
    assert qs.replace(b"-", b"") == q and ts.replace(b"-", b"") == t
    assert ts.count(b"-") == 400 and qs.count(b"-") == 0
    assert start == 11 and end == 10 + len(t)


def test_nothing_to_align(host):
    assert align(host, b"AAAAAAAA", b"CCCCCCCC", b"+", 100, 1, 9) is None
    assert align(host, b"", b"ACGT", b"+", 100, 1, 5) is None


def test_unrelated_sequences_stay_bounded(host):
    # no band is ever clear of the path here: the search must stop at the widest band its memory bound allows and
    # still hand back a well-formed (if meaningless) alignment, or nothing
    rng = random.Random(99)
    q = bytes(rng.choice(b"ACGT") for _ in range(6000))
    t = bytes(rng.choice(b"ACGT") for _ in range(5000))
    r = align(host, q, t, b"+", 9000, 1, 5001)
    if r is not None:
        qs, ts, start, end = r
        assert len(qs) == len(ts)
        assert qs.replace(b"-", b"") in q and ts.replace(b"-", b"") in t
