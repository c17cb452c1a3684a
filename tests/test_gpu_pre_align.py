"""`pbdagcon -a` on the device: k_pre_align (pbdagcon_b200/csrc/pre_align.cuh) against the host aligner
(pbdagcon_b200/host/simple_align.cpp), bit for bit -- gapped strings and Alignment::start -- including the reference's
one known-answer vector (test/cpp/SimpleAlignerTest.cpp:8-21), records that need wider bands, records that do not
align, and the CLI's two paths printing the same FASTA.  Parity with blasr_libcpp (the reference's own arithmetic for
this step) is unpinned: its source is not part of the reference tree."""
import os
import random
import subprocess

import pytest

from test_simple_align_cpu import align, mutate, revcomp, HOST_SO, host  # noqa: F401  (host is a fixture)

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "pbdagcon_b200", "host", "pbdagcon")


@pytest.fixture(scope="module")
def api():
    import pbdagcon_b200.api as api
    return api


def host_expect(host, q, t, strand, tlen, start):
    r = align(host, q, t, strand, tlen, start, start + len(t))
    return (b"", b"", start + 1) if r is None else (r[0], r[1], r[2])


def test_reference_known_answer_on_device(api, host):
    t = b"ACAGAGATGCAAGGTAAAGTACAATTGAAAAACTAACCTCTTCCAGCGAGACTTATAGCGA"
    q = b"ACAGAAGATGAAGGTAAATACAATGAAAAAACTACCTCGGTTCCAGCGAGAACTATAGCGA"
    eng = api.Consensus(min_cov=1, min_len=1, trim=0)
    got = eng.debug_pre_strings(api.PreBatch([(2092, [(q, t, "-", 765)])]))
    eng.close()
    assert got[0][1] == b"TCGCTATAAGT-CTCGCTGGAA--GAGGTTAGTTTTT-CAATTGTACTTTACCTTGCATCT-CTGT"
    assert got[0][2] == 1267
    assert got[0][0].replace(b"-", b"") == revcomp(q)


def test_device_alignments_equal_host_aligner(api, host):
    rng = random.Random(77)
    targets, expect = [], []
    for ti in range(20):
        L = rng.choice([400, 1500, 3000, 7000])
        tgt = bytes(rng.choice(b"ACGT") for _ in range(L))
        recs = []
        for r in range(10):
            span = rng.randint(L // 3, L)
            s = rng.randint(0, L - span)
            seg = tgt[s:s + span]
            kind = (ti * 10 + r) % 9
            q = mutate(rng, seg, rng.choice([0.05, 0.15, 0.25]))
            if kind == 5:      # a long insertion in the read: the path leaves the 64-column band and the band doubles
                k = len(q) // 2
                q = q[:k] + bytes(rng.choice(b"ACGT") for _ in range(rng.choice([90, 200, 600]))) + q[k:]
            if kind == 6:      # a long deletion
                k = len(q) // 3
                q = q[:k] + q[k + rng.choice([100, 300]):]
            if kind == 7:      # unrelated sequence
                q = bytes(rng.choice(b"ACGT") for _ in range(len(seg)))
            if kind == 8 and ti % 2:   # empty query segment
                q = b""
            if r % 2:
                rec = (revcomp(q), revcomp(seg), "-", L - (s + span))
            else:
                rec = (q, seg, "+", s)
            recs.append(rec)
            expect.append(host_expect(host, rec[0], rec[1], rec[2].encode(), L, rec[3]))
        targets.append((L, recs))
    eng = api.Consensus(min_cov=1, min_len=1, trim=0)
    got = eng.debug_pre_strings(api.PreBatch(targets))
    eng.close()
    assert len(got) == len(expect) == 200
    bad = [k for k, (g, e) in enumerate(zip(got, expect)) if g != e]
    assert not bad, (bad[:5], got[bad[0]][2], expect[bad[0]][2], len(got[bad[0]][0]), len(expect[bad[0]][0]))
    assert sum(1 for e in expect if e[0]) > 150 and sum(1 for e in expect if not e[0]) >= 1


def test_cli_device_and_host_align_print_the_same_fasta(api, tmp_path):
    from test_gpu_golden import _pre_workload
    lines, truth = _pre_workload(23, 16, 2500, 20)
    pre = str(tmp_path / "y.pre")
    open(pre, "w").write("\n".join(lines) + "\n")
    outs = []
    for extra in ([], ["--host-align"], ["--batch-targets", "3"]):
        r = subprocess.run([EXE, "-a", "-c", "8", "-j", "4"] + extra + [pre], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        outs.append(r.stdout)
    assert outs[0] == outs[1] == outs[2] and outs[0].count(">") >= 16
