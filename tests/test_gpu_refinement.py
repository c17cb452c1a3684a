"""BASELINE configs[4] at reduced count: a 40 kb backbone at 100x, three rounds of iterative refinement (the consensus
of round r is the backbone of round r+1, AlnGraphBoost(const std::string&) form).  Between rounds the reads are
re-aligned to the new backbone by tests/harness_align.c, a deterministic banded aligner shared by both sides, so
parity is checked per round on identical alignments.  40 kb also exceeds the build kernel's
shared-memory accumulators, so this is the test of its global-accumulator path."""
import ctypes as C
import os
import subprocess

import pytest

from helpers import Oracle, MODE_PIPELINE

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def hal(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("hal") / "libhal.so")
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(HERE, "harness_align.c")], check=True)
    lib = C.CDLL(so)
    lib.hal_align.restype = C.c_int64
    lib.hal_align.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, C.c_int64, C.c_char_p, C.c_char_p]

    def align(q, t, band):
        qo, to = C.create_string_buffer(len(q) + len(t) + 1), C.create_string_buffer(len(q) + len(t) + 1)
        n = lib.hal_align(q, len(q), t, len(t), band, qo, to)
        assert n >= 0, "band too narrow"
        return qo.raw[:n], to.raw[:n]
    return align


def _position_map(align, old_bb, new_bb):
    """old backbone position -> new backbone position (monotone), from a global alignment of the two."""
    a, b = align(old_bb, new_bb, 4096)
    m, i, j = [0] * (len(old_bb) + 1), 0, 0
    for x, y in zip(a, b):
        if x != 0x2D:
            m[i] = j
            i += 1
        if y != 0x2D:
            j += 1
    m[len(old_bb)] = len(new_bb)
    return m


def test_iterative_refinement_40kb_100x(hal):
    import pbdagcon_b200.api as api
    import pbdagcon_b200.synth as synth
    L, cov, rounds = 40000, 100, 3
    sb = synth.gen_batch(seed=5, L=L, cov=cov, t0=0, nt=1)
    alns = sb.target_alns(0)
    # the reads and the span of the (implicit) round-1 backbone each of them covers
    reads, bb = [], bytearray(b"N" * L)
    for q, t, start in alns:
        tb = t.replace(b"-", b"")
        bb[start - 1:start - 1 + len(tb)] = tb
        reads.append((q.replace(b"-", b""), start - 1, start - 1 + len(tb)))
    o = Oracle()
    backbone, cur_bb, cur_L = None, bytes(bb), L
    lengths = []
    for r in range(rounds):
        eng = api.Consensus(min_cov=8, min_len=500, trim=50)
        res = eng(api.PackedBatch.from_synth(synth.from_python([(cur_L, backbone, alns)], with_backbone=backbone is not None)))
        eng.close()
        exp = o.consensus(cur_L, alns, backbone=backbone, mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500)
        assert res.status[0] == api.T_OK
        assert res.frags(0) == exp["frags"], f"round {r + 1}"
        frag = max(res.frags(0), key=lambda f: len(f[2]))[2]
        assert len(frag) > 0.9 * L
        lengths.append(len(frag))
        if r + 1 == rounds:
            break
        # the consensus becomes the backbone; re-align every read to the window its old span maps to
        pm = _position_map(hal, cur_bb, frag)
        new_alns, new_reads = [], []
        for seq, s, e in reads:
            ns, ne = pm[s], pm[e]
            if ne - ns < 600:
                continue
            qa, ta = hal(seq, frag[ns:ne], 256)
            new_alns.append((qa, ta, ns + 1))
            new_reads.append((seq, ns, ne))
        assert len(new_alns) > 0.9 * len(alns)
        alns, reads, backbone, cur_bb, cur_L = new_alns, new_reads, frag, frag, len(frag)
    assert max(lengths) - min(lengths) < 0.05 * L   # the ends erode by a few hundred bases per round (-t 50, -c 8)
