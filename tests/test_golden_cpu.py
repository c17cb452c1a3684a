"""CPU suite: the oracle against the committed golden vectors (generated from the unmodified reference by
tests/golden/make_golden.py), and the C++ host boundary (m5 parser, scanner/packer, provider, CLI) without a GPU."""
import ctypes as C
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from helpers import Oracle, MODE_PIPELINE, MODE_DUMP_PRE, MODE_DUMP_POST

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
HOST_DIR = os.path.join(ROOT, "pbdagcon_b200", "host")


def md5(b):
    return hashlib.md5(b if isinstance(b, bytes) else b.encode()).hexdigest()


def load(name):
    return json.load(open(os.path.join(GOLD, name)))


# ------------------------------------------------------------------ oracle vs golden
def test_oracle_matches_golden_targets():
    o = Oracle()
    for i, c in enumerate(load("golden_targets.json")["cases"]):
        alns = [(q.encode(), t.encode(), s) for q, t, s in c["alns"]]
        bb = c["backbone"].encode() if c["backbone"] is not None else None
        for (q, t, s), exp in zip(alns, c["norm"]):
            got = o.normalize_trim(q, t, s, c["trim"])
            assert [got[0].decode(), got[1].decode(), got[2]] == exp, f"case {i}"
        r = o.consensus(c["tlen"], alns, backbone=bb, mode=MODE_PIPELINE | MODE_DUMP_PRE | MODE_DUMP_POST, trim=c["trim"],
                        min_weight=c["min_weight"], min_len=c["min_len"])
        assert md5(r["dump"]) == c["dump_md5"], f"case {i}: graph dump differs from the reference"
        assert [[a, b, s.decode()] for a, b, s in r["frags"]] == c["frags"], f"case {i}"


def test_oracle_matches_golden_synth():
    import pbdagcon_b200.synth as synth
    o = Oracle()
    for w in load("golden_synth.json")["workloads"]:
        b = synth.gen_batch(w["seed"], w["L"], w["cov"], 0, w["targets"])
        for t in range(w["targets"]):
            r = o.consensus(w["L"], b.target_alns(t), mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500)
            assert [[a, e, md5(s)] for a, e, s in r["frags"]] == w["frags"][t], f"seed {w['seed']} target {t}"


# ------------------------------------------------------------------ host boundary
@pytest.fixture(scope="module")
def host():
    import __graft_entry__ as ge
    ge.build_product()
    ge.build_host()
    lib = C.CDLL(os.path.join(HOST_DIR, "libdagcon_host.so"))
    lib.dgh_pack_file.restype = C.c_void_p
    lib.dgh_pack_file.argtypes = [C.c_char_p, C.c_int, C.c_uint32, C.c_int, C.c_uint32, C.c_uint64]
    lib.dgh_pack_error.restype = C.c_char_p
    lib.dgh_pack_error.argtypes = [C.c_void_p]
    lib.dgh_pack_counts.restype = C.c_uint32
    lib.dgh_pack_counts.argtypes = [C.c_void_p, C.c_int]
    lib.dgh_pack_arena_bytes.restype = C.c_uint64
    lib.dgh_pack_arena_bytes.argtypes = [C.c_void_p]
    lib.dgh_pack_array.restype = C.c_void_p
    lib.dgh_pack_array.argtypes = [C.c_void_p, C.c_int]
    lib.dgh_pack_free.argtypes = [C.c_void_p]
    lib.dgh_parse_line.argtypes = [C.c_char_p, C.c_uint64, C.c_int] + [C.POINTER(C.c_uint32)] * 3 + [C.c_char_p] + [C.POINTER(C.c_void_p)] * 4
    lib.dgh_free.argtypes = [C.c_void_p]
    lib.dgh_provider_groups.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_uint32), C.c_int]
    return lib


def parse_line(lib, line, fmt=0):
    tlen, start, end = C.c_uint32(), C.c_uint32(), C.c_uint32()
    strand = C.create_string_buffer(2)
    ptrs = [C.c_void_p() for _ in range(4)]
    rc = lib.dgh_parse_line(line, len(line), fmt, C.byref(tlen), C.byref(start), C.byref(end), strand, *[C.byref(p) for p in ptrs])
    if rc != 1:
        return rc
    vals = [C.string_at(p.value) for p in ptrs]
    for p in ptrs:
        lib.dgh_free(p)
    return {"id": vals[0], "sid": vals[1], "qstr": vals[2], "tstr": vals[3], "tlen": tlen.value, "start": start.value,
            "end": end.value, "strand": strand.raw[0]}


def unpack(lib, h):
    T, A = lib.dgh_pack_counts(h, 0), lib.dgh_pack_counts(h, 1)

    def arr(which, n, dt):
        p = lib.dgh_pack_array(h, which)
        return np.frombuffer(C.string_at(p, n * np.dtype(dt).itemsize), dtype=dt).copy() if n else np.zeros(0, dt)

    nb = lib.dgh_pack_arena_bytes(h)
    return {"tlen": arr(0, T, np.uint32), "aln_begin": arr(1, T + 1, np.uint32), "start": arr(2, A, np.uint32),
            "len": arr(3, A, np.uint32), "q_off": arr(4, A, np.uint64), "t_off": arr(5, A, np.uint64), "arena": arr(6, nb, np.uint8),
            "ids": C.string_at(lib.dgh_pack_array(h, 7)).decode().split("\n")[:-1],
            "batch_targets": arr(8, lib.dgh_pack_counts(h, 2), np.uint32), "dropped": lib.dgh_pack_counts(h, 3)}


def test_m5_parser_matches_reference_fixtures(host):
    """parseM5 of the reference's shipped fixtures (AlignmentTest.cpp:49-78 ParseBasic/ParseQuery use the same files)."""
    g = load("golden_m5.json")
    for name, f in g["files"].items():
        for line, exp in zip(f["lines"], f["parsed"]):
            got = parse_line(host, line.encode())
            for k in ("id", "sid", "qstr", "tstr"):
                assert got[k].decode() == exp[k], (name, k)
            assert got["tlen"] == exp["tlen"] and got["start"] == exp["start"] and got["strand"] == exp["strand"]
    # ParseBasic's own assertions (AlignmentTest.cpp:49-66)
    a = parse_line(host, g["files"]["basic.m5"]["lines"][0].encode())
    assert a["start"] >= 1 and len(a["qstr"]) == len(a["tstr"])


def test_parser_edge_cases(host):
    assert parse_line(host, b"") == 0
    assert parse_line(host, b"   ") == 0
    assert parse_line(host, b"a b c") == -1
    pre = b"q1 t1 - 100 5 50 ACGT ACGA"
    p = parse_line(host, pre, fmt=1)
    assert (p["sid"], p["id"], p["strand"], p["tlen"], p["start"], p["end"], p["qstr"], p["tstr"]) == \
        (b"q1", b"t1", ord("-"), 100, 5, 50, b"ACGT", b"ACGA")   # parsePre: no +1, no revComp (Alignment.cpp:95-111)
    m5 = b"q/0_4  4 0 4 + t 10 2 6 - -20 4 0 0 0 254 AC-GN |||*| ACTGT"   # double space, '-' strand, non-ACGT passes through
    p = parse_line(host, m5)
    assert p["start"] == 3 and p["tlen"] == 10 and p["qstr"] == b"NC-GT" and p["tstr"] == b"ACAGT"


def test_pack_file_matches_generator(host, tmp_path):
    import pbdagcon_b200.synth as synth
    path = str(tmp_path / "s.m5")
    synth.write_m5(7, 3000, 20, 0, 9, path)
    b = synth.gen_batch(7, 3000, 20, 0, 9)
    for threads, bt in ((1, 4), (3, 2), (2, 1000)):
        h = host.dgh_pack_file(path.encode(), 0, 8, threads, bt, 0)
        assert host.dgh_pack_error(h) == b""
        u = unpack(host, h)
        host.dgh_pack_free(h)
        assert u["ids"] == [f"synth/{t}/0_3000" for t in range(9)]
        assert list(u["tlen"]) == [3000] * 9 and list(u["aln_begin"]) == list(b.tgt_aln_begin)
        assert list(u["start"]) == list(b.aln_start) and list(u["len"]) == list(b.aln_len)
        assert int(u["batch_targets"].sum()) == 9 and (bt >= 9 or len(u["batch_targets"]) > 1)
        for k in range(b.n_alns):
            n = int(b.aln_len[k])
            assert u["arena"][int(u["q_off"][k]):int(u["q_off"][k]) + n].tobytes() == b.arena[int(b.aln_q_off[k]):int(b.aln_q_off[k]) + n].tobytes()
            assert u["arena"][int(u["t_off"][k]):int(u["t_off"][k]) + n].tobytes() == b.arena[int(b.aln_t_off[k]):int(b.aln_t_off[k]) + n].tobytes()
            assert int(u["q_off"][k]) % 16 == 0 and int(u["t_off"][k]) % 16 == 0


def test_reader_filter_and_errors(host, tmp_path):
    import pbdagcon_b200.synth as synth
    path = str(tmp_path / "s.m5")
    synth.write_m5(3, 1500, 6, 0, 3, path)   # 10 alignments per target
    h = host.dgh_pack_file(path.encode(), 0, 11, 2, 0, 0)   # Reader drops groups below minCov (main.cpp:66-72)
    assert host.dgh_pack_counts(h, 0) == 0 and host.dgh_pack_counts(h, 3) == 3
    host.dgh_pack_free(h)
    txt = open(path).read()
    open(path, "w").write(txt + "\n\n")                      # trailing blank lines are harmless
    h = host.dgh_pack_file(path.encode(), 0, 1, 2, 0, 0)
    assert host.dgh_pack_error(h) == b"" and host.dgh_pack_counts(h, 0) == 3
    host.dgh_pack_free(h)
    lines = txt.split("\n")
    open(path, "w").write("\n".join(lines[:5]) + "\n\n" + "\n".join(lines[5:]))   # a blank line mid-stream is refused
    h = host.dgh_pack_file(path.encode(), 0, 1, 2, 0, 0)
    assert b"blank line" in host.dgh_pack_error(h)
    host.dgh_pack_free(h)
    h = host.dgh_pack_file(str(tmp_path / "missing.m5").encode(), 0, 1, 2, 0, 0)
    assert b"Error opening file" in host.dgh_pack_error(h)
    host.dgh_pack_free(h)


def test_provider_grouping(host, tmp_path):
    """BlasrM5AlnProvider::nextTarget call pattern (BlasrM5AlnProvider.cpp:34-55): the first call yields an empty
    vector, then one target per call; the last call returns false but still fills dest."""
    import pbdagcon_b200.synth as synth
    path = str(tmp_path / "s.m5")
    synth.write_m5(3, 1500, 6, 0, 3, path)
    sizes = (C.c_uint32 * 16)()
    n = host.dgh_provider_groups(path.encode(), 0, sizes, 16)
    assert n == 4 and list(sizes[:4]) == [0, 10, 10, 10]


def test_cli_flags_and_no_cpu_fallback(host, tmp_path):
    import torch
    exe = os.path.join(HOST_DIR, "pbdagcon")
    r = subprocess.run([exe, "--help"], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0 and all(f in r.stdout for f in ("--threads", "--min-coverage", "--min-length", "--trim", "--align"))
    assert subprocess.run([exe], stderr=subprocess.PIPE).returncode == 1
    assert subprocess.run([exe, "-q", "x"], stderr=subprocess.PIPE).returncode == 1
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    p = tmp_path / "x.m5"
    p.write_text("")
    r = subprocess.run([exe, "-j", "2", "-c", "8", str(p)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr and r.stdout == ""
