"""GPU suite, dazcon (BASELINE configs[2]): the drop-in `dazcon` binary -- LAS/DB provider on the host, consensus on
the GPU -- against `dazcon_ref -j 1`, the header-exact oracle: byte-identical FASTA on the
committed fixture (goldens made from the unmodified reference) and, where oracle/_ref travelled, against the
reference binary live, also on a larger pair generated on the box with the dazzler tools."""
import os
import re
import subprocess

import pytest

import daz_helpers as dh

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(dh.DAZCON), reason="dazcon front not built")]
REFDIR = os.path.join(dh.ROOT, "oracle", "_ref")


def run(exe, args, timeout=600, env=None):
    r = subprocess.run([exe] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout, env=env)
    assert r.returncode == 0, r.stderr.decode()[-500:]
    return r.stdout


def test_dazcon_fixture_goldens():
    for g in dh.gold()["runs"]:
        out = run(dh.DAZCON, ["-j", "4", "-s", str(dh.DB), "-a", str(dh.LAS)] + g["flags"])
        lines = out.split(b"\n")
        recs = [[lines[i].decode(), dh.md5(lines[i + 1])] for i in range(0, len(lines) - 1, 2)]
        assert recs == g["records"], g["name"]
        assert len(out) == g["bytes"] and dh.md5(out) == g["md5"], g["name"]


@pytest.mark.skipif(not os.path.exists(dh.DAZCON_REF), reason="oracle/_ref/dazcon_ref did not travel")
def test_dazcon_fixture_live_against_reference_binary():
    for flags in (["-c", "8", "-t", "20"], ["-x", "-m", "12", "-l", "1000"]):
        ref = run(dh.DAZCON_REF, ["-j", "1", "-s", str(dh.DB), "-a", str(dh.LAS)] + flags)
        for extra in ([], ["--batch-targets", "7", "-j", "3"]):
            assert run(dh.DAZCON, ["-s", str(dh.DB), "-a", str(dh.LAS)] + flags + extra) == ref, (flags, extra)
    # several handles (shared device on a one-GPU box): same bytes, same well counter
    env = dict(os.environ, DGC_SHARE_GPUS="1")
    one = run(dh.DAZCON, ["-s", str(dh.DB), "-a", str(dh.LAS), "--batch-targets", "9"])
    assert run(dh.DAZCON, ["-s", str(dh.DB), "-a", str(dh.LAS), "--batch-targets", "9", "--gpus", "3"], env=env) == one


@pytest.mark.skipif(not os.path.exists(os.path.join(REFDIR, "daligner")), reason="dazzler tools did not travel")
def test_dazcon_generated_pair_against_reference(tmp_path):
    """~460 reads x 10 kb at 25x made on the box by the dazzler simulator and daligner; `dazcon_ref -j N` prints the records
    in a nondeterministic order with a per-thread counter, so records are compared with the counter field masked."""
    def tool(name, args, **kw):
        subprocess.run([os.path.join(REFDIR, name)] + args, cwd=tmp_path, check=True, **kw)
    with open(tmp_path / "reads.fasta", "wb") as f:
        tool("simulator", ["0.2", "-c25", "-m10000", "-s2000", "-x4000", "-e.15", "-r3"], stdout=f)
    tool("fasta2DB", ["reads", "reads.fasta"])
    tool("DBsplit", ["-x500", "-s200", "reads"])
    tool("daligner", ["reads", "reads"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    parts = sorted(p for p in os.listdir(tmp_path) if p.startswith("reads.reads.") and p.endswith(".las"))
    tool("LAsort", parts)
    tool("LAmerge", ["reads.las"] + [p[:-4] + ".S.las" for p in parts])
    db, las = str(tmp_path / "reads.db"), str(tmp_path / "reads.las")
    nthr = str(min(32, os.cpu_count() or 4))
    ref = run(dh.DAZCON_REF, ["-j", nthr, "-s", db, "-a", las], timeout=1200)
    out = run(dh.DAZCON, ["-j", nthr, "-s", db, "-a", las])

    def canon(text):
        lines = text.split(b"\n")
        return sorted(re.sub(rb"^(>[^/]+)/\d+/", rb"\1/*/", lines[i]) + b"\n" + lines[i + 1] for i in range(0, len(lines) - 1, 2))
    assert len(canon(out)) > 400 and canon(out) == canon(ref)
    # our own counter is the running record index in input order
    heads = [ln for ln in out.split(b"\n") if ln.startswith(b">")]
    assert [int(h.split(b"/")[1]) for h in heads] == list(range(len(heads)))
    ids = [int(h.split(b"/")[0][1:]) for h in heads]
    assert ids == sorted(ids)


def test_device_realignment_strings_equal_host_provider():
    """the gapped strings the DEVICE builds from the trace points (k_daz_iter = iter_np per segment,
    k_daz_emit = decodeAlignment) against the host provider, which runs the dazzler library's own Compute_Trace_PTS and
    is itself pinned to the reference provider (tests/test_daz_cpu.py): every overlap, byte for byte."""
    h = dh.host()
    for kw in (dict(), dict(min_cov=2, max_hits=6, sort_cov=1, proper=1)):
        dev = h.dump_device(dh.DB, dh.LAS, **kw)
        assert not dev.startswith(b"ERROR"), dev[:200]
        host = dh.parse_dump(h.dump(str(dh.DB), str(dh.LAS), **kw))
        got = dh.parse_dump(b"\n".join(ln + b" ." if ln.startswith(b"T ") else ln for ln in dev.split(b"\n")))
        assert len(got) == len(host) and len(host) > 100
        n = 0
        for (hid, hl, _, ha), (gid, gl, _, ga) in zip(host, got):
            assert (hid, hl) == (gid, gl)
            assert ha == ga, f"target {hid}"
            n += len(ha)
        assert n > 500


def test_dazcon_device_and_host_realignment_print_the_same_fasta():
    a = run(dh.DAZCON, ["-j", "4", "-s", str(dh.DB), "-a", str(dh.LAS)])
    b = run(dh.DAZCON, ["-j", "4", "-s", str(dh.DB), "-a", str(dh.LAS), "--host-realign"])
    assert a == b and a.count(b">") > 100
