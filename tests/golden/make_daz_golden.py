#!/usr/bin/env python
"""Generates the dazcon fixture and its goldens FROM THE UNMODIFIED REFERENCE, in this container:
  tests/golden/daz/{small.db,small.idx,small.bps,small.las}   a synthetic DAZZ_DB + daligner LAS pair
      (DAZZ_DB simulator | fasta2DB; DBsplit; daligner; LAsort; LAmerge -- SURVEY.md 8(d) cfg 3, built by oracle/Makefile)
  tests/golden/golden_dazcon.json   per flag set: FASTA md5 + (header, md5(seq)) of `dazcon_ref -j 1` (the header-exact
      oracle) and the md5 of the reference DazAlnProvider's dump (oracle/_ref/libdaz_ref.so)
Run from the repo root after `make -C oracle ref`.  Test infrastructure only."""
import ctypes as C
import hashlib
import json
import os
import shutil
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "oracle", "_ref")
OUT = os.path.join(ROOT, "tests", "golden", "daz")

RUNS = [
    {"name": "defaults", "flags": []},
    {"name": "proper_covsort", "flags": ["-o", "-x", "-c", "4", "-l", "300", "-t", "5", "-m", "20"]},
    {"name": "targets", "flags": ["-c", "3", "3", "17", "100", "138"]},
]
DUMPS = [(6, 85, 0, 0, []), (6, 85, 1, 1, []), (3, 10, 1, 0, []), (4, 85, 0, 0, [2, 50, 138]), (1, 2, 0, 1, [])]


def ref_dump(db, las, mc, mh, sc, pr, tg):
    lib = C.CDLL(os.path.join(REF, "libdaz_ref.so"))
    lib.dazref_dump.restype = C.c_void_p
    lib.dazref_dump.argtypes = [C.c_char_p, C.c_char_p, C.c_uint, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]
    lib.dazref_text.restype = C.c_void_p
    lib.dazref_text.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.dazref_free.argtypes = [C.c_void_p]
    arr = (C.c_int * len(tg))(*tg)
    h = lib.dazref_dump(db.encode(), las.encode(), mc, mh, sc, pr, arr, len(tg))
    n = C.c_uint64()
    s = C.string_at(lib.dazref_text(h, C.byref(n)), n.value)
    lib.dazref_free(h)
    return s


def main():
    os.makedirs(OUT, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        def run(cmd, **kw):
            subprocess.run(cmd, cwd=tmp, check=True, **kw)
        with open(os.path.join(tmp, "small.fasta"), "wb") as f:
            run([os.path.join(REF, "simulator"), "0.03", "-c25", "-m5000", "-s1000", "-x2000", "-e.15", "-r7"], stdout=f)
        run([os.path.join(REF, "fasta2DB"), "small", "small.fasta"])
        run([os.path.join(REF, "DBsplit"), "-x500", "-s200", "small"])
        run([os.path.join(REF, "daligner"), "small", "small"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        parts = sorted(p for p in os.listdir(tmp) if p.startswith("small.small.") and p.endswith(".las"))
        run([os.path.join(REF, "LAsort")] + parts)
        run([os.path.join(REF, "LAmerge"), "small.las"] + [p[:-4] + ".S.las" for p in parts])
        # the DB's hidden companions are committed under visible names (snapshots of the tree skip dot files); the tests
        # put them back as .small.idx / .small.bps in a scratch directory (tests/daz_helpers.py: fixture_dir)
        for f, g in (("small.db", "small.db"), (".small.idx", "small.idx"), (".small.bps", "small.bps"), ("small.las", "small.las")):
            shutil.copy(os.path.join(tmp, f), os.path.join(OUT, g))
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import daz_helpers
    db, las = daz_helpers.fixture_paths()
    gold = {"recipe": "simulator 0.03 -c25 -m5000 -s1000 -x2000 -e.15 -r7 | fasta2DB; DBsplit -x500 -s200; daligner; LAsort; LAmerge",
            "runs": [], "dumps": []}
    for r in RUNS:
        out = subprocess.run([os.path.join(REF, "dazcon_ref"), "-j", "1", "-s", db, "-a", las] + r["flags"],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout
        lines = out.split(b"\n")
        recs = [[lines[i].decode(), hashlib.md5(lines[i + 1]).hexdigest()] for i in range(0, len(lines) - 1, 2)]
        gold["runs"].append({"name": r["name"], "flags": r["flags"], "md5": hashlib.md5(out).hexdigest(), "bytes": len(out), "records": recs})
    for mc, mh, sc, pr, tg in DUMPS:
        s = ref_dump(db, las, mc, mh, sc, pr, tg)
        gold["dumps"].append({"min_cov": mc, "max_hits": mh, "sort_cov": sc, "proper": pr, "targets": tg, "md5": hashlib.md5(s).hexdigest(),
                              "bytes": len(s), "n_targets": s.count(b"\nT ") + s.startswith(b"T ")})
    json.dump(gold, open(os.path.join(ROOT, "tests", "golden", "golden_dazcon.json"), "w"), indent=0)
    print({r["name"]: (r["md5"], len(r["records"])) for r in gold["runs"]})


if __name__ == "__main__":
    main()
