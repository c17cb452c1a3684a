#!/usr/bin/env python
"""Generates tests/golden/*.json from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/Makefile).  Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

golden_m5.json      the reference's shipped m5 fixtures (test/cpp/basic.m5, parsequery.m5): every line, what the
                    reference's parseM5 makes of it (ref_parse_line), and the FASTA pbdagcon_ref prints for the
                    flag sets of SURVEY.md App. B.3 (config 1).
golden_targets.json small fuzzed targets (inputs stored verbatim) with the reference's normalizeGaps/trimAln
                    strings, graph-dump digests before/after mergeNodes and consensus fragments.
golden_synth.json   digests of the reference's FASTA on the first targets of the bench workloads (inputs are
                    regenerated from the seed by pbdagcon_b200/synth).
"""
import hashlib
import json
import os
import random
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
from helpers import Ref, rand_target, MODE_PIPELINE, MODE_DUMP_PRE, MODE_DUMP_POST  # noqa: E402

REF_TESTS = "/root/reference/test/cpp"


def md5(b):
    return hashlib.md5(b if isinstance(b, bytes) else b.encode()).hexdigest()


def canon_fasta(text):
    """Records sorted (the reference's -j>=2 output order is nondeterministic)."""
    lines = text.split("\n")
    recs = sorted(lines[i] + "\n" + lines[i + 1] + "\n" for i in range(0, len(lines) - 1, 2))
    return "".join(recs)


def ref_fasta(path, flags):
    r = subprocess.run([helpers.REF_BIN, "-j", "2"] + flags + [path], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True)
    return r.stdout.decode()


def golden_m5(ref):
    out = {"files": {}}
    for name in ("basic.m5", "parsequery.m5"):
        raw = open(os.path.join(REF_TESTS, name), "rb").read()
        lines = [ln for ln in raw.split(b"\n") if ln]
        parsed = []
        for ln in lines:
            p = ref.parse_line(ln, 0)
            parsed.append({k: (v.decode() if isinstance(v, bytes) else v) for k, v in p.items()})
        runs = []
        for flags in (["-c", "1", "-m", "1", "-t", "0"], ["-c", "1", "-m", "1", "-t", "50"], ["-c", "1", "-m", "100", "-t", "50"], []):
            fa = ref_fasta(os.path.join(REF_TESTS, name), flags)
            runs.append({"flags": flags, "fasta": fa, "md5": md5(fa)})
        out["files"][name] = {"lines": [ln.decode() for ln in lines], "parsed": parsed, "runs": runs}
    return out


def golden_targets(ref):
    rng = random.Random(20261104)
    cases = []
    for it in range(60):
        L, bb, alns = rand_target(rng, L=rng.randint(20, 160), n_aln=rng.randint(1, 9), err=rng.choice([0.05, 0.15, 0.3]),
                                  full=(it % 4 == 0), alphabet="ACGT" if it % 5 else "ACGTN", dots=(it % 7 == 0))
        trim, minw, minlen = rng.choice([0, 3, 10]), rng.choice([0, 1, 2, 4]), rng.choice([0, 5, 20])
        use_bb = it % 2 == 1
        r = ref.consensus(L, alns, backbone=bb if use_bb else None, mode=MODE_PIPELINE | MODE_DUMP_PRE | MODE_DUMP_POST,
                          trim=trim, min_weight=minw, min_len=minlen)
        norm = [ref.normalize_trim(q, t, s, trim) for q, t, s in alns]
        cases.append({
            "tlen": L, "backbone": bb.decode() if use_bb else None, "trim": trim, "min_weight": minw, "min_len": minlen,
            "alns": [[q.decode(), t.decode(), s] for q, t, s in alns],
            "norm": [[q.decode(), t.decode(), s] for q, t, s in norm],
            "dump_md5": md5(r["dump"]), "frags": [[a, b, s.decode()] for a, b, s in r["frags"]],
        })
    return {"cases": cases}


def golden_synth(ref):
    import pbdagcon_b200.synth as synth
    out = {"workloads": []}
    for seed, L, cov, nt in ((2, 8000, 30, 6), (4, 20000, 50, 2), (7, 3000, 20, 8)):
        b = synth.gen_batch(seed, L, cov, 0, nt)
        recs = []
        for t in range(nt):
            r = ref.consensus(L, b.target_alns(t), mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500)
            recs.append([[a, e, md5(s)] for a, e, s in r["frags"]])
        tmp = os.path.join("/tmp", f"golden_{seed}.m5")
        synth.write_m5(seed, L, cov, 0, nt, tmp)
        fa = canon_fasta(ref_fasta(tmp, ["-c", "8"]))
        os.unlink(tmp)
        out["workloads"].append({"seed": seed, "L": L, "cov": cov, "targets": nt, "flags": "-c 8 -m 500 -t 50",
                                 "frags": recs, "fasta_sorted_md5": md5(fa), "fasta_bases": sum(len(x) for x in fa.split("\n")[1::2])})
    return out


def main():
    assert helpers.have_ref() and os.path.exists(REF_TESTS), "needs oracle/_ref and /root/reference (build container only)"
    ref = Ref()
    for name, fn in (("golden_m5.json", golden_m5), ("golden_targets.json", golden_targets), ("golden_synth.json", golden_synth)):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(fn(ref), f, indent=1)
        print("wrote", name)


if __name__ == "__main__":
    main()
