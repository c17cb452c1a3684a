#!/usr/bin/env python
"""Per-source-line summary of an ncu capture: joins `ncu --page source --csv` (SASS rows, in program order) with
`nvdisasm --print-line-info` of the same cubin (same instruction order) and prints executed warp instructions and
stall samples per line of dagcon_kernels.cuh.

usage: ncu_by_line.py <report.ncu-rep> <kernel name> [top N]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "pbdagcon_b200", "csrc", "libdagcon_b200.so")


def sass_lines(kernel):
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=d, check=True, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info", cubin], cwd=d, check=True, stdout=subprocess.PIPE, text=True).stdout
    out, cur, infn = [], ("?", 0), False
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            infn = kernel in ln
            continue
        if not infn:
            continue
        m = re.search(r'//## File "(.*?)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            out.append(cur)
    return out


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ci, cs, ct = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    body = []
    for r in rows[h + 1:]:
        if r and r[0] == "Address":
            break  # a second launch of the same kernel
        if len(r) > ct and r[0].startswith("0x"):
            body.append(r)
    lines = sass_lines(kernel)
    if len(lines) != len(body):
        print(f"warning: {len(lines)} SASS instructions in the cubin vs {len(body)} rows in the report", file=sys.stderr)
    agg = {}
    for ln, r in zip(lines, body):
        a = agg.setdefault(ln, [0, 0, 0, 0])
        a[0] += int(r[ci] or 0); a[1] += int(r[cs] or 0); a[2] += int(r[ct] or 0); a[3] += 1
    srcs = {}

    def text(key):
        f, n = key
        if f not in srcs:
            try:
                srcs[f] = open(os.path.join(ROOT, "pbdagcon_b200", "csrc", f)).read().splitlines()
            except OSError:
                srcs[f] = []
        return srcs[f][n - 1].strip()[:100] if 0 < n <= len(srcs[f]) else ""
    ti = sum(a[0] for a in agg.values()) or 1
    ts = sum(a[1] for a in agg.values()) or 1
    print(f"{kernel}: {ti} warp instructions, {ts} stall samples, {len(body)} SASS instructions")
    print(f"{'file:line':>20} {'inst%':>6} {'stall%':>6} {'thr/inst':>8} {'sass':>4}  source")
    for ln, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        s = text(ln)
        print(f"{ln[0][:14]:>14}:{ln[1]:<5d} {100 * a[0] / ti:6.2f} {100 * a[1] / ts:6.2f} {a[2] / max(1, a[0]):8.1f} {a[3]:4d}  {s}")


if __name__ == "__main__":
    main()
