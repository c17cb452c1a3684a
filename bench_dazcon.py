"""bench.py --config 3: dazcon on a synthetic DAZZ_DB + LAS pair (BASELINE configs[2]: "50k reads x 10 kb, 25x,
daligner-format alignments"; the default run uses --targets 4000 reads so that it ends within minutes, the full size
is `--targets 50000`).

The pair is made on the box by the reference's own tooling (built into oracle/_ref by oracle/Makefile; it travels with
the snapshot): DAZZ_DB `simulator` (15 % error: ins .733 / del .2 / sub .067) | fasta2DB; DBsplit; daligner per block;
LAsort; LAmerge -- SURVEY.md 8(d) cfg 3.  Generating inputs is test tooling, not the product.

  value   consensus Mbp/s of the GPU path on the HBM-resident batch the dazcon front packed (dgc_upload + dgc_run)
  e2e     the drop-in `dazcon` binary, whole process wall clock (LAS/DB read, hit selection, trace-point realignment,
          H2D, kernels, D2H, FASTA), best of 3
  parity  `dazcon` against `dazcon_ref -j 1` on a sample of target ids, byte for byte (header counter included)
  cpu_baseline / --impl reference   `dazcon_ref -j <host threads>` whole process on a bounded sample of target ids
"""
import ctypes as C
import hashlib
import json
import os
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(ROOT, "oracle", "_ref")
DAZCON = os.path.join(ROOT, "pbdagcon_b200", "host", "dazcon")
HOST_SO = os.path.join(ROOT, "pbdagcon_b200", "host", "libdagcon_host.so")
RANK = int(os.environ.get("RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
FLAGS = ["-c", "6", "-l", "500", "-t", "10", "-m", "85"]      # dazcon defaults (reference dazcon.cpp:122-192)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def make_pair(n_reads, seed, workdir, threads):
    """reads.db + reads.las for about n_reads reads of 10 kb at 25x; returns (db, las, seconds)."""
    t0 = time.perf_counter()
    os.makedirs(workdir, exist_ok=True)
    genome_mb = n_reads * 10000 / 25 / 1e6

    def tool(name, args, **kw):
        subprocess.run([os.path.join(REF, name)] + args, cwd=workdir, check=True, **kw)
    with open(os.path.join(workdir, "reads.fasta"), "wb") as f:
        tool("simulator", [f"{genome_mb:.4f}", "-c25", "-m10000", "-s2000", "-x4000", "-e.15", f"-r{seed}"], stdout=f)
    tool("fasta2DB", ["reads", "reads.fasta"])
    os.unlink(os.path.join(workdir, "reads.fasta"))
    tool("DBsplit", ["-x500", "-s200", "reads"])
    nblocks = 1
    for ln in open(os.path.join(workdir, "reads.db")):
        if ln.startswith("blocks ="):
            nblocks = int(ln.split("=")[1])
    if nblocks == 1:
        tool("daligner", ["reads", "reads"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    else:
        for i in range(1, nblocks + 1):      # the HPCdaligner plan: block i against blocks 1..i
            tool("daligner", [f"reads.{i}"] + [f"reads.{j}" for j in range(1, i + 1)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    parts = sorted(p for p in os.listdir(workdir) if p.endswith(".las") and p.startswith("reads."))
    for i in range(0, len(parts), 64):
        tool("LAsort", parts[i:i + 64])
    for p in parts:
        os.unlink(os.path.join(workdir, p))
    sorted_parts = [p[:-4] + ".S.las" for p in parts]
    # LAmerge takes up to 252 files; merge in groups when there are more
    level, cur = 0, sorted_parts
    while len(cur) > 1:
        nxt = []
        for i in range(0, len(cur), 200):
            out = f"merge.{level}.{i}.las" if len(cur) > 200 else "reads.las"
            tool("LAmerge", [out] + cur[i:i + 200])
            nxt.append(out)
        for p in cur:
            os.unlink(os.path.join(workdir, p))
        cur, level = nxt, level + 1
    if cur[0] != "reads.las":
        os.rename(os.path.join(workdir, cur[0]), os.path.join(workdir, "reads.las"))
    return os.path.join(workdir, "reads.db"), os.path.join(workdir, "reads.las"), time.perf_counter() - t0


def fasta_stats(out):
    lines = out.split(b"\n")
    return len(lines) // 2, sum(len(lines[i + 1]) for i in range(0, len(lines) - 1, 2))


def timed(cmd, reps=1, env=None):
    best, out = None, b""
    for _ in range(reps):
        t0 = time.perf_counter()
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            raise RuntimeError(f"{os.path.basename(cmd[0])} failed: {r.stderr.decode()[-300:]}")
        best, out = (dt if best is None else min(best, dt)), r.stdout
    return best, out


def n_reads_of(db):
    for ln in open(db):
        f = ln.split()
        if len(f) == 3 and f[0].isdigit():
            return int(f[0])
    return 0


def pack_daz(db, las, threads, ids=()):
    """The dazcon front (LAS scan, hit selection, trace-point realignment on `threads` host threads) -> PackedBatch.
    ids: restrict to these 1-based target ids (dazcon's positional arguments); empty = every target."""
    import pbdagcon_b200.api as api
    lib = C.CDLL(HOST_SO)
    lib.dgh_pack_daz.restype = C.c_void_p
    lib.dgh_pack_daz.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_double),
                                 C.POINTER(C.c_int), C.c_int]
    lib.dgh_pack_counts.restype = C.c_uint32
    lib.dgh_pack_counts.argtypes = [C.c_void_p, C.c_int]
    lib.dgh_pack_arena_bytes.restype = C.c_uint64
    lib.dgh_pack_arena_bytes.argtypes = [C.c_void_p]
    lib.dgh_pack_array.restype = C.c_void_p
    lib.dgh_pack_array.argtypes = [C.c_void_p, C.c_int]
    lib.dgh_pack_error.restype = C.c_char_p
    lib.dgh_pack_error.argtypes = [C.c_void_p]
    lib.dgh_pack_free.argtypes = [C.c_void_p]
    sec = C.c_double()
    idarr = (C.c_int * len(ids))(*[int(i) for i in ids])
    h = lib.dgh_pack_daz(db.encode(), las.encode(), 6, 85, 0, 0, threads, 4096, C.byref(sec), idarr, len(ids))
    err = lib.dgh_pack_error(h)
    if err:
        raise RuntimeError(err.decode())
    T, A, nb = lib.dgh_pack_counts(h, 0), lib.dgh_pack_counts(h, 1), lib.dgh_pack_arena_bytes(h)

    def arr(which, n, dt):
        p = lib.dgh_pack_array(h, which)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), (n * np.dtype(dt).itemsize,)).view(dt).copy()
    pb = api.PackedBatch(arr(0, T, np.uint32), arr(1, T + 1, np.uint32), arr(2, A, np.uint32), arr(3, A, np.uint32),
                         arr(4, A, np.uint64), arr(5, A, np.uint64), arr(6, nb, np.uint8), arr(9, T, np.uint64))
    lib.dgh_pack_free(h)
    return pb, sec.value


def sample_ids(n, k):
    stride = max(1, n // max(1, k))
    return [str(i) for i in range(1, n + 1, stride)][:k]


def main(args):
    if RANK != 0:
        return        # the dazcon binary drives every GPU itself (one process, C++ dispatcher): no per-rank work
    threads = host_threads()
    n_reads = args.targets or 4000
    work = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else "/tmp", f"dgc_daz_{os.getpid()}")
    try:
        db, las, gen_s = make_pair(n_reads, args.seed or 3, work, threads)
        n = n_reads_of(db)
        cfg = {"workload": f"BASELINE configs[2]: dazcon on a synthetic DAZZ_DB/LAS pair, {n} reads x 10 kb, 25x (simulator -e.15 | fasta2DB; "
                           f"DBsplit; daligner; LAsort; LAmerge), flags {' '.join(FLAGS)}",
               "config": 3, "reads": n, "las_bytes": os.path.getsize(las), "gen_seconds": round(gen_s, 1),
               "l2": "the packed alignment strings (>= 1 GB) exceed the 126 MB L2; no flush needed"}
        k_cpu = min(n, max(64, 48 * threads))
        ids = sample_ids(n, k_cpu)
        if args.impl == "reference":
            times, bases, nrec = [], 0, 0
            for i in range(max(1, min(args.warmup, 1)) + args.steps):
                sec, out = timed([os.path.join(REF, "dazcon_ref"), "-j", str(threads), "-s", db, "-a", las] + FLAGS + ids)
                if i >= max(1, min(args.warmup, 1)):
                    times.append(sec)
                    nrec, bases = fasta_stats(out)
            sec = sum(times) / len(times)
            val = bases / sec / 1e6
            print(json.dumps({"impl": "reference", "metric": "consensus_mbp_per_s", "value": val, "unit": "Mbp/s", "n_gpus": args.gpus,
                              "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
                              "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic", "config": cfg,
                              "targets_per_s": len(ids) / sec,
                              "cpu_baseline": {"value": val, "unit": "Mbp/s", "cores": threads, "kind": "reference",
                                               "sample": f"dazcon_ref -j {threads} whole process on {len(ids)} of {n} target ids (the LAS is read in full)"},
                              "e2e": {"value": val, "unit": "Mbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
            return
        import torch
        import pbdagcon_b200.api as api
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
        sys.path.insert(0, ROOT)
        from bench import ClockSampler, roofline_block
        # ---- e2e: the dazcon binary, whole process
        sec_cli, out = timed([DAZCON, "-j", str(threads), "-s", db, "-a", las] + FLAGS + ["--gpus", str(max(1, args.gpus))], 3)
        nrec, bases_cli = fasta_stats(out)
        # ---- parity: byte-identical to `dazcon_ref -j 1` on a sample of ids (and the 1-GPU run equals the N-GPU run)
        pid = sample_ids(n, min(n, 48))
        _, ours = timed([DAZCON, "-j", str(threads), "-s", db, "-a", las] + FLAGS + pid)
        _, theirs = timed([os.path.join(REF, "dazcon_ref"), "-j", "1", "-s", db, "-a", las] + FLAGS + pid)
        parity = {"checker": "reference", "targets": len(pid), "ok": ours == theirs, "records": fasta_stats(ours)[0],
                  "what": "dazcon vs dazcon_ref -j 1 on a strided sample of target ids, byte for byte (header counter included)",
                  "md5": hashlib.md5(ours).hexdigest(), "reference_md5": hashlib.md5(theirs).hexdigest()}
        assert parity["ok"], "dazcon FASTA differs from dazcon_ref -j 1"
        if args.gpus > 1:
            _, one = timed([DAZCON, "-j", str(threads), "-s", db, "-a", las] + FLAGS + ["--gpus", "1"])
            parity["same_bytes_as_1gpu"] = one == out
            assert one == out
        # ---- CPU baseline: dazcon_ref with every host thread on a bounded sample of ids
        cpu_sec, cpu_out = timed([os.path.join(REF, "dazcon_ref"), "-j", str(threads), "-s", db, "-a", las] + FLAGS + ids)
        cpu_rec, cpu_bases = fasta_stats(cpu_out)
        cpu = {"value": cpu_bases / cpu_sec / 1e6, "unit": "Mbp/s", "cores": threads, "kind": "reference", "targets_per_s": len(ids) / cpu_sec,
               "sample": f"dazcon_ref -j {threads} whole process on {len(ids)} of {n} target ids, {cpu_sec:.1f} s wall"}
        # ---- resident arm: what the front packs, through dgc_upload / dgc_run
        # (at most RESIDENT_CAP targets: 50 000 reads are 50 GB of strings + 100 GB of normalised columns in one batch)
        rcap = min(n, 6000)
        rids = list(range(1, rcap + 1)) if rcap < n else []
        pb, pack_s = pack_daz(db, las, threads, rids)
        eng = api.Consensus(min_cov=6, min_len=500, trim=10, device=0)
        dbh = eng.upload(pb)
        for _ in range(args.warmup):
            eng.run(dbh, copy=False)
        sampler = ClockSampler(0)
        sampler.start()
        phases, launches, span = {}, 0, []
        for _ in range(args.steps):
            eng.run(dbh, copy=False)
            tm = eng.timings()
            span.append(tm["ms_span"])
            launches += tm["launches"]
            for kk, v in tm.items():
                if kk.startswith("ms_"):
                    phases[kk] = phases.get(kk, 0.0) + v
        clocks = sampler.stop()
        res = eng.run(dbh)
        bases = res.total_bases()
        if not rids:
            assert bases == bases_cli, "the resident run and the dazcon binary disagree on the consensus"
        frag_bytes = bases + 12 * int(res.seq_len.shape[0])
        eng.free(dbh)
        eng.close()
        dev_s = sum(span) / 1e3
        value = bases * args.steps / dev_s / 1e6
        args.config, args.targets = 3, pb.n_targets
        roofline = roofline_block(args, phases, args.steps, tm["chunks"], pb.alg_bytes() + frag_bytes)
        h2d = int(pb.arena.size) + 4 * (2 * pb.n_targets + 1) + 8 * pb.n_targets + 32 * pb.n_alns
        line = {"metric": "consensus_mbp_per_s", "value": value, "unit": "Mbp/s", "n_gpus": max(1, args.gpus), "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
                "config": dict(cfg, resident_targets=pb.n_targets, targets_packed=pb.n_targets, alignments=pb.n_alns, host_front_seconds=round(pack_s, 2),
                               host_front_threads=threads, host_front_mbp_per_s=round(bases / pack_s / 1e6, 2),
                               note="value = the GPU path on the HBM-resident batch (one GPU); e2e = the dazcon binary, whole process, on "
                                    "--gpus GPUs; host_front_* = LAS scan + hit selection + trace-point realignment on the host threads alone"),
                "targets_per_s": pb.n_targets * args.steps / dev_s,
                "e2e": {"value": bases_cli / sec_cli / 1e6, "unit": "Mbp/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": frag_bytes + 16 * pb.n_targets,
                        "targets_per_s": n / sec_cli, "ms_per_step": sec_cli * 1e3, "records": nrec,
                        "what": "dazcon binary, whole process incl. CUDA init, best of 3"},
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity}
        print(json.dumps(line), flush=True)
    finally:
        shutil.rmtree(work, ignore_errors=True)
