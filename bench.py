#!/usr/bin/env python
"""bench.py -- throughput of the DAGCon consensus hot path on B200 (see DESIGN.md "Measurement").

A step = one pass of the hot path (normalize/trim -> graph build -> mergeNodes -> bestPath/consensus) over one
batch of synthetic PacBio-like targets.  `--config N` picks the workload (SURVEY.md 8(d) numbering, 1-based;
BASELINE.json configs[N-1]):

  2 (default)  10 000 targets x 8 kb x 30x, flags -c 8 -m 500 -t 50           -- the headline
  3            dazcon on a synthetic DAZZ_DB + LAS pair (bench_dazcon.py)
  4            20 kb x 50x targets streamed through dgc_submit/dgc_collect from a recycled pool, weak
               scaling: --targets-total over all ranks
  5            256 targets x 40 kb x 100x, 3 refinement rounds (consensus fed back as the backbone)

  value     corrected Mbp/s with the packed batch already resident in HBM (dgc_upload + dgc_run)
  e2e       the same metric through the reference-facing call (dgc_submit/dgc_collect) from pinned HOST buffers:
            staged H2D of the batch, kernels, D2H of the CnsResults inside the timed region
  parity    per-target digests (status, ranges, sequences) of the GPU results against the reference's own code
            (oracle/_ref) on a sample of every rank's targets, and e2e-vs-resident over the whole batch
  cli       whole-process wall clock of the drop-in `pbdagcon` binary against `pbdagcon_ref` on the same m5 file
  --impl reference   the reference's own CPU code (oracle/_ref = its unmodified sources) on all host threads, on a
            bounded sample of the same workload.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

RANK = int(os.environ.get("RANK", "0"))
LOCAL_RANK = int(os.environ.get("LOCAL_RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))

MIN_COV, MIN_LEN, TRIM = 8, 500, 50
CLI = os.path.join(ROOT, "pbdagcon_b200", "host", "pbdagcon")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "pbdagcon_ref")

SHAPES = {2: (8000, 30.0, 10000, 2), 4: (20000, 50.0, 4800, 4), 5: (40000, 100.0, 256, 5)}   # L, cov, targets/GPU, seed


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def gen_workload(seed, L, cov, t0, nt, sub, threads):
    """nt targets starting at global index t0, generated in parallel in pieces of <= 125 targets and grouped into
    sub-batches of `sub` targets.  Returns a list of synth.Batch."""
    import pbdagcon_b200.synth as synth
    piece = max(1, min(125, int(125 * 8000 * 30 / (L * cov)))) if nt >= 125 else nt
    jobs = [(t0 + s, min(piece, nt - s)) for s in range(0, nt, piece)]
    with ThreadPoolExecutor(max_workers=max(1, threads)) as ex:
        pieces = list(ex.map(lambda j: synth.gen_batch(seed, L, cov, j[0], j[1]), jobs))
    per = max(1, -(-sub // piece))
    return [synth.concat(pieces[i:i + per]) for i in range(0, len(pieces), per)]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md 'clocks line')."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ref_library():
    """The CPU arm: oracle/_ref (the reference's own sources) when built, else the C restatement."""
    import helpers
    if helpers.have_ref():
        return helpers.Ref(), "reference"
    return helpers.Oracle(), "port"


def frag_digest(status, frags):
    """One target's canonical digest: status, then (range0, range1, md5(seq)) per fragment."""
    h = hashlib.md5()
    h.update(b"%d|" % int(status))
    for a, b, s in frags:
        h.update(b"%d_%d:" % (a, b))
        h.update(hashlib.md5(s).digest())
    return h.hexdigest()


def cpu_consensus(checker, work, threads, backbones=None):
    """work: [(tlen, alns)] -- the CPU implementation on `threads` host threads (ctypes releases the GIL).
    Returns (seconds, [frags per target])."""
    import helpers

    def one(i):
        L, a = work[i]
        bb = backbones[i] if backbones is not None else None
        return checker.consensus(L, a, backbone=bb, mode=helpers.MODE_PIPELINE, trim=TRIM, min_weight=MIN_COV, min_len=MIN_LEN)["frags"]

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        out = list(ex.map(one, range(len(work))))
    return time.perf_counter() - t0, out


def sample_work(batches, idxs, L):
    """[(tlen, alns)] of the global (within this rank) target indices idxs; unpacking is never timed."""
    starts, acc = [], 0
    for b in batches:
        starts.append(acc)
        acc += b.n_targets
    out = []
    for t in idxs:
        bi = max(i for i, s in enumerate(starts) if s <= t)
        out.append((L, batches[bi].target_alns(t - starts[bi])))
    return out


def results_digest(res):
    """md5 over a whole Results object (every status, range and consensus byte, in target order)."""
    h = hashlib.md5()
    for a in (res.status, res.frag_begin, res.range0, res.range1, res.seq_len):
        h.update(np.ascontiguousarray(a).tobytes())
    for f in range(int(res.seq_len.shape[0])):     # fragments in order; seq_off is an arena detail, not part of the result
        o, n = int(res.seq_off[f]), int(res.seq_len[f])
        h.update(res.seq[o:o + n].tobytes())
    return h.hexdigest()


def canon_md5(text):
    """md5 of the sorted 2-line FASTA records (the reference's -j >= 2 order is nondeterministic)."""
    lines = text.split(b"\n")
    recs = sorted(lines[i] + b"\n" + lines[i + 1] + b"\n" for i in range(0, len(lines) - 1, 2))
    bases = sum(len(lines[i + 1]) for i in range(0, len(lines) - 1, 2))
    return hashlib.md5(b"".join(recs)).hexdigest(), len(recs), bases


def timed_cli(cmd, reps):
    best, out = None, b""
    for _ in range(reps):
        t0 = time.perf_counter()
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            raise RuntimeError(f"{cmd[0]} failed: {r.stderr.decode()[-400:]}")
        out = r.stdout
        best = dt if best is None else min(best, dt)
    return best, out


def cli_block(args, n_gpus, threads, with_reference=True):
    """Whole-process wall clock (parse + consensus + FASTA) of the drop-in CLI on an m5 file on tmpfs, best of 3,
    against the reference binary with -j = host threads on the same file; canonical-record md5 must agree."""
    import pbdagcon_b200.synth as synth
    nt = min(args.targets, args.cli_targets)
    if nt <= 0 or not os.path.exists(CLI):
        return None
    tmp = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    path = os.path.join(tmp, f"dgc_bench_{os.getpid()}.m5")
    try:
        nbytes = synth.write_m5(args.seed, args.L, args.cov, 0, nt, path)
        flags = ["-c", str(MIN_COV), "-j", str(threads)]
        sec, out = timed_cli([CLI] + flags + ["--gpus", str(n_gpus), path], 3)
        md5, nrec, bases = canon_md5(out)
        blk = {"targets": nt, "m5_bytes": nbytes, "threads": threads, "gpus": n_gpus, "seconds": sec, "records": nrec,
               "value": bases / sec / 1e6, "unit": "Mbp/s", "targets_per_s": nt / sec, "md5": md5,
               "what": "pbdagcon CLI, whole process incl. CUDA init, m5 parse, FASTA write; best of 3"}
        if n_gpus > 1:      # the C++ dispatcher over N GPUs must write the same bytes as over one (input-target order)
            _, out1 = timed_cli([CLI] + flags + ["--gpus", "1", path], 1)
            blk["same_bytes_as_1gpu"] = out1 == out
            assert out1 == out, "pbdagcon --gpus N output differs from --gpus 1"
        if with_reference and os.path.exists(REF_CLI):
            rsec, rout = timed_cli([REF_CLI] + flags + [path], args.cli_ref_reps)
            rmd5, rrec, rbases = canon_md5(rout)
            blk.update({"reference_seconds": rsec, "reference_value": rbases / rsec / 1e6, "ratio": rsec / sec,
                        "reference_md5": rmd5, "identical": rmd5 == md5,
                        "reference_what": f"oracle/_ref/pbdagcon_ref -j {threads}, best of {args.cli_ref_reps}"})
            assert rmd5 == md5, "CLI FASTA differs from the reference binary's (canonical record order)"
        return blk
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass


def workload_config(args, **extra):
    c = {"workload": f"BASELINE configs[{args.config - 1}]: {args.targets} targets/GPU x {args.L} bp backbone x {args.cov:g}x "
                     f"(A={round(args.cov / 0.6)} alignments/target), 15% ins/del/sub error, flags -c {MIN_COV} -m {MIN_LEN} -t {TRIM}",
         "config": args.config, "targets_per_gpu": args.targets, "backbone_len": args.L, "coverage": args.cov, "seed": args.seed,
         "l2": "inputs (>= 1 GB of alignment strings per GPU) exceed the 126 MB L2; no flush needed"}
    c.update(extra)
    return c


def run_reference(args):
    """--impl reference: the reference's own code on the host cores.  Hot path over a bounded sample through
    oracle/_ref/libdagcon_ref.so (value), plus the whole-process pbdagcon_ref CLI on an m5 file (cli)."""
    if RANK != 0:
        return
    threads = host_threads()
    checker, kind = ref_library()
    per_target = args.L * args.cov / (8000 * 30.0)
    sample = int(min(args.targets, max(32, threads * 6 / max(1.0, per_target ** 1.3))))
    batches = gen_workload(args.seed, args.L, args.cov, 0, sample, sample, threads)
    work = sample_work(batches, range(sample), args.L)
    times, bases = [], 0
    for i in range(args.warmup + args.steps):
        dt, frags = cpu_consensus(checker, work, threads)
        if i >= args.warmup:
            times.append(dt)
            bases = sum(len(f[2]) for fr in frags for f in fr)
    sec = sum(times) / len(times)
    val = bases / sec / 1e6
    cli = None
    if args.config == 2 and os.path.exists(REF_CLI) and args.cli_targets > 0:
        import pbdagcon_b200.synth as synth
        nt = min(args.targets, args.cli_targets)
        path = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else "/tmp", f"dgc_bench_ref_{os.getpid()}.m5")
        try:
            synth.write_m5(args.seed, args.L, args.cov, 0, nt, path)
            rsec, rout = timed_cli([REF_CLI, "-c", str(MIN_COV), "-j", str(threads), path], args.cli_ref_reps)
            md5, nrec, rb = canon_md5(rout)
            cli = {"targets": nt, "threads": threads, "seconds": rsec, "value": rb / rsec / 1e6, "unit": "Mbp/s", "records": nrec,
                   "md5": md5, "what": f"oracle/_ref/pbdagcon_ref -j {threads} whole process, best of {args.cli_ref_reps}"}
        finally:
            try:
                os.unlink(path)
            except OSError:
                pass
    line = {"impl": "reference", "metric": "consensus_mbp_per_s", "value": val, "unit": "Mbp/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic", "targets_per_s": sample / sec,
            "config": workload_config(args, sample_targets=sample),
            "cpu_baseline": {"value": val, "unit": "Mbp/s", "cores": threads, "kind": kind,
                             "sample": f"{sample} of the workload's targets per step, hot path only (no parsing), "
                                       f"{threads} threads"},
            "e2e": {"value": val, "unit": "Mbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "cli": cli}
    print(json.dumps(line), flush=True)


class Dist:
    """torch.distributed plumbing: barrier + timing reductions + the ordered gather of the ranks' parity records
    (pbdagcon_b200/shard.py).  NCCL carries no consensus data: targets shard with no data-path collective."""

    def __init__(self, dev):
        import torch
        self.torch = torch
        self.dist = None
        if WORLD > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
            self.dist = dist

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _red(self, x, op):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def allmax(self, x):
        return self._red(x, self.dist.ReduceOp.MAX if self.dist else None)

    def allsum(self, x):
        return self._red(x, self.dist.ReduceOp.SUM if self.dist else None)

    def gather(self, records):
        from pbdagcon_b200.shard import gather_ordered
        return gather_ordered(records, self.dist)

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


def kernel_src_sha():
    """Hash of the kernel sources (csrc/*.cuh): profiles/traffic.json records the one its ncu capture was taken on."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "pbdagcon_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith(".cuh"):
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def bind_numa(dev):
    """Multi-rank runs: keep this rank's page-locked staging buffers (and, where the cpuset allows it, its threads) on
    the GPU's NUMA node.  Eight ranks uploading 5.4 GB each from one node's memory share that node's controllers and
    the socket interconnect (e2e efficiency 0.82 at N=8, 14.5 GB/s per GPU)."""
    if WORLD <= 1:
        return None
    info = {}
    try:
        import ctypes
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        node = -1
        for cand in (bus.lower(), bus.lower()[4:] if len(bus) > 12 else bus.lower()):
            try:
                node = int(open(f"/sys/bus/pci/devices/{cand}/numa_node").read())
                break
            except OSError:
                continue
        info["node"] = node
        if node >= 0:
            # memory first: set_mempolicy(MPOL_PREFERRED, {node}) -- cudaHostAlloc pages then come from that node
            mask = ctypes.c_ulong(1 << node)
            rc = ctypes.CDLL(None, use_errno=True).syscall(238, 1, ctypes.byref(mask), ctypes.c_ulong(64))
            info["mempolicy"] = "preferred" if rc == 0 else f"errno {ctypes.get_errno()}"
            cpus = set()
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                a, _, b = part.partition("-")
                cpus |= set(range(int(a), int(b or a) + 1))
            allowed = sorted(cpus & set(os.sched_getaffinity(0)))
            if allowed:
                os.sched_setaffinity(0, allowed)
                info["cpus"] = len(allowed)
            else:
                info["cpus"] = 0      # the cpuset keeps every rank on another node's cores: memory policy only
    except Exception as e:   # no NVML / sysfs / permission: run unbound, and say so
        info["error"] = str(e)[:80]
    return info


def pin_batches(api, subs, packed4=True):
    """Page-locked staging buffers for the e2e arm.  packed4: the strings go to the device as 4-bit codes, two columns
    per byte (DGC_BATCH_PACKED4) -- what the C++ packers of the CLI send; half the bytes on the link."""
    import pbdagcon_b200.synth as synth
    pinned, sub_pb = [], []
    for b in subs:
        r = synth.pack4(b) if packed4 else None
        arena, qo, to = r if r is not None else (b.arena, b.aln_q_off, b.aln_t_off)
        pa = api.PinnedArray((arena.size,), np.uint8)
        pa.array[:] = arena
        pinned.append(pa)
        sub_pb.append(api.PackedBatch(b.tgt_tlen, b.tgt_aln_begin, b.aln_start, b.aln_len, qo, to, pa.array, packed4=r is not None))
        assert sub_pb[-1].arena.ctypes.data == pa.ptr     # a view of the pinned array, not a copy
    return pinned, sub_pb


def e2e_pass(api, eng, sub_pb, keep=False):
    """Every sub-batch through dgc_submit/dgc_collect from pinned host buffers; returns (bases, d2h bytes, results)."""
    nb = d2h = 0
    nsub, sent, got = len(sub_pb), 0, 0
    kept = []
    while got < nsub:
        while sent < nsub and sent - got < api.MAX_INFLIGHT:
            eng.submit(sub_pb[sent])
            sent += 1
        r = eng.collect(copy=keep)
        got += 1
        if keep:
            kept.append(r)
            nb += r.total_bases()
        else:
            nb += int(np.ctypeslib.as_array(r.frag_seq_len, (r.n_frags,)).astype(np.int64).sum()) if r.n_frags else 0
            d2h += int(r.seq_bytes) + 8 * r.n_frags + 16 * r.n_targets
    return nb, d2h, kept


def roofline_block(args, phases, steps, chunks, alg_bytes_step):
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kern = {"normalize_trim": phases.get("ms_normalize", 0), "build_graph": phases.get("ms_build", 0),
            "merge_nodes": phases.get("ms_merge", 0), "best_path": phases.get("ms_bestpath", 0)}
    if kern["best_path"] == 0:      # default: both sweeps run as one launch; --flags 512 splits them
        kern = {"normalize_trim": kern["normalize_trim"], "build_graph": kern["build_graph"], "merge_best": kern["merge_nodes"]}
    top = max(kern, key=kern.get)
    n_launch_top = steps * max(1, chunks) if top != "normalize_trim" else steps
    avg_ms = kern[top] / n_launch_top
    per_launch = alg_bytes_step / max(1, chunks) if top != "normalize_trim" else alg_bytes_step
    achieved = per_launch / (avg_ms / 1e3) / 1e9
    traffic, traffic_note = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("config") != args.config or tj.get("targets") != args.targets:
            traffic_note = "profiles/traffic.json was captured on another workload: not quoted"
        elif tj.get("src_sha") != kernel_src_sha():
            traffic_note = "profiles/traffic.json was captured on other kernel sources (src_sha differs): refused as stale"
        else:
            traffic = tj.get(top)
            traffic_note = tj.get("_source")
    except (OSError, ValueError):
        pass
    return {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "traffic_source": traffic_note, "peak_source": peak_src, "avg_launch_ms": avg_ms,
            "alg_bytes_per_launch": per_launch, "kernel_ms_per_step": {k: v / steps for k, v in kern.items()},
            "whole_path_frac": alg_bytes_step / (sum(kern.values()) / steps / 1e3) / 1e9 / peak}


def parity_block(dd, res_frags, res_status, subs, n_targets, L, threads, rank_t0, extra_idxs=()):
    """GPU results of this rank against the reference's own code on a strided sample of its targets; every rank
    checks its own range and rank 0 gathers the verdicts in rank order."""
    checker, kind = ref_library()
    want = max(16, min(n_targets, threads * 4))
    stride = max(1, n_targets // want)
    idxs = sorted(set(list(range(0, n_targets, stride))[:want]) | set(extra_idxs))
    work = sample_work(subs, idxs, L)
    sec, exp = cpu_consensus(checker, work, threads)
    bad = [rank_t0 + t for t, e in zip(idxs, exp) if frag_digest(0, e) != frag_digest(res_status(t), res_frags(t))]
    mine = [(RANK, len(idxs), len(bad), bad[:5], hashlib.md5("".join(frag_digest(res_status(t), res_frags(t)) for t in idxs).encode()).hexdigest())]
    allr = dd.gather(mine)
    if allr is None:
        return None, sec, exp, idxs
    return {"checker": kind, "targets": sum(r[1] for r in allr), "mismatches": sum(r[2] for r in allr),
            "ok": all(r[2] == 0 for r in allr), "ranks": len(allr), "first_bad": [b for r in allr for b in r[3]][:5],
            "digest_of_digests": hashlib.md5("".join(r[4] for r in allr).encode()).hexdigest(),
            "what": "per-target md5(status, range0, range1, md5(seq)) of the GPU results vs oracle/_ref on a strided sample of every rank's targets"}, sec, exp, idxs


def run_gpu(args):
    import torch
    import pbdagcon_b200.api as api
    from pbdagcon_b200.shard import target_range
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    dev = LOCAL_RANK
    torch.cuda.set_device(dev)
    numa = bind_numa(dev)
    dd = Dist(dev)
    cores = host_threads()
    threads = max(1, host_threads() // max(1, WORLD))

    # ---- workload: every rank owns a static range of the target stream (weak scaling, no collective on the data path)
    t_gen = time.perf_counter()
    rank_t0, rank_t1 = target_range(args.targets * WORLD, RANK, WORLD)
    subs = gen_workload(args.seed, args.L, args.cov, rank_t0, rank_t1 - rank_t0, args.sub, threads)
    import pbdagcon_b200.synth as synth
    pinned, sub_pb = pin_batches(api, subs, packed4=not args.ascii_wire)
    if len(subs) > 1:
        wb = synth.concat(subs)
        whole = (None if args.ascii_wire else api.PackedBatch.from_synth_packed4(wb)) or api.PackedBatch.from_synth(wb)
        del wb
    else:
        whole = sub_pb[0]
    gen_s = time.perf_counter() - t_gen
    alg_in = sum(b.alg_bytes() for b in subs)
    h2d_bytes = sum(int(p.arena.size) + 4 * (2 * p.n_targets + 1) + 24 * p.n_alns + 8 * p.n_alns for p in sub_pb)

    eng = api.Consensus(min_cov=MIN_COV, min_len=MIN_LEN, trim=TRIM, device=dev, scratch_bytes=args.scratch_gb << 30, flags=args.flags)
    # ---- arm 1: HBM-resident
    db = eng.upload(whole)
    if len(subs) > 1:
        whole.arena = None                      # the HBM copy is what the resident arm reads
        whole.c.arena = None
    sampler = ClockSampler(dev)
    phases = {}
    launches = 0
    for _ in range(args.warmup):
        eng.run(db, copy=False)
    dd.barrier()
    sampler.start()
    span_ms = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.run(db, copy=False)
        tm = eng.timings()
        span_ms.append(tm["ms_span"])
        launches += tm["launches"]
        for k, v in tm.items():
            if k.startswith("ms_"):
                phases[k] = phases.get(k, 0.0) + v
    dd.barrier()
    wall_s = time.perf_counter() - t0
    clocks = sampler.stop()
    res = eng.run(db)                       # a copy of the last results for the parity check / base count
    bases = res.total_bases()
    n_ok = int((res.status == api.T_OK).sum())
    frag_bytes = int(res.seq_len.astype(np.int64).sum()) + 12 * int(res.seq_len.shape[0])
    resident_digest = results_digest(res)
    eng.free(db)
    dev_s = dd.allmax(sum(span_ms) / 1e3)      # CUDA events on the launching stream, max over ranks
    wall_s = dd.allmax(wall_s)
    tot_bases = dd.allsum(bases)
    tot_targets = dd.allsum(args.targets)
    value = tot_bases * args.steps / dev_s / 1e6

    # ---- arm 2: end to end from pinned host buffers through submit/collect
    for _ in range(max(1, args.warmup - 1)):
        e2e_pass(api, eng, sub_pb)
    dd.barrier()
    t0 = time.perf_counter()
    e2e_bases = 0
    d2h_bytes = 0
    for _ in range(args.steps):
        nb, d2h_bytes, _ = e2e_pass(api, eng, sub_pb)
        e2e_bases += nb
    dd.barrier()
    e2e_s = dd.allmax(time.perf_counter() - t0)
    e2e_val = dd.allsum(e2e_bases) / e2e_s / 1e6
    # the e2e arm must reproduce the resident arm bit for bit, over the WHOLE batch (untimed extra pass)
    e2e_same = None
    if len(sub_pb) == 1:
        _, _, kept = e2e_pass(api, eng, sub_pb, keep=True)
        e2e_same = results_digest(kept[0]) == resident_digest
        assert e2e_same, "e2e and resident arms disagree on the consensus"
    assert e2e_bases == bases * args.steps, "e2e and resident arms disagree on the consensus"

    # ---- parity on every rank against the reference's own code; timed CPU baseline on rank 0 at N=1
    nt_local = rank_t1 - rank_t0
    cpu = None
    extra = []
    if WORLD == 1 and not args.no_cpu:
        per_target = args.L * args.cov / (8000 * 30.0)
        extra = list(range(int(min(nt_local, max(16, cores * 12 / max(1.0, per_target ** 1.3))))))
    parity, _, _, _ = parity_block(dd, res.frags, lambda t: res.status[t], subs, nt_local, args.L, threads, rank_t0) \
        if not args.no_cpu or WORLD > 1 else (None, 0, None, None)
    if extra:
        checker, kind = ref_library()
        sec, exp = cpu_consensus(checker, sample_work(subs, extra, args.L), cores)
        nb = sum(len(f[2]) for fr in exp for f in fr)
        cpu = {"value": nb / sec / 1e6, "unit": "Mbp/s", "cores": cores, "kind": kind, "targets_per_s": len(extra) / sec,
               "sample": f"first {len(extra)} targets of the workload, hot path only, {cores} threads, {sec:.1f} s wall"}
        bad = [t for t, e in zip(extra, exp) if e != res.frags(t)]
        if parity is not None:
            parity["targets"] += len(extra); parity["mismatches"] += len(bad); parity["ok"] = parity["ok"] and not bad
    eng.close()
    for p in pinned:
        p.free()
    dd.barrier()          # every rank has released its GPU memory: rank 0 may now drive all GPUs through the C++ dispatcher

    if RANK != 0:
        dd.barrier()      # wait for rank 0's CLI block (it drives every GPU)
        dd.close()
        return

    cli = None
    if args.config == 2 and args.cli_targets > 0:
        try:
            cli = cli_block(args, WORLD, cores, with_reference=(WORLD == 1 and not args.no_cpu))
        except AssertionError:
            raise
        except Exception as e:
            cli = {"error": str(e)[:300]}
    dd.barrier()
    alg_bytes_step = alg_in + frag_bytes                 # B_alg summed over the batch
    roofline = roofline_block(args, phases, args.steps, tm["chunks"], alg_bytes_step)
    if parity is not None:
        parity["e2e_equals_resident_whole_batch"] = e2e_same
        assert parity["ok"], f"GPU results differ from the reference on the sample: {parity}"
    line = {"metric": "consensus_mbp_per_s", "value": value, "unit": "Mbp/s", "n_gpus": WORLD, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
            "config": workload_config(args, sub_batch_targets=args.sub, targets_ok=n_ok, gen_seconds=round(gen_s, 1),
                                      chunks_per_step=tm["chunks"], numa=numa,
                                      wire="ASCII" if args.ascii_wire else "4-bit codes (DGC_BATCH_PACKED4), packed on the host side of the C ABI"),
            "targets_per_s": tot_targets * args.steps / dev_s, "wall_ms_per_step": wall_s * 1e3 / args.steps,
            "e2e": {"value": e2e_val, "unit": "Mbp/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "targets_per_s": tot_targets * args.steps / e2e_s, "ms_per_step": e2e_s * 1e3 / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "cli": cli}
    print(json.dumps(line), flush=True)
    dd.close()


def run_stream(args):
    """config 4: the pre-assembly shape streamed from a recycled pool (1 M x 20 kb x 50x is 2.2 TB and
    cannot be materialised).  Every rank owns `pool` distinct targets (seeded by global index) and pushes them through
    dgc_submit/dgc_collect pass after pass, two passes in flight, until its share of --targets-total is done; a step
    = one pass.  Memory is bounded by the pool and the handle's arenas."""
    import torch
    import pbdagcon_b200.api as api
    from pbdagcon_b200.shard import target_range
    dev = LOCAL_RANK
    torch.cuda.set_device(dev)
    numa = bind_numa(dev)
    dd = Dist(dev)
    cores = host_threads()
    threads = max(1, host_threads() // max(1, WORLD))
    pool = args.targets
    t0g, t1g = target_range(args.targets_total, RANK, WORLD)
    share = t1g - t0g
    passes = max(1, -(-share // pool))
    t_gen = time.perf_counter()
    subs = gen_workload(args.seed, args.L, args.cov, RANK * pool, pool, args.sub, threads)
    pinned, sub_pb = pin_batches(api, subs, packed4=not args.ascii_wire)
    gen_s = time.perf_counter() - t_gen
    alg_in = sum(b.alg_bytes() for b in subs)
    h2d_bytes = sum(int(p.arena.size) + 4 * (2 * p.n_targets + 1) + 32 * p.n_alns for p in sub_pb)
    eng = api.Consensus(min_cov=MIN_COV, min_len=MIN_LEN, trim=TRIM, device=dev, scratch_bytes=args.scratch_gb << 30, flags=args.flags)
    _, _, kept = e2e_pass(api, eng, sub_pb, keep=True)         # warm-up pass, kept for parity
    for _ in range(max(0, args.warmup - 1)):
        e2e_pass(api, eng, sub_pb)
    sampler = ClockSampler(dev)
    dd.barrier()
    sampler.start()
    t0 = time.perf_counter()
    nb = d2h = 0
    phases, launches, chunks = {}, 0, 1
    for _ in range(passes):
        b, d2h, _ = e2e_pass(api, eng, sub_pb)
        nb += b
        tm = eng.timings()
        launches += tm["launches"] * len(sub_pb)
        chunks = tm["chunks"]
        for k, v in tm.items():
            if k.startswith("ms_"):
                phases[k] = phases.get(k, 0.0) + v * len(sub_pb)
    dd.barrier()
    sec = dd.allmax(time.perf_counter() - t0)
    clocks = sampler.stop()
    tot_bases = dd.allsum(nb)
    tot_targets = dd.allsum(passes * pool)
    # parity: the warm-up pass against the reference on a sample of every rank's pool
    starts = np.cumsum([0] + [r.n_targets for r in kept])

    def loc(t):
        i = int(np.searchsorted(starts, t, side="right") - 1)
        return kept[i], t - int(starts[i])
    parity, cpu_sec, exp, idxs = parity_block(dd, lambda t: loc(t)[0].frags(loc(t)[1]), lambda t: loc(t)[0].status[loc(t)[1]],
                                              subs, pool, args.L, threads, RANK * pool)
    eng.close()
    if RANK != 0:
        dd.close()
        return
    cpu_bases = sum(len(f[2]) for fr in exp for f in fr)
    cpu = {"value": cpu_bases / cpu_sec / 1e6, "unit": "Mbp/s", "cores": threads, "kind": "reference" if parity["checker"] == "reference" else "port",
           "targets_per_s": len(idxs) / cpu_sec,
           "sample": f"{len(idxs)} targets of rank 0's pool (strided), hot path only, {threads} threads, {cpu_sec:.1f} s wall; targets are "
                     f"independent, so the full job extrapolates linearly"}
    frag_bytes = nb // max(1, passes) + 12 * sum(int(r.seq_len.shape[0]) for r in kept)
    roofline = roofline_block(args, phases, passes * len(sub_pb), chunks, (alg_in + frag_bytes) / len(sub_pb))
    assert parity["ok"], f"GPU results differ from the reference on the sample: {parity}"
    val = tot_bases / sec / 1e6
    line = {"metric": "consensus_mbp_per_s", "value": val, "unit": "Mbp/s", "n_gpus": WORLD, "steps": passes, "warmup": args.warmup,
            "ms_per_step": sec * 1e3 / passes, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/int32/f32", "data": "synthetic (recycled pool, SURVEY 0.10)",
            "config": workload_config(args, targets_total=int(tot_targets), pool_targets_per_gpu=pool, passes=passes,
                                      sub_batch_targets=args.sub, gen_seconds=round(gen_s, 1), numa=numa,
                                      note="value is measured through dgc_submit/dgc_collect from pinned host buffers (H2D inside "
                                           "the timed region): for a streamed job there is no HBM-resident arm, value == e2e"),
            "targets_per_s": tot_targets / sec,
            "e2e": {"value": val, "unit": "Mbp/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h,
                    "targets_per_s": tot_targets / sec, "ms_per_step": sec * 1e3 / passes},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity}
    print(json.dumps(line), flush=True)
    dd.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5], help="SURVEY 8(d) workload number (BASELINE configs[N-1])")
    ap.add_argument("--targets", type=int, default=0, help="targets per GPU (config 4: pool size per GPU)")
    ap.add_argument("--targets-total", type=int, default=0, help="config 4: targets streamed over all ranks (default 4 passes of the pool)")
    ap.add_argument("--L", type=int, default=0)
    ap.add_argument("--cov", type=float, default=0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--sub", type=int, default=0, help="targets per dgc_submit in the e2e arm (0 = the whole step in one "
                    "submit: the library stages the upload itself)")
    ap.add_argument("--scratch-gb", type=int, default=0, help="dgc_opts.scratch_bytes in GiB (0 = the library default)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--ascii-wire", action="store_true", help="send the strings as ASCII (default: 4-bit codes, DGC_BATCH_PACKED4)")
    ap.add_argument("--cli-targets", type=int, default=2000, help="targets of the whole-process CLI comparison (0 = skip)")
    ap.add_argument("--cli-ref-reps", type=int, default=1)
    ap.add_argument("--flags", type=int, default=0, help="extra dgc_opts.flags (tuning, e.g. 256 = no prefetch)")
    ap.add_argument("--rounds", type=int, default=3, help="config 5: refinement rounds")
    args = ap.parse_args()
    if args.config == 3:
        import bench_dazcon
        return bench_dazcon.main(args)
    L, cov, nt, seed = SHAPES[args.config]
    args.L = args.L or L
    args.cov = args.cov or cov
    args.targets = args.targets or nt
    args.seed = args.seed or seed
    if args.sub <= 0:
        args.sub = args.targets
    if args.config == 4 and args.targets_total <= 0:
        args.targets_total = 4 * args.targets * WORLD
    if args.impl == "reference":
        run_reference(args)
    elif args.config == 4:
        run_stream(args)
    elif args.config == 5:
        import bench_refine
        bench_refine.main(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
