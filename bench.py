#!/usr/bin/env python
"""bench.py -- throughput of the DAGCon consensus hot path on B200 (see DESIGN.md "Measurement").

A step = one pass of the hot path (normalize/trim -> graph build -> mergeNodes -> bestPath/consensus) over one
batch of synthetic PacBio-like targets.  Default workload = BASELINE.json configs[1]: 10 000 targets x 8 kb
backbone x 30x coverage, 15 % indel-dominated error, pbdagcon flags -c 8 (-m 500 -t 50).

  value     corrected Mbp/s with the packed batch already resident in HBM (dgc_upload + dgc_run)
  e2e       the same metric through the reference-facing call (dgc_submit/dgc_collect) from pinned HOST buffers:
            staged H2D of the batch, kernels, D2H of the CnsResults inside the timed region
  --impl reference   the reference's own CPU code (oracle/_ref/libdagcon_ref.so = its unmodified sources) on all
            host threads, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
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

RANK = int(os.environ.get("RANK", "0"))
LOCAL_RANK = int(os.environ.get("LOCAL_RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))

MIN_COV, MIN_LEN, TRIM = 8, 500, 50


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def gen_workload(seed, L, cov, t0, nt, sub, threads):
    """nt targets starting at global index t0, generated in parallel in pieces of <= 125 targets and grouped into
    sub-batches of `sub` targets.  Returns a list of synth.Batch."""
    import pbdagcon_b200.synth as synth
    piece = 125 if nt >= 125 else nt
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
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    if helpers.have_ref():
        return helpers.Ref(), "reference"
    return helpers.Oracle(), "port"


def cpu_consensus(checker, batches, n_targets, L, threads):
    """Run the CPU implementation over the first n_targets targets on `threads` host threads (ctypes releases the
    GIL).  Returns (seconds, consensus bases)."""
    import helpers
    work = []
    for b in batches:
        for t in range(b.n_targets):
            if len(work) < n_targets:
                work.append((b, t))
    alns = [b.target_alns(t) for b, t in work]   # unpacking is not part of the timed region

    def one(a):
        r = checker.consensus(L, a, mode=helpers.MODE_PIPELINE, trim=TRIM, min_weight=MIN_COV, min_len=MIN_LEN)
        return sum(len(f[2]) for f in r["frags"])

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        bases = sum(ex.map(one, alns))
    return time.perf_counter() - t0, bases, len(alns)


def run_reference(args):
    if RANK != 0:
        return
    threads = host_threads()
    checker, kind = ref_library()
    sample = min(args.targets, max(threads * 6, 32))
    batches = gen_workload(args.seed, args.L, args.cov, 0, sample, sample, threads)
    times, bases = [], 0
    for i in range(args.warmup + args.steps):
        dt, nb, nt = cpu_consensus(checker, batches, sample, args.L, threads)
        if i >= args.warmup:
            times.append(dt); bases = nb
    sec = sum(times) / len(times)
    val = bases / sec / 1e6
    line = {"impl": "reference", "metric": "consensus_mbp_per_s", "value": val, "unit": "Mbp/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic", "targets_per_s": sample / sec,
            "config": workload_config(args, sample_targets=sample),
            "cpu_baseline": {"value": val, "unit": "Mbp/s", "cores": threads, "kind": kind,
                             "sample": f"{sample} of the workload's targets per step, hot path only (no parsing), "
                                       f"{threads} threads"},
            "e2e": {"value": val, "unit": "Mbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, **extra):
    c = {"workload": f"BASELINE configs[1]: {args.targets} targets x {args.L} bp backbone x {args.cov}x "
                     f"(A={round(args.cov / 0.6)} alignments/target), 15% ins/del/sub error, flags -c {MIN_COV} -m {MIN_LEN} -t {TRIM}",
         "targets_per_gpu": args.targets, "backbone_len": args.L, "coverage": args.cov, "seed": args.seed,
         "l2": "inputs (5.4 GB of alignment strings per GPU at the default size) exceed the 126 MB L2; no flush needed"}
    c.update(extra)
    return c


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import pbdagcon_b200.api as api
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    dev = LOCAL_RANK
    torch.cuda.set_device(dev)
    if WORLD > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
    threads = max(1, host_threads() // max(1, WORLD))

    def barrier():
        torch.cuda.synchronize()
        if WORLD > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if WORLD == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        if WORLD == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- workload: every rank owns a static range of the target stream (weak scaling, no collective on the data path)
    t_gen = time.perf_counter()
    subs = gen_workload(args.seed, args.L, args.cov, RANK * args.targets, args.targets, args.sub, threads)
    import pbdagcon_b200.synth as synth
    # pinned staging buffers for the e2e arm: the arena dominates; small arrays ride along pageable
    pinned = []
    sub_pb = []
    for b in subs:
        pa = api.PinnedArray((b.arena.size,), np.uint8)
        pa.array[:] = b.arena
        b.arena = pa.array                      # drop the pageable copy
        pinned.append(pa)
        sub_pb.append(api.PackedBatch(b.tgt_tlen, b.tgt_aln_begin, b.aln_start, b.aln_len, b.aln_q_off, b.aln_t_off, pa.array))
        # PackedBatch made a contiguous view of the pinned array (no copy): keep pointer identity
        assert sub_pb[-1].arena.ctypes.data == pa.ptr
    whole = api.PackedBatch.from_synth(synth.concat(subs)) if len(subs) > 1 else sub_pb[0]
    gen_s = time.perf_counter() - t_gen
    alg_in = whole.alg_bytes()
    h2d_bytes = sum(int(p.arena.size) + 4 * (2 * p.n_targets + 1) + 24 * p.n_alns + 8 * p.n_alns for p in sub_pb)

    eng = api.Consensus(min_cov=MIN_COV, min_len=MIN_LEN, trim=TRIM, device=dev, targets_per_warp=args.tpw,
                        scratch_bytes=args.scratch_gb << 30, flags=args.flags)
    # ---- arm 1: HBM-resident
    db = eng.upload(whole)
    if len(subs) > 1:
        whole.arena = None                      # the HBM copy is what the resident arm reads
        whole.c.arena = None
    sampler = ClockSampler(dev)
    phases = {}
    launches = 0
    res = None
    for _ in range(args.warmup):
        res = eng.run(db)
    barrier()
    sampler.start()
    span_ms = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = eng.run(db, copy=False)
        tm = eng.timings()
        span_ms.append(tm["ms_span"])
        launches += tm["launches"]
        for k, v in tm.items():
            if k.startswith("ms_"):
                phases[k] = phases.get(k, 0.0) + v
    barrier()
    wall_s = time.perf_counter() - t0
    clocks = sampler.stop()
    res = eng.run(db)                       # a copy of the last results for the parity spot-check / base count
    bases = res.total_bases()
    n_ok = int((res.status == api.T_OK).sum())
    frag_bytes = int(res.seq_len.astype(np.int64).sum()) + 12 * int(res.seq_len.shape[0])
    eng.free(db)
    dev_s = allmax(sum(span_ms) / 1e3)      # CUDA events on the launching stream, max over ranks
    wall_s = allmax(wall_s)
    tot_bases = allsum(bases)
    tot_targets = allsum(args.targets)
    value = tot_bases * args.steps / dev_s / 1e6

    # ---- arm 2: end to end from pinned host buffers through submit/collect (two sub-batches in flight)
    e2e_t0 = [0.0]

    def e2e_pass():
        e2e_t0[0] = time.perf_counter()
        nb = 0
        d2h = 0
        nsub, sent, got = len(sub_pb), 0, 0
        while got < nsub:
            while sent < nsub and sent - got < api.MAX_INFLIGHT:
                eng.submit(sub_pb[sent])
                sent += 1
            r = eng.collect(copy=False)
            got += 1
            if args.e2e_trace:
                tmx = eng.timings()
                print("e2e sub-batch", got, "chunks", tmx["chunks"], "launches", tmx["launches"], {k: round(v, 1) for k, v in tmx.items() if k.startswith("ms_")},
                      "t=%.1f ms" % ((time.perf_counter() - e2e_t0[0]) * 1e3), file=sys.stderr)
            nb += int(np.ctypeslib.as_array(r.frag_seq_len, (r.n_frags,)).astype(np.int64).sum()) if r.n_frags else 0
            d2h += int(r.seq_bytes) + 8 * r.n_frags + 16 * r.n_targets
        return nb, d2h

    for _ in range(max(1, args.warmup - 1)):
        e2e_pass()
    barrier()
    t0 = time.perf_counter()
    e2e_bases = 0
    d2h_bytes = 0
    for _ in range(args.steps):
        nb, d2h_bytes = e2e_pass()
        e2e_bases += nb
    barrier()
    e2e_s = allmax(time.perf_counter() - t0)
    e2e_val = allsum(e2e_bases) / e2e_s / 1e6
    assert e2e_bases == bases * args.steps, "e2e and resident arms disagree on the consensus"

    if RANK != 0:
        eng.close()
        if WORLD > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (HBM-bound integer/byte work; no tensor cores on this path)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kern = {"normalize_trim": phases.get("ms_normalize", 0), "build_graph": phases.get("ms_build", 0),
            "merge_nodes": phases.get("ms_merge", 0), "best_path": phases.get("ms_bestpath", 0)}
    if kern["best_path"] == 0:      # default: both sweeps run as one launch (k_merge_best2); --flags 512 splits them
        kern = {"normalize_trim": kern["normalize_trim"], "build_graph": kern["build_graph"], "merge_best": kern["merge_nodes"]}
    top = max(kern, key=kern.get)
    n_launch_top = args.steps * max(1, tm["chunks"]) if top != "normalize_trim" else args.steps
    alg_bytes_step = alg_in + frag_bytes                 # B_alg summed over the batch
    avg_ms = kern[top] / n_launch_top
    achieved = (alg_bytes_step / max(1, tm["chunks"]) if top != "normalize_trim" else alg_bytes_step) / (avg_ms / 1e3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(top)
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "avg_launch_ms": avg_ms,
                "alg_bytes_per_launch": alg_bytes_step / max(1, tm["chunks"]),
                "kernel_ms_per_step": {k: v / args.steps for k, v in kern.items()},
                "whole_path_frac": alg_bytes_step / (sum(kern.values()) / args.steps / 1e3) / 1e9 / peak}

    # ---- CPU baseline on this box's host cores: bounded sample of the same workload
    cpu = None
    if WORLD == 1 and not args.no_cpu:
        checker, kind = ref_library()
        cores = host_threads()
        sample = min(args.targets, max(cores * 12, 32))
        sec, nb, nt = cpu_consensus(checker, subs, sample, args.L, cores)
        cpu = {"value": nb / sec / 1e6, "unit": "Mbp/s", "cores": cores, "kind": kind, "targets_per_s": nt / sec,
               "sample": f"first {nt} targets of the workload, hot path only, {cores} threads, {sec:.1f} s wall"}
        # parity spot check while we are here: the sample's consensus bases must match the GPU's
        gpu_nb = int(sum(len(f[2]) for t in range(nt) for f in res.frags(t)))
        assert gpu_nb == nb, f"GPU and CPU consensus differ on the sample: {gpu_nb} vs {nb}"
    eng.close()

    line = {"metric": "consensus_mbp_per_s", "value": value, "unit": "Mbp/s", "n_gpus": WORLD, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int32/f32", "data": "synthetic",
            "config": workload_config(args, sub_batch_targets=args.sub, targets_ok=n_ok, gen_seconds=round(gen_s, 1),
                                      targets_per_warp=args.tpw, chunks_per_step=tm["chunks"]),
            "targets_per_s": tot_targets * args.steps / dev_s, "wall_ms_per_step": wall_s * 1e3 / args.steps,
            "e2e": {"value": e2e_val, "unit": "Mbp/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "targets_per_s": tot_targets * args.steps / e2e_s, "ms_per_step": e2e_s * 1e3 / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
    print(json.dumps(line), flush=True)
    if WORLD > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--targets", type=int, default=10000, help="targets per GPU")
    ap.add_argument("--L", type=int, default=8000)
    ap.add_argument("--cov", type=float, default=30)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--sub", type=int, default=0, help="targets per dgc_submit in the e2e arm (0 = the whole step in one "
                    "submit: the library stages the upload itself)")
    ap.add_argument("--tpw", type=int, default=0, help="targets per warp in the serial kernels (0 = auto)")
    ap.add_argument("--scratch-gb", type=int, default=112)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-trace", action="store_true", help="print the device timings of every collected sub-batch")
    ap.add_argument("--flags", type=int, default=0, help="extra dgc_opts.flags (tuning, e.g. 256 = no prefetch)")
    args = ap.parse_args()
    if args.sub <= 0:
        args.sub = args.targets
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
