"""bench.py --config 5: the high-coverage stress of BASELINE configs[4] -- 256 targets x 40 kb x 100x, three rounds of
iterative backbone refinement: the consensus of round r is the backbone of round r+1 (AlnGraphBoost(const std::string&)
form from round 2 on).

Between rounds the reads have to be expressed against the new backbone.  That step belongs to neither side of the
comparison (the reference has no refinement driver): a deterministic harness shared by both (tests/harness_align.c) aligns
old backbone to new consensus (banded global DP) and carries every read's alignment over through that map (hal_lift).
Its time is excluded from the timed region and reported (`config.harness_seconds`).  The three rounds' batches are
prepared once; a step = the three rounds back to back on the GPU.

  value   consensus Mbp/s (bases of all three rounds / device time of all three), batches resident in HBM
  e2e     the same through dgc_submit/dgc_collect from pinned host buffers
  parity  every round: a sample of targets against the reference's own code on the identical alignments
"""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
RANK = int(os.environ.get("RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))


def load_harness():
    so = os.path.join(tempfile.mkdtemp(prefix="dgc_hal_"), "libhal.so")
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "harness_align.c")], check=True)
    lib = C.CDLL(so)
    lib.hal_align.restype = C.c_int64
    lib.hal_align.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, C.c_int64, C.c_char_p, C.c_char_p]
    lib.hal_lift.restype = C.c_int64
    lib.hal_lift.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p,
                             C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    return lib


def backbone_map(lib, old, new, band):
    """map_new[p], next_new[p] of hal_lift from a banded global alignment of the two backbones."""
    qo, to = C.create_string_buffer(len(old) + len(new) + 1), C.create_string_buffer(len(old) + len(new) + 1)
    n = lib.hal_align(old, len(old), new, len(new), band, qo, to)
    if n < 0:
        raise RuntimeError("harness band too narrow")
    a = np.frombuffer(qo.raw[:n], np.uint8)
    b = np.frombuffer(to.raw[:n], np.uint8)
    ia, ib = a != 0x2D, b != 0x2D
    newpos = np.cumsum(ib) - 1                     # index of the new base in / before every column
    before = np.cumsum(ib) - ib                    # new bases strictly in front of the column
    cols = np.nonzero(ia)[0]                       # the column of every old base
    map_new = np.where(ib[cols], newpos[cols], -1).astype(np.int32)
    next_new = np.concatenate([before[cols], [int(ib.sum())]]).astype(np.int32)
    return np.ascontiguousarray(map_new), np.ascontiguousarray(next_new)


def lift_target(lib, alns, old_bb, new_bb, band=2048, min_span=600):
    map_new, next_new = backbone_map(lib, old_bb, new_bb, band)
    out = []
    ns, ne = C.c_int64(), C.c_int64()
    for q, t, start in alns:
        cap = len(q) + len(new_bb) // 8 + 4096
        qo, to = C.create_string_buffer(cap), C.create_string_buffer(cap)
        n = lib.hal_lift(q, t, len(q), start - 1, map_new.ctypes.data, next_new.ctypes.data, new_bb, qo, to, cap, C.byref(ns), C.byref(ne))
        if n < 0:
            raise RuntimeError("hal_lift capacity")
        if ne.value - ns.value < min_span:
            continue
        out.append((qo.raw[:n], to.raw[:n], ns.value + 1))
    return out


def main(args):
    import torch
    import pbdagcon_b200.api as api
    import pbdagcon_b200.synth as synth
    sys.path.insert(0, ROOT)
    from bench import (ClockSampler, roofline_block, gen_workload, host_threads, ref_library, cpu_consensus, frag_digest, Dist,
                       workload_config, MIN_COV, MIN_LEN, TRIM)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    dd = Dist(dev)
    cores = host_threads()
    threads = max(1, cores // max(1, WORLD))
    T, L, rounds = args.targets, args.L, args.rounds
    lib = load_harness()
    t_gen = time.perf_counter()
    sb = synth.concat(gen_workload(args.seed, L, args.cov, RANK * T, T, T, threads))
    # round 1: the generator's alignments, AlnGraphBoost(size_t) form; the backbone the reads were drawn against is the
    # coordinate system the harness maps from
    tgt_alns, tgt_bb = [sb.target_alns(t) for t in range(T)], [None] * T
    old_bb = [synth.backbone(args.seed, L, args.cov, RANK * T + t) for t in range(T)]
    gen_s = time.perf_counter() - t_gen
    eng = api.Consensus(min_cov=MIN_COV, min_len=MIN_LEN, trim=TRIM, device=dev, scratch_bytes=args.scratch_gb << 30, flags=args.flags)
    checker, kind = ref_library()
    harness_s, parity_bad, parity_n, cpu_sec, cpu_bases, cpu_n = 0.0, 0, 0, 0.0, 0, 0
    pbs, pins, round_info = [], [], []
    for r in range(rounds):
        with_bb = r > 0
        b = synth.from_python([(len(tgt_bb[t]) if with_bb else L, tgt_bb[t], tgt_alns[t]) for t in range(T)], with_backbone=with_bb)
        pa = api.PinnedArray((b.arena.size,), np.uint8)
        pa.array[:] = b.arena
        pb = api.PackedBatch(b.tgt_tlen, b.tgt_aln_begin, b.aln_start, b.aln_len, b.aln_q_off, b.aln_t_off, pa.array, b.tgt_bb_off)
        pins.append(pa); pbs.append(pb)
        res = eng(pb)
        assert all(int(s) == api.T_OK for s in res.status), f"round {r + 1}: a target was rejected"
        # parity on a sample, against the reference on the identical alignments (timed on round 1 as the CPU baseline)
        idx = list(range(0, T, max(1, T // max(1, min(T, 2 * threads)))))[:2 * threads]
        work = [(len(tgt_bb[t]) if with_bb else L, tgt_alns[t]) for t in idx]
        sec, exp = cpu_consensus(checker, work, threads, backbones=[tgt_bb[t] for t in idx] if with_bb else None)
        if r == 0:
            cpu_sec, cpu_bases, cpu_n = sec, sum(len(f[2]) for fr in exp for f in fr), len(idx)
        parity_bad += sum(frag_digest(0, e) != frag_digest(res.status[t], res.frags(t)) for t, e in zip(idx, exp))
        parity_n += len(idx)
        frags = [max(res.frags(t), key=lambda f: len(f[2]))[2] for t in range(T)]
        round_info.append({"round": r + 1, "bases": res.total_bases(), "alignments": pb.n_alns, "mean_consensus_len": float(np.mean([len(f) for f in frags]))})
        if r + 1 == rounds:
            break
        th = time.perf_counter()
        with ThreadPoolExecutor(max_workers=threads) as ex:
            new_alns = list(ex.map(lambda t: lift_target(lib, tgt_alns[t], old_bb[t], frags[t]), range(T)))
        harness_s += time.perf_counter() - th
        tgt_alns, tgt_bb, old_bb = new_alns, frags, frags
    alg_in = sum(p.alg_bytes() for p in pbs)
    h2d = sum(int(p.arena.size) + 4 * (2 * p.n_targets + 1) + 8 * p.n_targets + 32 * p.n_alns for p in pbs)
    # ---- resident arm
    dbs = [eng.upload(p) for p in pbs]
    for _ in range(args.warmup):
        for d in dbs:
            eng.run(d, copy=False)
    dd.barrier()
    sampler = ClockSampler(dev)
    sampler.start()
    phases, launches, span, chunks = {}, 0, 0.0, 1
    for _ in range(args.steps):
        for d in dbs:
            eng.run(d, copy=False)
            tm = eng.timings()
            span += tm["ms_span"]; launches += tm["launches"]; chunks = tm["chunks"]
            for k, v in tm.items():
                if k.startswith("ms_"):
                    phases[k] = phases.get(k, 0.0) + v
    dd.barrier()
    clocks = sampler.stop()
    bases = sum(eng.run(d).total_bases() for d in dbs)
    for d in dbs:
        eng.free(d)
    dev_s = dd.allmax(span / 1e3)
    value = dd.allsum(bases) * args.steps / dev_s / 1e6
    # ---- e2e arm
    def e2e():
        nb = d2h = 0
        for p in pbs:
            eng.submit(p)
            r = eng.collect(copy=False)
            nb += int(np.ctypeslib.as_array(r.frag_seq_len, (r.n_frags,)).astype(np.int64).sum()) if r.n_frags else 0
            d2h += int(r.seq_bytes) + 8 * r.n_frags + 16 * r.n_targets
        return nb, d2h
    for _ in range(max(1, args.warmup - 1)):
        e2e()
    dd.barrier()
    t0 = time.perf_counter()
    eb = 0
    for _ in range(args.steps):
        nb, d2h = e2e()
        eb += nb
    dd.barrier()
    e2e_s = dd.allmax(time.perf_counter() - t0)
    assert eb == bases * args.steps
    eng.close()
    par = dd.gather([(RANK, parity_n, parity_bad)])
    if RANK != 0:
        dd.close()
        return
    frag_bytes = bases + 12 * 3 * T
    roofline = roofline_block(args, phases, args.steps * rounds, chunks, (alg_in + frag_bytes) / rounds)
    parity = {"checker": kind, "targets": sum(p[1] for p in par), "mismatches": sum(p[2] for p in par), "ok": all(p[2] == 0 for p in par),
              "ranks": len(par), "what": "every round, a strided sample of targets vs oracle/_ref on the identical (harness-lifted) alignments"}
    assert parity["ok"], parity
    line = {"metric": "consensus_mbp_per_s", "value": value, "unit": "Mbp/s", "n_gpus": WORLD, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32/f32",
            "data": "synthetic",
            "config": workload_config(args, rounds=rounds, round_info=round_info, harness_seconds=round(harness_s, 1), gen_seconds=round(gen_s, 1),
                                      harness="tests/harness_align.c: banded global DP old backbone -> new consensus, reads carried over through that map "
                                              "(hal_lift); shared by both sides, excluded from the timed region",
                                      step="the three rounds back to back (batches prepared once)"),
            "targets_per_s": T * WORLD * rounds * args.steps / dev_s,
            "e2e": {"value": dd.allsum(eb) / e2e_s / 1e6, "unit": "Mbp/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h * rounds,
                    "ms_per_step": e2e_s * 1e3 / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "cpu_baseline": {"value": cpu_bases / cpu_sec / 1e6, "unit": "Mbp/s", "cores": threads, "kind": kind, "targets_per_s": cpu_n / cpu_sec,
                             "sample": f"round 1, {cpu_n} of {T} targets, hot path only, {threads} threads, {cpu_sec:.1f} s wall"},
            "parity": parity}
    print(json.dumps(line), flush=True)
    dd.close()
