"""Evidence tooling (run under gpurun): whole-process wall clock of the drop-in CLIs.
  m5:   10 000 targets x 8 kb x 30x (8 GB of text) through pbdagcon, and through the reference binary (-j nproc)
  pre:  -a on 'pre' records (ungapped segments), aligned on the device (default) and on the host threads (--host-align)
usage: python profiles/r02_cli_probe.py [m5_targets] [pre_targets] [run_reference 0|1]"""
import hashlib, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pbdagcon_b200.synth as synth
from bench import canon_md5, host_threads
CLI = os.path.join(ROOT, "pbdagcon_b200", "host", "pbdagcon")
REF = os.path.join(ROOT, "oracle", "_ref", "pbdagcon_ref")
nm5 = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
npre = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
runref = int(sys.argv[3]) if len(sys.argv) > 3 else 1
thr = host_threads()


def timed(cmd):
    t0 = time.perf_counter()
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    dt = time.perf_counter() - t0
    assert r.returncode == 0, r.stderr.decode()[-400:]
    return dt, r.stdout, r.stderr.decode().strip().split("\n")[-1]


if nm5:
    path = "/dev/shm/probe.m5"
    t0 = time.time(); nb = synth.write_m5(2, 8000, 30, 0, nm5, path)
    print(f"m5: {nm5} targets, {nb / 1e9:.2f} GB, generated in {time.time() - t0:.1f} s", flush=True)
    best = None
    for rep in range(3):
        dt, out, st = timed([CLI, "-c", "8", "-j", str(thr), "--stats", path])
        md5, nrec, bases = canon_md5(out)
        print(f"  pbdagcon -j {thr}: wall {dt:.2f} s = {bases / dt / 1e6:.1f} Mbp/s whole process; {st}", flush=True)
        best = dt if best is None else min(best, dt)
    if runref:
        dt, out, _ = timed([REF, "-c", "8", "-j", str(thr), path])
        rmd5, rrec, rbases = canon_md5(out)
        print(f"  pbdagcon_ref -j {thr}: wall {dt:.2f} s = {rbases / dt / 1e6:.2f} Mbp/s; records identical: {rmd5 == md5}; ratio {dt / best:.1f}x", flush=True)
    os.unlink(path)
if npre:
    path = "/dev/shm/probe.pre"
    t0 = time.time()
    with open(path, "wb") as f:
        for t in range(npre):
            b = synth.gen_batch(2, 8000, 30, t, 1)
            for k, (q, tt, start) in enumerate(b.target_alns(0)):
                qs, ts = q.replace(b"-", b""), tt.replace(b"-", b"")
                f.write(b"r%d_%d synth/%d/0_8000 + 8000 %d %d %s %s\n" % (t, k, t, start - 1, start - 1 + len(ts), qs, ts))
    print(f"pre: {npre} targets, {os.path.getsize(path) / 1e9:.2f} GB, generated in {time.time() - t0:.1f} s", flush=True)
    outs = []
    for extra in ([], ["--host-align"]):
        for rep in range(2):
            dt, out, st = timed([CLI, "-a", "-c", "8", "-j", str(thr), "--stats"] + extra + [path])
        md5, nrec, bases = canon_md5(out)
        outs.append(md5)
        print(f"  pbdagcon -a {' '.join(extra) or '(device align)'}: wall {dt:.2f} s = {bases / dt / 1e6:.1f} Mbp/s; {st}", flush=True)
    print("  same FASTA:", outs[0] == outs[1])
    os.unlink(path)
