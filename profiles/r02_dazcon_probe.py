"""Evidence tooling: where the dazcon binary's wall time goes on a generated DB/LAS pair (run under gpurun).
usage: python profiles/r02_dazcon_probe.py [reads] [batch-targets ...]"""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_dazcon as bd
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
work = "/dev/shm/dgc_daz_probe"
db, las, gen = bd.make_pair(n, 3, work, bd.host_threads())
print("pair", n, "reads, gen", round(gen, 1), "s, las", os.path.getsize(las) >> 20, "MiB", flush=True)
for bt in (sys.argv[2:] or ["512", "4096"]):
    for extra in ([], ["--host-realign"]):
        t0 = time.time()
        r = subprocess.run([bd.DAZCON, "-j", str(bd.host_threads()), "-s", db, "-a", las, "--stats", "--batch-targets", bt] + extra,
                           stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
        print("batch-targets", bt, extra, "wall", round(time.time() - t0, 2), r.stderr.strip().split("\n")[-1], flush=True)
r = subprocess.run([bd.DAZCON, "-v", "-j", str(bd.host_threads()), "-s", db, "-a", las, "--batch-targets", "4096"], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
print(r.stderr[-3000:])
