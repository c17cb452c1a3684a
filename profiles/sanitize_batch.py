"""One 200-target fuzz batch through the C ABI, checked against the oracle -- the program compute-sanitizer runs
(profiles/r02_capture.sh: --tool memcheck, --tool racecheck, --tool initcheck).  Test/evidence tooling only."""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import pbdagcon_b200.api as api  # noqa: E402
import pbdagcon_b200.synth as synth  # noqa: E402

rng = random.Random(2024)
targets = []
for i in range(180):
    L, bb, alns = helpers.rand_target(rng, err=rng.choice([0.05, 0.15, 0.3]), full=(i % 4 == 0), dots=(i % 7 == 0))
    targets.append((L, None, alns))
head = synth.from_python(targets)
tail = synth.gen_batch(seed=3, L=2500, cov=25, t0=0, nt=20)      # bubbles, out-merges, pool lists
wide = synth.gen_batch(seed=4, L=9000, cov=12, t0=0, nt=2)       # beyond the shared-memory accumulators (gacc path)
pb = api.PackedBatch.from_synth(synth.concat([head, tail, wide]))
eng = api.Consensus(min_cov=2, min_len=20, trim=3)
res = eng(pb)
os.environ["DGC_STAGE_MIN_BYTES"] = "1"
os.environ["DGC_STAGE_PIECE_BYTES"] = str(1 << 18)
eng2 = api.Consensus(min_cov=2, min_len=20, trim=3, scratch_bytes=8 << 20)   # staged upload, several chunks
eng2.submit(pb)
res2 = eng2.collect()
o = helpers.checker()
bad = 0
for t, (L, _, alns) in enumerate(targets):
    if len(alns) < 2:
        continue
    exp = o.consensus(L, alns, mode=helpers.MODE_PIPELINE, trim=3, min_weight=2, min_len=20)["frags"]
    bad += res.frags(t) != exp or res2.frags(t) != exp
for t in range(20):
    exp = o.consensus(2500, tail.target_alns(t), mode=helpers.MODE_PIPELINE, trim=3, min_weight=2, min_len=20)["frags"]
    bad += res.frags(180 + t) != exp or res2.frags(180 + t) != exp
for t in range(2):
    exp = o.consensus(9000, wide.target_alns(t), mode=helpers.MODE_PIPELINE, trim=3, min_weight=2, min_len=20)["frags"]
    bad += res.frags(200 + t) != exp or res2.frags(200 + t) != exp
eng.close()
eng2.close()
print(f"sanitize_batch: {res.n_targets} targets, {res.total_bases()} consensus bases, mismatches vs reference: {bad}")
sys.exit(1 if bad else 0)
