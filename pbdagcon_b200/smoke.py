"""smoke(): one small invocation of the hot path on cuda:0 through the C ABI, checked against the oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import Oracle, MODE_PIPELINE  # the checker, never the product path
    from . import api, synth
    sb = synth.gen_batch(seed=7, L=3000, cov=20, t0=0, nt=4)
    eng = api.Consensus(min_cov=8, min_len=500, trim=50, device=0)
    res = eng(api.PackedBatch.from_synth(sb))
    tm = eng.timings()
    eng.close()
    o = Oracle()
    for t in range(4):
        exp = o.consensus(3000, sb.target_alns(t), mode=MODE_PIPELINE, trim=50, min_weight=8, min_len=500)
        assert res.frags(t) == exp["frags"], f"smoke: target {t} differs from the oracle"
    print(f"smoke ok: 4 targets, {res.total_bases()} consensus bases, {tm['launches']} kernel launches, "
          f"{tm['ms_total']:.2f} ms device time")
