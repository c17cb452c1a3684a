"""dgc_pre_align_noalloc (include/dagcon_b200.h): the `-a` aligner with its scratch in a static __device__ array, for
callers that must not allocate.  No counterpart in the reference.

Checked through the C ABI against the existing entry point (dgc_debug_pre_strings, i.e. dgc_submit_pre's aligner):
same gapped strings, same Alignment::start, byte for byte (integer work: no tolerance), on calls in a row that share
the static array -- different batches, so the second call runs over the first one's leftovers.  Inputs and outputs
are device memory the TEST owns (torch tensors); the entry point itself allocates nothing.

Collected last on purpose."""
import ctypes as C
import random

import numpy as np
import pytest

from test_simple_align_cpu import mutate, revcomp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import pbdagcon_b200.api as api
    return api


def make_targets(seed, n_targets, per_target, lengths):
    rng = random.Random(seed)
    targets = []
    for ti in range(n_targets):
        L = rng.choice(lengths)
        tgt = bytes(rng.choice(b"ACGT") for _ in range(L))
        recs = []
        for r in range(per_target):
            span = rng.randint(L // 3, L)
            s = rng.randint(0, L - span)
            seg = tgt[s:s + span]
            kind = (ti * per_target + r) % 9
            q = mutate(rng, seg, rng.choice([0.05, 0.15, 0.25]))
            if kind == 5:      # a long insertion in the read: the band has to double
                k = len(q) // 2
                q = q[:k] + bytes(rng.choice(b"ACGT") for _ in range(rng.choice([90, 200, 600]))) + q[k:]
            if kind == 6:      # a long deletion
                k = len(q) // 3
                q = q[:k] + q[k + rng.choice([100, 300]):]
            if kind == 7:      # unrelated sequence
                q = bytes(rng.choice(b"ACGT") for _ in range(len(seg)))
            if kind == 8 and ti % 2:   # empty query segment
                q = b""
            recs.append((revcomp(q), revcomp(seg), "-", L - (s + span)) if r % 2 else (q, seg, "+", s))
        targets.append((L, recs))
    return targets


class DevCall:
    """One dgc_pre_align_noalloc call: device tensors for every input and output, then the results as
    [(q, t, start1)] in dgc_debug_pre_strings' form."""

    def __init__(self, api, targets):
        import torch
        self.torch = torch
        pb = api.PreBatch(targets)
        tgt_tlen, rec_begin, start, strand, qlen, tseglen, q_off, t_off = pb.a
        R = pb.n_recs
        rec_tlen = np.repeat(tgt_tlen, np.diff(rec_begin.astype(np.int64))).astype(np.uint32)
        cap = (qlen.astype(np.int64) + tseglen.astype(np.int64) + 15) & ~15
        out_q_off = np.zeros(R, np.uint64)
        out_t_off = np.zeros(R, np.uint64)
        pos = 0
        for k in range(R):
            out_q_off[k] = pos; pos += int(cap[k])
            out_t_off[k] = pos; pos += int(cap[k])
        dev = lambda x: torch.from_numpy(x.view(np.int32 if x.dtype == np.uint32 else np.int64 if x.dtype == np.uint64 else x.dtype).copy()).cuda()
        self.R = R
        self.keep = [dev(x) for x in (start, strand, qlen, tseglen, rec_tlen, q_off, t_off, pb.arena, out_q_off, out_t_off)]
        self.out_arena = torch.full((pos + 64,), 0x58, dtype=torch.uint8, device="cuda")
        self.outs = [torch.full((R + 1,), -7, dtype=torch.int32, device="cuda") for _ in range(5)]   # len, start, tbegin, tend, state
        a = api.CPreDev()
        a.n_recs = R
        (a.rec_start, a.rec_strand, a.rec_qlen, a.rec_tseglen, a.rec_tlen, a.rec_q_off, a.rec_t_off, a.arena, a.out_q_off,
         a.out_t_off) = [t.data_ptr() for t in self.keep]
        a.out_arena = self.out_arena.data_ptr()
        a.out_len, a.out_start, a.out_tbegin, a.out_tend, a.out_state = [t.data_ptr() for t in self.outs]
        self.args = a
        self.out_q_off, self.out_t_off = out_q_off, out_t_off
        torch.cuda.synchronize()

    def run(self, api):
        err = C.create_string_buffer(512)
        rc = api.load_library().dgc_pre_align_noalloc(0, C.byref(self.args), None, err, 512)
        return rc, err.value.decode()

    def results(self):
        self.torch.cuda.synchronize()
        ln, st, _, _, state = [t.cpu().numpy().view(np.uint32) for t in self.outs]
        arena = self.out_arena.cpu().numpy()
        assert set(state[:self.R].tolist()) <= {1, 3}, sorted(set(state[:self.R].tolist()))
        out = []
        for k in range(self.R):
            qo, to, n = int(self.out_q_off[k]), int(self.out_t_off[k]), int(ln[k])
            out.append((arena[qo:qo + n].tobytes(), arena[to:to + n].tobytes(), int(st[k])))
        return out


def existing(api, targets):
    eng = api.Consensus(min_cov=1, min_len=1, trim=0)
    try:
        return eng.debug_pre_strings(api.PreBatch(targets))
    finally:
        eng.close()


def test_noalloc_entry_point_matches_the_existing_one_on_calls_in_a_row(api):
    # batch A: 200 records up to 7 kb (their move matrices exceed the 64 MiB scratch together: several waves; some need bands
    # 128..1024); batch B: other records, other sizes, run straight after A over A's leftovers; then A again.
    tA = make_targets(77, 20, 10, [400, 1500, 3000, 7000])
    tB = make_targets(78, 30, 5, [300, 900, 5000])
    wantA, wantB = existing(api, tA), existing(api, tB)
    assert sum(1 for e in wantA if e[0]) > 150 and sum(1 for e in wantA if not e[0]) >= 1
    for name, targets, want in (("A", tA, wantA), ("B", tB, wantB), ("A again", tA, wantA)):
        call = DevCall(api, targets)
        rc, err = call.run(api)
        assert rc == 0, (name, rc, err)
        got = call.results()
        assert len(got) == len(want)
        bad = [k for k, (g, w) in enumerate(zip(got, want)) if g != w]
        assert not bad, (name, bad[:5], got[bad[0]][2], want[bad[0]][2], len(got[bad[0]][0]), len(want[bad[0]][0]))


def test_noalloc_entry_point_states_its_limits(api):
    import torch
    lib = api.load_library()
    err = C.create_string_buffer(512)
    a = api.CPreDev()
    a.n_recs = api.NOALLOC_MAX_RECS + 1
    assert lib.dgc_pre_align_noalloc(0, C.byref(a), None, err, 512) == -1 and b"DGC_NOALLOC_MAX_RECS" in err.value
    a.n_recs = 3
    assert lib.dgc_pre_align_noalloc(0, C.byref(a), None, err, 512) == -1 and b"null device pointer" in err.value
    a.n_recs = 0
    assert lib.dgc_pre_align_noalloc(0, C.byref(a), None, err, 512) == 0
    # one record whose move matrix alone is larger than the static scratch: (n + 1) * 129 bytes at band 64 > 64 MiB
    n = 530_000
    rng = np.random.default_rng(5)
    seq = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, n)].tobytes()
    call = DevCall(api, [(n, [(seq, seq, "+", 0)])])
    rc, msg = call.run(api)
    assert rc == -1 and "beyond the static scratch" in msg and "dgc_submit_pre" in msg, (rc, msg)
    torch.cuda.synchronize()
    # ... and the entry point still works after a refused call
    t = make_targets(5, 2, 4, [500])
    call = DevCall(api, t)
    rc, msg = call.run(api)
    assert rc == 0, msg
    assert call.results() == existing(api, t)
