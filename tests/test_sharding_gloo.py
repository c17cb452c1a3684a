"""World-size-2 run of the multi-GPU host logic on CPU (gloo): static target ranges + ordered gather.
The consensus itself is the oracle here (no GPU in this suite); what is under test is pbdagcon_b200/shard.py."""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from pbdagcon_b200.shard import target_range, chunk_owner, gather_ordered  # noqa: E402

N_TARGETS, L, COV = 7, 1200, 10   # odd on purpose: rank 0 gets 4 targets, rank 1 gets 3


def records_for(t0, t1):
    import helpers
    import pbdagcon_b200.synth as synth
    o = helpers.Oracle()
    out = []
    if t1 > t0:
        sb = synth.gen_batch(seed=77, L=L, cov=COV, t0=t0, nt=t1 - t0)
        for i in range(t1 - t0):
            r = o.consensus(L, sb.target_alns(i), mode=helpers.MODE_PIPELINE, trim=20, min_weight=4, min_len=100)
            out.extend((t0 + i, a, b, s) for a, b, s in r["frags"])
    return out


def _worker(rank, world, port, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        t0, t1 = target_range(N_TARGETS, rank, world)
        got = gather_ordered(records_for(t0, t1), dist)
        if rank == 0:
            q.put(got)
        else:
            assert got is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_target_ranges_cover_the_stream_once():
    for n in (0, 1, 7, 8, 10000, 1000003):
        for world in (1, 2, 3, 4, 8):
            ranges = [target_range(n, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
                assert a1 == b0 and a0 <= a1 and b0 <= b1
            assert max(b - a for a, b in ranges) - min(b - a for a, b in ranges) <= -(-n // world)
    with pytest.raises(ValueError):
        target_range(10, 2, 2)
    assert [chunk_owner(c, 4) for c in range(6)] == [0, 1, 2, 3, 0, 1]


def test_gather_single_process_is_identity():
    assert gather_ordered([(0, 1, 2, b"A")]) == [(0, 1, 2, b"A")]


def test_world_size_2_gloo_gather_matches_single_rank():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    want = records_for(0, N_TARGETS)
    assert got == want
    assert [r[0] for r in got] == sorted(r[0] for r in got), "records must be in input-target order"
    assert len({r[0] for r in got}) == N_TARGETS
