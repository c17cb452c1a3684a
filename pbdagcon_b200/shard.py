"""Static sharding of the target stream over ranks and the host-side ordered gather.

Targets are independent (reference main.cpp:130-138: every consumer owns its graph), so a multi-GPU run has no
data-path collective: rank r of G takes a contiguous range of ceil(T/G) targets and rank 0 concatenates the ranks'
FASTA records in rank order, which is input-target order.  The only communication is that gather; under torchrun it
runs over whatever backend the process group has (nccl on the GPU box, gloo in the CPU tests).
"""


def target_range(n_targets, rank, world):
    """[t0, t1) of rank `rank`: contiguous ranges of ceil(T/G) targets; trailing ranks may be short or empty."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    per = -(-n_targets // world)
    t0 = min(n_targets, rank * per)
    return t0, min(n_targets, t0 + per)


def chunk_owner(chunk_index, world):
    """Round-robin dealing of fixed-size chunks (for streams with skewed target sizes, e.g. dazcon reads)."""
    return chunk_index % world


def gather_ordered(local_records, dist=None):
    """Every rank passes the records of its range (a list, in target order); rank 0 gets the concatenation in rank
    order, the other ranks None.  `dist` is torch.distributed (initialised) or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return list(local_records)
    world, rank = dist.get_world_size(), dist.get_rank()
    parts = [None] * world if rank == 0 else None
    dist.gather_object(list(local_records), parts, dst=0)
    if rank != 0:
        return None
    out = []
    for p in parts:
        out.extend(p)
    return out
