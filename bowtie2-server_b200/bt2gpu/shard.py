"""Read-batch sharding across ranks (one process per GPU, index replicated, no collective on
the data path: SURVEY.md 8e).  Reads are independent -- a read's result is a pure function of
(sequence, qualities, name, --seed), pat.cpp:47-87 -- so rank r takes a contiguous slice and
results are concatenated in rank order (= read order, what OutputQueue restores, outq.cpp:61)."""


def shard_bounds(n, rank, world):
    """Contiguous slice [lo, hi) of n reads owned by `rank`: sizes differ by at most one."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_in_order(local, world, dist=None):
    """Concatenate per-rank result lists in rank order (host-side ordering, the only exchange)."""
    if world == 1 or dist is None:
        return list(local)
    parts = [None] * world
    dist.all_gather_object(parts, list(local))
    out = []
    for p in parts:
        out.extend(p)
    return out
