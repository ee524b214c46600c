"""CPU suite, world_size 2 over gloo: sharding the read batch across ranks (index replicated,
no data-path collective) gives the same per-read results, in read order, as a single rank.
The per-rank engine here is the CPU oracle (this is a test of the host-side sharding logic)."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
from bt2gpu import fmio, scoring, shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _seed_ranges(recs):
    sc = scoring.Scoring(local=False)
    eng = H.FmOracle(H.load_index_fixture("lambda_index.npz"), sc.to_struct())
    pool = fmio.ReadPool([r["seq"] for r in recs], [r["qual"] for r in recs])
    eng.stage_reads(pool)
    jobs, _ = fmio.instantiate_exact_seeds(pool, 22)
    hits = eng.seed_exact(jobs)
    sw = eng.exact_sweep()
    per_read = [[] for _ in recs]
    for j, h in zip(jobs, hits):
        per_read[int(j["read"])].append((int(j["fw"]), int(j["off"]), int(h["topf"]), int(h["botf"]), int(h["topb"]), int(h["botb"])))
    return [(per_read[i], sw[i].tobytes()) for i in range(len(recs))]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    recs = fmio.parse_seed_dump(os.path.join(H.GOLDEN, "fm_lambda_r1.txt"), limit=61)
    lo, hi = shard.shard_bounds(len(recs), rank, world)
    mine = _seed_ranges(recs[lo:hi])
    allr = shard.gather_in_order(mine, world, dist)
    if rank == 0:
        q.put(allr)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_equals_single_rank():
    recs = fmio.parse_seed_dump(os.path.join(H.GOLDEN, "fm_lambda_r1.txt"), limit=61)
    single = _seed_ranges(recs)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = q.get(timeout=120)
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got == single


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 64, 65537):
        for w in (1, 2, 3, 8):
            b = [shard.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
