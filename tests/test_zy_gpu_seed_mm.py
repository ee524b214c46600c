"""GPU: bt2gpu_seed_mm_batch -- the multi-zone seeds of -N 1 / -N 2 (Seed::oneMmSeeds / twoMmSeeds searched by
searchSeedBi with its mismatch branches) -- through the C ABI, bit for bit against the dumps of the unmodified reference
(ref_seed_dump --seedmms: every BW range of every seed offset in both indexes, the reference string each stands for, in
the order of the reference's reportHit calls, and the read's bwops total) and against the CPU oracle on random reads.
-N 0 through the same entry point must agree with bt2gpu_seed_exact_batch."""
import os

import numpy as np
import pytest

import helpers as H
from bt2gpu import fmio, lib, scoring, structs
from test_oracle_fm import MM_CASES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


_IDX = {}


def dev_index(name):
    if name not in _IDX:
        ix = H.load_index_fixture(name)
        _IDX[name] = (ix, lib.DeviceIndex(ix, 0))
    return _IDX[name]


def _jobs_of(recs, seed_len):
    """One job per (read, strand, offset) of the dump + where each read's jobs start."""
    jobs, start = [], []
    for ri, r in enumerate(recs):
        sl, interval, noffs, _ = r["SD"]
        start.append(len(jobs))
        L = min(seed_len, len(r["seq"]))
        for fw in (1, 0):
            for i in range(noffs):
                jobs.append((ri, i * interval, L, fw))
    start.append(len(jobs))
    a = np.zeros(len(jobs), dtype=structs.seed_job)
    for k, (ri, off, L, fw) in enumerate(jobs):
        a[k] = (ri, off, L, fw)
    return a, start


@pytest.mark.parametrize("dump,idx,mms,seed_len", MM_CASES)
def test_cuda_multizone_seeds_match_reference_dump(ctx, dump, idx, mms, seed_len):
    sc = scoring.Scoring(local=False)
    ctx.set_scoring(sc)
    ix, dix = dev_index(idx)
    ctx.bind_index(dix)
    recs = fmio.parse_seed_dump(os.path.join(H.GOLDEN, dump))
    pool = fmio.ReadPool([r["seq"] for r in recs], [r["qual"] for r in recs])
    ctx.stage_reads(pool)
    jobs, start = _jobs_of(recs, seed_len)
    res, hits = ctx.seed_mm(jobs, mms)
    nranges = 0
    for ri, r in enumerate(recs):
        sl, interval, noffs, bwops = r["SD"]
        if noffs == 0:
            continue
        S = {(fw, i): rng for fw, i, off, rng in r["S"]}
        K = {(fw, i): keys for fw, i, keys in r["K"]}
        a, b = start[ri], start[ri + 1]
        assert int(res["bwops"][a:b].sum()) == bwops, (ri, int(res["bwops"][a:b].sum()), bwops)
        for j in range(a, b):
            key = (int(jobs["fw"][j]), (j - a) % noffs)
            want = [rg + (k[0],) for rg, k in zip(S[key], K[key])]
            h = hits[int(res["hit_off"][j]):int(res["hit_off"][j]) + int(res["nhits"][j])]
            got = [(int(x["topf"]), int(x["botf"]), int(x["topb"]), int(x["botb"]), int(x["key"])) for x in h]
            assert got == want, (ri, key, got, want)
            nranges += len(want)
    assert nranges > 300


def test_cuda_multizone_seeds_match_oracle_random_and_exact_entry_point(ctx):
    rng = np.random.default_rng(99)
    sc = scoring.Scoring(local=False)
    ctx.set_scoring(sc)
    ix, dix = dev_index("rep_index.npz")
    ctx.bind_index(dix)
    recs = fmio.parse_seed_dump(os.path.join(H.GOLDEN, "fm_rep.txt"))
    seqs, quals = [], []
    for i in range(400):                     # pieces of reads that hit the repeat genome, with 0..3 substitutions and the odd N
        s = list(recs[int(rng.integers(0, len(recs)))]["seq"])
        for _ in range(int(rng.integers(0, 4))):
            s[int(rng.integers(0, len(s)))] = "ACGTN"[int(rng.integers(0, 5))]
        seqs.append("".join(s))
        quals.append("".join(chr(33 + int(x)) for x in rng.integers(2, 41, size=len(s))))
    pool = fmio.ReadPool(seqs, quals)
    ctx.stage_reads(pool)
    orc = H.FmOracle(ix, sc.to_struct())
    orc.stage_reads(pool)
    for mms, sl, per in ((1, 22, 9), (2, 20, 6), (1, 12, 5), (2, 9, 4), (0, 22, 7), (1, 32, 11)):
        sj, _ = fmio.instantiate_exact_seeds(pool, sl, lambda n: per)      # (read, off, len, fw) of every offset, both strands
        want, ops = orc.seed_mm(sj, mms, cap=512 * len(sj))
        res, hits = ctx.seed_mm(sj, mms, hit_cap=512 * len(sj))
        assert int(res["bwops"].astype(np.int64).sum()) == ops, (mms, sl)
        for j in range(len(sj)):
            h = hits[int(res["hit_off"][j]):int(res["hit_off"][j]) + int(res["nhits"][j])]
            got = [(int(x["topf"]), int(x["botf"]), int(x["topb"]), int(x["botb"]), int(x["key"])) for x in h]
            assert got == want[j], (mms, sl, j, got, want[j])
        if mms == 0:
            ex = ctx.seed_exact(sj)
            for j in range(len(sj)):
                n = int(res["nhits"][j])
                assert n == (1 if ex["botf"][j] > ex["topf"][j] else 0)
                if n:
                    x = hits[int(res["hit_off"][j])]
                    assert (int(x["topf"]), int(x["botf"]), int(x["topb"]), int(x["botb"])) == tuple(int(ex[k][j]) for k in ("topf", "botf", "topb", "botb"))
    # a pool too small for an offset's hits is refused with the error the shim retries on
    sj, _ = fmio.instantiate_exact_seeds(pool, 12, lambda n: 5)
    with pytest.raises(lib.Bt2GpuError, match="seed hit pool overflow"):
        ctx.seed_mm(sj, 2, hit_cap=len(sj))
