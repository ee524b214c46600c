"""GPU: the chained hot path (exactSweep -> seeds -> extend -> locate -> DP) on a freshly built
index of a random genome: every stage equals the CPU oracle on the same inputs, and reads are
located at their simulated origin."""
import numpy as np
import pytest

import helpers as H
from bt2gpu import build, fmio, lib, pipeline, scoring, structs, synth, dpio

pytestmark = pytest.mark.gpu


def test_hot_path_chain_matches_oracle():
    rng = np.random.default_rng(42)
    genome = synth.random_genome(rng, 1 << 20)
    ix, _ = build.build_index([genome], ["chr1"])
    sc = scoring.Scoring(local=False)
    ctx = lib.Context(0)
    dix = lib.DeviceIndex(ix, 0)
    hp = pipeline.HotPath(ctx, dix, genome, sc)
    reads, quals, pos, _ = synth.simulate_reads(rng, genome, 3000, 150)
    rc = rng.random(3000) < 0.5
    reads[rc] = np.array([3, 2, 1, 0, 4], np.uint8)[reads[rc][:, ::-1]]
    quals[rc] = quals[rc][:, ::-1]
    res = hp.step(np.ascontiguousarray(reads), np.ascontiguousarray(quals))
    c, L = hp.counters, hp.last
    assert c["dp_jobs"] > 600 and c["aligned"] > 0.95 * c["dp_jobs"]   # simulate.pl reads are noisy: ~1/3 keep a clean exact seed
    assert (np.abs(L["start"] - pos[L["rid"]]) <= 64).mean() > 0.98
    # oracle on the same stages
    orc = H.FmOracle(ix, sc.to_struct())
    pool = fmio.ReadPool.__new__(fmio.ReadPool)
    pool.seq, pool.qual = reads.reshape(-1), quals.reshape(-1)
    pool.reads = np.zeros(len(reads), dtype=structs.read)
    pool.reads["rd_off"] = np.arange(len(reads)) * 150
    pool.reads["rd_len"] = 150
    orc.stage_reads(pool)
    ctx.stage_reads(pool)
    assert orc.exact_sweep().tobytes() == ctx.exact_sweep().tobytes()
    sj = hp.make_seed_jobs(reads)
    ref_jobs, _ = fmio.instantiate_exact_seeds(pool, 22)
    assert sj.tobytes() == ref_jobs.tobytes()          # vectorised instantiateSeeds == the per-read one
    assert orc.seed_exact(sj).tobytes() == ctx.seed_exact(sj).tobytes()
    # DP results of the chain == oracle DP on the chain's jobs
    jobs = L["jobs"]
    o = np.where(L["fw"][:, None], reads[L["rid"]], np.array([3, 2, 1, 0, 4], np.uint8)[reads[L["rid"]][:, ::-1]])
    q = np.where(L["fw"][:, None], quals[L["rid"]], quals[L["rid"]][:, ::-1])
    cols = jobs["refl"][:, None] + np.arange(211)[None, :]
    G = len(genome)
    codes = np.where((cols >= 0) & (cols < G), genome[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
    ref = H.oracle_dp(sc.to_struct(), np.ascontiguousarray(o.reshape(-1)), np.ascontiguousarray(q.reshape(-1)),
                      np.ascontiguousarray((1 << codes).astype(np.uint8).reshape(-1)), jobs, max_aln=2)
    got = ctx.dp_fetch()          # the chain's batch is still staged: fetch it again with the candidates
    for i in range(len(jobs)):
        assert dpio.results_to_py(*ref, i) == dpio.results_to_py(got[0], got[1], got[2], got[3], i), i
    # the multi-threaded C++ glue (bt2gpu_hotpath_round0_batch) makes the same choices and gets the same answers
    n = len(reads)
    minsc = np.full(n, sc.score_min.f_int(150), dtype=np.int32)
    nceil = np.full(n, sc.nceil(150), dtype=np.int32)
    dref = lib.DeviceRef(genome)
    out, st = ctx.hotpath_round0(dref, np.ascontiguousarray(reads.reshape(-1)), np.ascontiguousarray(quals.reshape(-1)),
                                 pool.reads, minsc, nceil, nthreads=4)
    nj = out["njobs"]
    assert nj == len(jobs) and st["dp_jobs"] == nj and st["seeds"] == len(sj)
    assert np.array_equal(out["job_read"][:nj], L["rid"]) and np.array_equal(out["job_start"][:nj], L["start"])
    assert np.array_equal(out["job_fw"][:nj].astype(bool), L["fw"])
    assert out["sweep"].tobytes() == orc.exact_sweep().tobytes()
    got2 = ctx.dp_fetch()
    for i in range(nj):
        assert dpio.results_to_py(*ref, i) == dpio.results_to_py(got2[0], got2[1], got2[2], got2[3], i), i
    assert st["aligned"] == c["aligned"] and st["seed_bwops"] == c["seed_bwops"] and st["locate_steps"] == c["locate_steps"]
    assert ctx.gather_probe("fw", 1 << 22) > 0
    ctx.close()
