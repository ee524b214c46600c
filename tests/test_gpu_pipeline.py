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


def _oracle_round0(ix, genome, sc, pool, minsc, nceil, seed_len=22, maxgap=15, max_aln=2, pair_boost=False):
    """The batch runtime's chain restated with the CPU oracle stage by stage (ragged reads).
    genome: one code array, or the list of reference sequences (codes 0..3, 4 = N) the index was built over."""
    seqs_ref = genome if isinstance(genome, list) else [genome]
    orc = H.FmOracle(ix, sc.to_struct())
    orc.stage_reads(pool)
    sweep = orc.exact_sweep()
    # paired driver: interval boosted by 20 % (bt2_search.cpp:3392-3395)
    ival = (lambda L: max(int(fmio.seed_interval(L) * 1.2 + 0.5), 1)) if pair_boost else fmio.seed_interval
    sj, _ = fmio.instantiate_exact_seeds(pool, seed_len, interval_fn=ival)
    sh = orc.seed_exact(sj) if len(sj) else np.zeros(0, dtype=structs.seed_hit)
    hit = np.nonzero(sh["botf"] > sh["topf"])[0]
    xj = np.zeros(len(hit), dtype=structs.extend_job)
    for f in ("read", "off", "len", "fw"):
        xj[f] = sj[f][hit]
    for f in ("topf", "botf", "topb", "botb"):
        xj[f] = sh[f][hit]
    xr = orc.extend(xj) if len(xj) else np.zeros(0, dtype=structs.extend_result)
    picks = []
    h = 0
    while h < len(xj):
        e, best = h, h
        while e < len(xj) and xj["read"][e] == xj["read"][h]:
            sz, bsz = int(xj["botf"][e]) - int(xj["topf"][e]), int(xj["botf"][best]) - int(xj["topf"][best])
            ex, bex = int(xr["nlex"][e]) + int(xr["nrex"][e]), int(xr["nlex"][best]) + int(xr["nrex"][best])
            if sz < bsz or (sz == bsz and ex > bex):
                best = e
            e += 1
        picks.append(best)
        h = e
    picks = np.array(picks, dtype=np.int64)
    offs, steps = orc.locate(xj["topf"][picks]) if len(picks) else (np.zeros(0, np.uint32), np.zeros(0, np.uint32))
    b_seq, b_qual, b_win, jobs, rid, starts, fws, tidxs = [], [], [], np.zeros(len(picks), dtype=structs.dp_job), [], [], [], []
    fo = H.FrameOracle()
    ro = wo = 0
    comp = np.array([3, 2, 1, 0, 4], np.uint8)
    for k, p in enumerate(picks):
        x = xj[p]
        ri = int(x["read"]); rdlen = int(pool.reads["rd_len"][ri]); o = int(pool.reads["rd_off"][ri])
        rd, qu = pool.seq[o:o + rdlen], pool.qual[o:o + rdlen]
        if not x["fw"]:
            rd, qu = comp[rd[::-1]], qu[::-1]
        seedoff = int(x["off"]) if x["fw"] else rdlen - int(x["off"]) - int(x["len"])
        # Ebwt::joinedToTextOff of the located hit (straddlers are kept outside eeMode), then the window out of that
        # sequence with N in its gaps and off its ends (BitPairReference::getStretch)
        tidx, toff, tlen, _ = fo.joined_to_text_off(ix.fw, int(x["len"]), int(offs[k]), 0)
        g = seqs_ref[tidx]
        G = len(g)
        assert tlen == G
        start = toff - seedoff
        ok, rect = fo.frame_seed_ext(start, rdlen, G, maxgap, maxgap, int(nceil[ri]), maxgap, 1)   # trimToRef (the reference default)
        assert ok
        refl, ncol = rect[0], rect[1] - rect[0] + 1
        cols = refl + np.arange(ncol + 1)
        codes = np.where((cols >= 0) & (cols < G), g[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
        b_seq.append(rd); b_qual.append(qu); b_win.append((1 << codes).astype(np.uint8))
        j = jobs[k]
        j["rd_off"], j["rd_len"], j["rf_off"], j["ncol"] = ro, rdlen, wo, ncol
        j["minsc"], j["nceil"], j["refl"], j["triml"] = minsc[ri], nceil[ri], refl, rect[4]
        j["corel"], j["corer"], j["fw"], j["enable8"] = rect[6], rect[7], x["fw"], 1
        ro += rdlen; wo += ncol + 1
        rid.append(ri); starts.append(start); fws.append(int(x["fw"])); tidxs.append(tidx)
    cat = lambda a: np.ascontiguousarray(np.concatenate(a)) if a else np.zeros(0, np.uint8)
    ref = H.oracle_dp(sc.to_struct(), cat(b_seq), cat(b_qual), cat(b_win), jobs, max_aln=max_aln) if len(jobs) else None
    return dict(sweep=sweep, nseeds=len(sj), rid=np.array(rid), start=np.array(starts), fw=np.array(fws), tidx=np.array(tidxs),
                dp=ref, njobs=len(jobs),
                steps=int(steps.sum()) if len(picks) else 0, jobs=jobs)


@pytest.mark.parametrize("local", [False, True])
def test_batch_runtime_ragged_reads_match_oracle(local):
    # reads of every length from 1 up (shorter than the seed, shorter than the ftab jump), all-N reads,
    # reads off both ends of the text: the device glue (seed instantiation, pick, compaction, framing,
    # device-side DP staging with several stripe classes) against the oracle chained stage by stage
    rng = np.random.default_rng(7)
    genome = synth.random_genome(rng, 1 << 18)
    ix, _ = build.build_index([genome], ["chr1"])
    sc = scoring.Scoring(local=local)      # local: G,20,8 minimum scores, match bonus 2, u8 with i16 re-runs
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    dix = lib.DeviceIndex(ix, 0)
    ctx.bind_index(dix)
    dref = lib.DeviceRef(genome)
    seqs, quals = [], []
    lens = list(range(1, 40)) + [int(x) for x in rng.integers(22, 420, size=300 if local else 700)]
    for i, L in enumerate(lens):
        p = int(rng.integers(0, len(genome) - L))
        if i % 97 == 0:
            p = 0 if (i // 97) % 2 == 0 else len(genome) - L          # windows hang off the text
        r = genome[p:p + L].copy()
        mm = rng.random(L) < 0.02
        r[mm] = rng.integers(0, 4, size=int(mm.sum()))
        if i % 11 == 0:
            r[rng.integers(0, L)] = 4
        if i % 53 == 0:
            r[:] = 4
        if rng.random() < 0.5:
            r = np.array([3, 2, 1, 0, 4], np.uint8)[r[::-1]]
        seqs.append(r.astype(np.uint8)); quals.append((rng.integers(2, 41, size=L) + 33).astype(np.uint8))
    pool = fmio.ReadPool.__new__(fmio.ReadPool)
    pool.seq, pool.qual = np.concatenate(seqs), np.concatenate(quals)
    pool.reads = np.zeros(len(lens), dtype=structs.read)
    pool.reads["rd_len"] = lens
    pool.reads["rd_off"] = np.concatenate([[0], np.cumsum(lens)[:-1]])
    minsc = np.array([max(1, sc.score_min.f_int(L)) if local else sc.score_min.f_int(L) for L in lens], dtype=np.int32)
    nceil = np.array([sc.nceil(L) for L in lens], dtype=np.int32)
    want = _oracle_round0(ix, genome, sc, pool, minsc, nceil)
    out, st = ctx.hotpath_round0(dref, pool.seq, pool.qual, pool.reads, minsc, nceil)
    nj = out["njobs"]
    assert out["sweep"].tobytes() == want["sweep"].tobytes()
    assert st["seeds"] == want["nseeds"] and nj == want["njobs"] and nj > (200 if local else 400)
    assert np.array_equal(out["job_read"][:nj], want["rid"]) and np.array_equal(out["job_start"][:nj], want["start"])
    assert np.array_equal(out["job_fw"][:nj], want["fw"]) and st["locate_steps"] == want["steps"]
    got = ctx.dp_fetch()
    for i in range(nj):
        assert dpio.results_to_py(*want["dp"], i) == dpio.results_to_py(got[0], got[1], got[2], got[3], i), i
    # an empty batch and a batch without a single seed are not errors
    e = np.zeros(0, dtype=np.uint8)
    out0, st0 = ctx.hotpath_round0(dref, np.zeros(1, np.uint8), np.zeros(1, np.uint8), np.zeros(0, dtype=structs.read),
                                   np.zeros(0, np.int32), np.zeros(0, np.int32))
    assert out0["njobs"] == 0
    rdN = np.zeros(2, dtype=structs.read); rdN["rd_len"] = 30; rdN["rd_off"] = [0, 30]
    outN, stN = ctx.hotpath_round0(dref, np.full(60, 4, np.uint8), np.full(60, 73, np.uint8), rdN,
                                   np.full(2, -18, np.int32), np.full(2, 1, np.int32))
    assert outN["njobs"] == 0 and stN["seeds"] == 0
    ctx.close()


def test_batch_runtime_multiple_sequences_with_gaps():
    """IDX-8: the device reference carries the `.3.bt2` records.  Five sequences, two of them with N stretches inside
    (several fragments each) and one with leading / trailing Ns; reads sampled across fragment and sequence ends:
    joinedToTextOff, per-sequence framing and the windows (N in the gaps and off the ends) against the oracle chain."""
    rng = np.random.default_rng(23)
    seqs = [synth.random_genome(rng, n).astype(np.uint8) for n in (60000, 41000, 90000, 3000, 52000)]
    seqs[1][15000:15040] = 4; seqs[1][30000:30003] = 4
    seqs[2][:25] = 4; seqs[2][-60:] = 4; seqs[2][50000:50200] = 4
    ix, extras = build.build_index(seqs, ["s%d" % i for i in range(len(seqs))])
    sc = scoring.Scoring(local=False)
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    dix = lib.DeviceIndex(ix, 0)
    ctx.bind_index(dix)
    dref = lib.DeviceRef(extras["joined"], 0, recs=extras["recs"])
    rs, qs, lens = [], [], []
    comp = np.array([3, 2, 1, 0, 4], np.uint8)
    spots = [(1, 14900), (1, 15030), (1, 29950), (2, 0), (2, 30), (2, 89850), (2, 49900), (2, 50150), (3, 0), (3, 2850), (0, 59850), (4, 0)]
    for i in range(900):
        L = int(rng.integers(30, 260))
        if i < 20 * len(spots):
            t, p = spots[i % len(spots)]
            p = max(0, min(p + int(rng.integers(-40, 40)), len(seqs[t]) - L))
        else:
            t = int(rng.integers(0, len(seqs))); p = int(rng.integers(0, len(seqs[t]) - L))
        r = seqs[t][p:p + L].copy()
        mm = rng.random(L) < 0.015
        r[mm] = rng.integers(0, 4, size=int(mm.sum()))
        if rng.random() < 0.5:
            r = comp[r[::-1]]
        rs.append(r.astype(np.uint8)); qs.append((rng.integers(2, 41, size=L) + 33).astype(np.uint8)); lens.append(L)
    pool = fmio.ReadPool.__new__(fmio.ReadPool)
    pool.seq, pool.qual = np.concatenate(rs), np.concatenate(qs)
    pool.reads = np.zeros(len(lens), dtype=structs.read)
    pool.reads["rd_len"] = lens
    pool.reads["rd_off"] = np.concatenate([[0], np.cumsum(lens)[:-1]])
    minsc = np.array([sc.score_min.f_int(L) for L in lens], dtype=np.int32)
    nceil = np.array([sc.nceil(L) for L in lens], dtype=np.int32)
    want = _oracle_round0(ix, seqs, sc, pool, minsc, nceil)
    out, st = ctx.hotpath_round0(dref, pool.seq, pool.qual, pool.reads, minsc, nceil)
    nj = out["njobs"]
    assert nj == want["njobs"] and nj > 500 and len(set(want["tidx"].tolist())) == len(seqs)
    assert np.array_equal(out["job_read"][:nj], want["rid"]) and np.array_equal(out["job_start"][:nj], want["start"])
    assert np.array_equal(ctx.hotpath_job_tidx(nj), want["tidx"])
    got = ctx.dp_fetch()
    for i in range(nj):
        assert dpio.results_to_py(*want["dp"], i) == dpio.results_to_py(got[0], got[1], got[2], got[3], i), i
    ctx.close()
