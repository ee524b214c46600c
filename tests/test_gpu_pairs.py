"""GPU: the paired-end batch runtime (bt2gpu_hotpath_pairs_batch) against the CPU oracle chained stage
by stage: anchor pass == the unpaired chain over all mates; then, for the first alignment of every
anchor job, otherMate -> frameFindMateRect (oracle/frame_oracle.c, pinned on the reference) -> window
-> oracle DP of the opposite mate -> peClassifyPair."""
import numpy as np
import pytest

import helpers as H
from bt2gpu import build, dpio, fmio, lib, scoring, structs, synth
from test_gpu_pipeline import _oracle_round0

pytestmark = pytest.mark.gpu

COMP = np.array([3, 2, 1, 0, 4], np.uint8)
RECT = ("refl", "refr", "refl_pretrim", "refr_pretrim", "triml", "trimr", "corel", "corer", "maxgap")


def _extent(rdlen, a, edits):
    es = edits[a["edit_off"]: a["edit_off"] + a["nedits"]]
    return rdlen - int(a["trim5"]) - int(a["trim3"]) - int((es["type"] == 2).sum()) + int((es["type"] == 1).sum())


def _make_pairs(rng, genome, npairs, lens):
    """FR pairs: mate 1 = fragment start, mate 2 = reverse complement of the fragment end; half the fragments flipped."""
    G = len(genome)
    seqs, quals = [], []
    for i in range(npairs):
        l1, l2 = int(lens[2 * i]), int(lens[2 * i + 1])
        frag = max(int(rng.normal(250, 45)), max(l1, l2) + 5)
        if i % 13 == 0:
            frag = max(l1, l2)                                # one mate contains the other
        if i % 40 == 0:
            p = int(rng.integers(0, 30))                      # close to the left end of the text
        elif i % 40 == 1:
            p = G - frag - int(rng.integers(0, 30))           # ... the right end
        else:
            p = int(rng.integers(0, G - frag))
        f = genome[p:p + frag].copy()
        nerr = int(rng.integers(0, 6))
        for _ in range(nerr):
            f[int(rng.integers(0, frag))] = rng.integers(0, 4)
        if i % 9 == 0:                                        # a deletion in the fragment: gapped mate alignments
            d = int(rng.integers(20, frag - 20))
            f = np.concatenate([f[:d], f[d + 2:]])
        if i % 2:
            f = COMP[f[::-1]]
        m1, m2 = f[:l1], COMP[f[::-1]][:l2]
        seqs += [m1, m2]
        quals += [rng.integers(35, 74, size=len(m1)).astype(np.uint8), rng.integers(35, 74, size=len(m2)).astype(np.uint8)]
    return seqs, quals


@pytest.mark.parametrize("flags,nseq", [((0, 0, 0, 1, 1, 1), 1), ((0, 1, 1, 1, 1, 1), 1), ((0, 0, 0, 0, 0, 1), 1), ((0, 0, 0, 1, 1, 1), 3)])
def test_paired_batch_runtime_matches_oracle(flags, nseq):
    rng = np.random.default_rng(11)
    genome = synth.random_genome(rng, 1 << 18)
    # nseq = 3: the same text cut into three reference sequences (mates are searched in the anchor's sequence,
    # windows stop at its ends; pairs sampled across a cut do not pair up)
    cuts = [0, len(genome)] if nseq == 1 else [0, 80000, 180000, len(genome)]
    refseqs = [genome[cuts[i]:cuts[i + 1]] for i in range(len(cuts) - 1)]
    ix, extras = build.build_index(refseqs, ["chr%d" % (i + 1) for i in range(len(refseqs))])
    sc = scoring.Scoring(local=False)
    scs = sc.to_struct().reshape(-1)[0]
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    dix = lib.DeviceIndex(ix, 0)
    ctx.bind_index(dix)
    dref = lib.DeviceRef(genome) if nseq == 1 else lib.DeviceRef(extras["joined"], 0, recs=extras["recs"])
    npairs = 400
    lens = rng.integers(60, 180, size=2 * npairs)
    seqs, quals = _make_pairs(rng, genome, npairs, lens)
    pool = fmio.ReadPool.__new__(fmio.ReadPool)
    pool.seq = np.ascontiguousarray(np.concatenate(seqs)); pool.qual = np.ascontiguousarray(np.concatenate(quals))
    pool.reads = np.zeros(len(seqs), dtype=structs.read)
    pool.reads["rd_len"] = [len(s) for s in seqs]
    pool.reads["rd_off"] = np.concatenate([[0], np.cumsum(pool.reads["rd_len"])[:-1]])
    rl = pool.reads["rd_len"]
    minsc = np.array([sc.score_min.f_int(int(L)) for L in rl], dtype=np.int32)
    nceil = np.array([sc.nceil(int(L)) for L in rl], dtype=np.int32)
    fo = H.FrameOracle()
    pol = fo.policy(3, 500, 0, flags)
    pol["trim_to_ref"] = 1
    maxhalf, max_aln = 15, 2

    out, st, ps = ctx.hotpath_pairs(dref, pool.seq, pool.qual, pool.reads, minsc, nceil, pol, maxhalf=maxhalf, max_aln=max_aln)
    mate_dp = ctx.dp_fetch()                                   # the mate pass is what is still staged
    want = _oracle_round0(ix, refseqs, sc, pool, minsc, nceil, max_aln=max_aln, pair_boost=True)
    nj = out["njobs"]
    assert np.array_equal(ctx.hotpath_job_tidx(nj), want["tidx"])
    assert nj == want["njobs"] and nj > 500
    assert np.array_equal(out["job_read"][:nj], want["rid"]) and np.array_equal(out["job_fw"][:nj], want["fw"])
    wres, _, walns, wedits = want["dp"]
    for i in range(nj):
        A = dpio.results_to_py(wres, want["dp"][1], walns, wedits, i)[1]
        got = out["alns"][i * max_aln: i * max_aln + int(out["results"][i]["naln"])]
        assert [(a["score"], a["refoff"], a["nedits"]) for a in A] == [(int(a["score"]), int(a["refoff"]), int(a["nedits"])) for a in got]

    # ---- the mate pass, restated ----
    b_seq, b_qual, b_win, info = [], [], [], []
    mjobs = []
    ro = wo = 0
    for k in range(nj):
        r = wres[k]
        if r["naln"] == 0:
            continue
        a = walns[r["aln_off"]]
        gseq = refseqs[int(want["tidx"][k])]
        G = len(gseq)
        ri = int(want["rid"][k]); mate = ri ^ 1
        is1, fw = int(ri % 2 == 0), int(want["fw"][k])
        len_a, len_o = int(rl[ri]), int(rl[mate])
        ordg = fo.lib.bt2o_max_read_gaps(int(minsc[mate]), len_o, int(scs["match_bonus"]), int(scs["rd_gap_open"]), int(scs["rd_gap_ext"]))
        orfg = fo.lib.bt2o_max_ref_gaps(int(minsc[mate]), len_o, int(scs["match_bonus"]), int(scs["rf_gap_open"]), int(scs["rf_gap_ext"]))
        found, oleft, o4, ofw = fo.other_mate(pol, is1, fw, int(a["refoff"]), len_o + ordg, G, len_a if is1 else len_o, len_o if is1 else len_a)
        if not found:
            continue
        ok, rect = fo.frame_mate(int(not oleft), *o4, len_o, G, ordg, orfg, int(nceil[mate]), maxhalf, 1)
        if not ok:
            continue
        o = int(pool.reads["rd_off"][mate])
        rd, qu = pool.seq[o:o + len_o], pool.qual[o:o + len_o]
        if not ofw:
            rd, qu = COMP[rd[::-1]], qu[::-1]
        refl, ncol = rect[0], rect[1] - rect[0] + 1
        cols = refl + np.arange(ncol + 1)
        codes = np.where((cols >= 0) & (cols < G), gseq[np.clip(cols, 0, G - 1)], 4).astype(np.uint8)
        b_seq.append(rd); b_qual.append(qu); b_win.append((1 << codes).astype(np.uint8))
        j = np.zeros(1, dtype=structs.dp_job)[0]
        j["rd_off"], j["rd_len"], j["rf_off"], j["ncol"] = ro, len_o, wo, ncol
        j["minsc"], j["nceil"], j["refl"], j["triml"] = minsc[mate], nceil[mate], refl, rect[4]
        j["corel"], j["corer"], j["fw"], j["enable8"] = rect[6], rect[7], ofw, 1
        mjobs.append(j)
        ro += len_o; wo += ncol + 1
        info.append(dict(anchor_job=k, read=mate, ofw=ofw, oleft=oleft, rect=rect, refoff=int(a["refoff"]),
                         extent=_extent(len_a, a, wedits), is1=is1, fw=fw))
    mjobs = np.array(mjobs, dtype=structs.dp_job)
    nm = out["nmates"]
    assert nm == len(mjobs) == ps["mate_jobs"] and nm > 400
    mref = H.oracle_dp(scs.reshape(1) if hasattr(scs, "reshape") else scs, np.ascontiguousarray(np.concatenate(b_seq)),
                       np.ascontiguousarray(np.concatenate(b_qual)), np.ascontiguousarray(np.concatenate(b_win)), mjobs, max_aln=max_aln)
    nconc = naligned = ntrim = 0
    classes = set()
    for m in range(nm):
        gi, wi = out["mates"][m], info[m]
        assert (int(gi["anchor_job"]), int(gi["read"]), int(gi["ofw"]), int(gi["oleft"])) == (wi["anchor_job"], wi["read"], wi["ofw"], wi["oleft"])
        assert tuple(int(gi["rect"][f]) for f in RECT) == wi["rect"]
        assert (int(gi["anchor_refoff"]), int(gi["anchor_extent"])) == (wi["refoff"], wi["extent"])
        ntrim += wi["rect"][4] > 0 or wi["rect"][5] > 0
        assert dpio.results_to_py(*mref, m) == dpio.results_to_py(mate_dp[0], mate_dp[1], mate_dp[2], mate_dp[3], m), m
        assert out["mresults"][m].tobytes() == mate_dp[0][m].tobytes()
        r = mref[0][m]
        cls = 0
        if r["naln"]:
            a = mref[2][r["aln_off"]]
            ext = _extent(int(rl[wi["read"]]), a, mref[3])
            mate1 = wi["read"] % 2 == 0
            args = (int(a["refoff"]), ext, wi["ofw"], wi["refoff"], wi["extent"], wi["fw"]) if mate1 else \
                   (wi["refoff"], wi["extent"], wi["fw"], int(a["refoff"]), ext, wi["ofw"])
            cls = fo.classify(pol, *args)
            naligned += 1
            nconc += cls != 5
            assert int(gi["mate_extent"]) == ext
        assert int(gi["cls"]) == cls, m
        classes.add(cls)
    assert ps["mate_aligned"] == naligned and ps["concordant"] == nconc
    assert naligned > 0.8 * nm and ntrim >= 5 and len(classes) >= 3
    if flags[3:5] == (1, 1):
        assert nconc > 0.7 * naligned
    ctx.close()
