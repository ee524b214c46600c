"""GPU: size-independent properties of the chained hot path at a full-size batch (65536 reads).

The oracle is too slow for this many reads, so every alignment the batch runtime returns is
checked against what it claims, from first principles:
  * decode: applying the edit list to the oriented read reproduces the genome text at `refoff`
    (mismatch / read-gap characters are the genome's, gaps have the stated positions);
  * the score re-derived from the edit list with the Scoring tables equals the reported score,
    is >= minsc and <= the window's best; ns / gaps agree with the edits;
  * the alignment overlaps the diagonal the seed hit was located on, and (indel shifts aside) the
    read's simulated origin;
  * the sweep's exact end-to-end ranges are consistent with the alignments (an exact hit <=> a
    zero-edit alignment exists for that read's window).
"""
import numpy as np
import pytest

from bt2gpu import build, lib, scoring, structs, synth

pytestmark = pytest.mark.gpu

MM, REF_GAP, READ_GAP = 3, 2, 1
CODE = {ord("A"): 0, ord("C"): 1, ord("G"): 2, ord("T"): 3, ord("N"): 4}


def _decode(read, qual, fw, aln, edits, sc):
    """-> (aligned reference codes, score, ns, gaps) from the edit list of one alignment."""
    comp = np.array([3, 2, 1, 0, 4], np.uint8)
    rd = read if fw else comp[read[::-1]]
    qu = qual if fw else qual[::-1]
    n = len(rd)
    t5, t3 = int(aln["trim5"]), int(aln["trim3"])
    tb, te = (t5, t3) if fw else (t3, t5)               # trims of the oriented read (left, right)
    rdex = n - tb - te
    es = [(int(e["pos"]), int(e["chr"]), int(e["qchr"]), int(e["type"])) for e in edits]
    if not fw:                                           # invertPoss (edit.cpp:50-78) is an involution
        es = [(rdex - p - (0 if t == READ_GAP else 1), c, q, t) for (p, c, q, t) in reversed(es)]
    out, score, ns, gaps = [], 0, 0, 0
    i = 0
    last_gap = None                                      # (type, read position) of the previous gap edit
    for (p, c, q, t) in es:
        assert 0 <= p <= rdex and p >= i, "edits must be ordered by read position"
        if p > i:
            last_gap = None
        while i < p:
            out.append(int(rd[tb + i])); assert rd[tb + i] < 4; i += 1
        qq = min(max(int(qu[tb + min(i, rdex - 1)]) - 33, 0), 255)
        if t == MM:
            assert chr(q) == "ACGTN"[rd[tb + i]]
            rc_ = CODE[c]
            out.append(rc_)
            isn = rd[tb + i] > 3 or rc_ > 3
            score -= int(sc["npens"][qq] if isn else sc["mmpens"][qq])
            ns += int(isn)
            i += 1; last_gap = None
        elif t == REF_GAP:                               # read character with no reference character
            assert c == ord("-") and chr(q) == "ACGTN"[rd[tb + i]]
            ext = last_gap == (REF_GAP, i - 1)
            score -= int(sc["rf_gap_ext"] if ext else sc["rf_gap_open"])
            gaps += 1
            last_gap = (REF_GAP, i); i += 1
        else:                                            # reference character with no read character
            assert t == READ_GAP and q == ord("-")
            out.append(CODE[c])
            ext = last_gap == (READ_GAP, i)
            score -= int(sc["rd_gap_ext"] if ext else sc["rd_gap_open"])
            gaps += 1
            last_gap = (READ_GAP, i)
    while i < rdex:
        out.append(int(rd[tb + i])); assert rd[tb + i] < 4; i += 1
    return np.array(out, np.uint8), score, ns, gaps


def test_alignments_decode_to_the_genome_at_full_batch_size():
    rng = np.random.default_rng(2026)
    genome = synth.random_genome(rng, 32 << 20)
    ix, _ = build.build_index([genome], ["chr1"])
    sc = scoring.Scoring(local=False)
    scs = sc.to_struct().reshape(-1)[0]
    ctx = lib.Context(0)
    ctx.set_scoring(sc)
    dix = lib.DeviceIndex(ix, 0)
    ctx.bind_index(dix)
    dref = lib.DeviceRef(genome)
    n, L = 65536, 150
    reads, quals, pos, _ = synth.simulate_reads(rng, genome, n, L)
    clean = rng.choice(n, size=4096, replace=False)                 # simulate.pl reads are noisy: plant error-free ones too
    cp = rng.integers(0, len(genome) - L, size=len(clean))
    reads[clean] = genome[cp[:, None] + np.arange(L)[None, :]]
    pos[clean] = cp
    rcflip = rng.random(n) < 0.5
    reads[rcflip] = np.array([3, 2, 1, 0, 4], np.uint8)[reads[rcflip][:, ::-1]]
    quals[rcflip] = quals[rcflip][:, ::-1]
    rd = np.zeros(n, dtype=structs.read)
    rd["rd_off"] = np.arange(n, dtype=np.int64) * L
    rd["rd_len"] = L
    minsc = np.full(n, sc.score_min.f_int(L), dtype=np.int32)
    nceil = np.full(n, sc.nceil(L), dtype=np.int32)
    out, st = ctx.hotpath_round0(dref, np.ascontiguousarray(reads.reshape(-1)), np.ascontiguousarray(quals.reshape(-1)),
                                 rd, minsc, nceil, max_aln=2)
    nj = out["njobs"]
    assert nj > 0.3 * n and st["aligned"] > 0.99 * nj
    G = len(genome)
    mg = 15
    nchk = nexact = 0
    sweep = out["sweep"]
    for k in range(nj):
        r = out["results"][k]
        ri = int(out["job_read"][k]); fw = bool(out["job_fw"][k]); start = int(out["job_start"][k])
        assert r["naln"] <= 2 and r["cural"] <= r["ncand"]
        has_exact = False
        for a in out["alns"][k * 2: k * 2 + int(r["naln"])]:
            es = out["edits"][a["edit_off"]: a["edit_off"] + a["nedits"]]
            ref, score, ns, gaps = _decode(reads[ri], quals[ri], fw, a, es, scs)
            ro = int(a["refoff"])
            lo, hi = max(ro, 0), min(ro + len(ref), G)
            want = np.full(len(ref), 4, np.uint8)
            want[lo - ro: hi - ro] = genome[lo:hi]
            assert np.array_equal(ref, want), (k, ri)
            assert score == a["score"] and ns == a["ns"] and gaps == a["gaps"], (k, score, int(a["score"]))
            assert minsc[ri] <= a["score"] <= r["best"] and ns <= nceil[ri]
            assert a["trim5"] == 0 and a["trim3"] == 0                      # end-to-end: whole read aligned
            assert abs(ro - start) <= 2 * mg + gaps                        # inside the framed parallelogram
            has_exact |= (a["nedits"] == 0)
            nchk += 1
        # SeedAligner::exactSweep reports an exact end-to-end hit for this strand <=> the best alignment has no edits
        s = 0 if fw else 1
        ex = (not sweep["done"][ri][s]) and sweep["nedit"][ri][s] == 0 and sweep["bot"][ri][s] > sweep["top"][ri][s]
        if ex:
            assert has_exact, (k, ri)
            nexact += 1
        elif r["naln"] > 0 and r["best"] == 0:
            assert False, "a zero-penalty alignment without an exact sweep hit"
    origin_ok = np.abs(out["job_start"][:nj] - pos[out["job_read"][:nj]]) <= 64
    assert origin_ok.mean() > 0.99
    assert nchk > 0.3 * n and nexact >= 4000
    ctx.close()
