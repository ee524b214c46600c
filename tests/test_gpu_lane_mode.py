"""GPU: a context configured the way the batched server's lanes configure theirs -- polling waits
(bt2gpu_ctx_set_wait), no event records (bt2gpu_ctx_set_timing), reads staged without qualities where no call reads
them -- gives bit for bit what a default context gives, through the same C-ABI batch calls, call after call."""
import numpy as np
import pytest

import helpers as H
from bt2gpu import build, fmio, lib, scoring, structs, synth

pytestmark = pytest.mark.gpu


def _reads(rng, genome, n, lens):
    seqs, quals = [], []
    for _ in range(n):
        ln = int(rng.integers(lens[0], lens[1]))
        p = int(rng.integers(0, len(genome) - ln))
        s = genome[p:p + ln].copy()
        mm = rng.random(ln) < 0.02
        s[mm] = rng.integers(0, 5, size=int(mm.sum()))
        seqs.append(s.astype(np.uint8))
        quals.append((rng.integers(2, 41, size=ln) + 33).astype(np.uint8))
    return fmio.ReadPool(seqs, quals)


def _dp_canon(r):
    """Per problem: the scalars, its candidates, its alignments with their edits -- without the pool offsets, which
    depend on the order in which thread blocks claimed their slices."""
    results, cands, alns, edits = r
    out = []
    for x in results:
        c = cands[int(x["cand_off"]):int(x["cand_off"]) + int(x["ncand"])].tobytes()
        al = []
        for a in alns[int(x["aln_off"]):int(x["aln_off"]) + int(x["naln"])]:
            e = edits[int(a["edit_off"]):int(a["edit_off"]) + int(a["nedits"])].tobytes()
            al.append((tuple(int(a[k]) for k in a.dtype.names if k != "edit_off"), e))
        out.append((tuple(int(x[k]) for k in ("best", "aligned", "width", "ncand", "naln", "tail_tries", "cural")), c, al))
    return out


def test_lane_mode_context_matches_default_context():
    rng = np.random.default_rng(2026)
    genome = rng.integers(0, 4, size=300000, dtype=np.uint8)
    genome[5000:5200] = genome[90000:90200]                     # a repeat: seed ranges wider than one row
    ix, _ = build.build_index([genome], ["g"])
    dix = lib.DeviceIndex(ix, 0)
    sc = scoring.Scoring(local=False)
    a, b = lib.Context(0), lib.Context(0)
    try:
        for c in (a, b):
            c.set_scoring(sc)
            c.bind_index(dix)
        b.set_wait(20000)
        b.set_timing(False)
        for rep in range(3):                      # consecutive calls reuse the context's arenas
            pool = _reads(rng, genome, 300 + 200 * rep, (30, 200))
            a.stage_reads(pool)
            b.stage_reads(pool, quals=False)
            assert a.exact_sweep().tobytes() == b.exact_sweep().tobytes()
            sj, _ = fmio.instantiate_exact_seeds(pool, 20, lambda n: 6)
            ha, hb = a.seed_exact(sj), b.seed_exact(sj)
            assert ha.tobytes() == hb.tobytes()
            hit = np.nonzero(ha["botf"] > ha["topf"])[0]
            xj = np.zeros(len(hit), dtype=structs.extend_job)
            for k in ("topf", "botf", "topb", "botb"):
                xj[k] = ha[k][hit]
            for k in ("read", "off", "len", "fw"):
                xj[k] = sj[k][hit]
            assert a.extend(xj).tobytes() == b.extend(xj).tobytes()
            rows = rng.integers(0, ix.fw.len + 1, size=700).astype(np.uint32)
            (oa, sa), (ob, sb) = a.locate(rows), b.locate(rows)
            assert np.array_equal(oa, ob) and np.array_equal(sa, sb)
            # DP: seed-extension windows around the reads' true diagonals
            seq, qual, win, jobs = synth.dp_extend_batch(rng, genome, 400 + 100 * rep, 120, sc)[:4]
            ra, rb = a.dp_align_batch_abi(seq, qual, win, jobs, max_aln=4), b.dp_align_batch_abi(seq, qual, win, jobs, max_aln=4)
            assert _dp_canon(ra) == _dp_canon(rb)
            assert (ra[0]["naln"] > 0).sum() > 100
            assert (ha["botf"] > ha["topf"]).sum() > 100 and (sa > 0).sum() > 100
        with pytest.raises(lib.Bt2GpuError):
            b.set_wait(200000000)                 # above 100 ms: refused
    finally:
        a.close(); b.close(); dix.close()
